// obj_loader.cpp -- see obj_loader.hpp.  Behavioural restatement of the tinyobjloader 2.0.0
// paths Scene::loadObjects exercises (external/include/tiny_obj_loader.h: number parsing
// :913-1050, index fix-up :835-866, triple parsing :1218-1269, quad split :1555-1663,
// LoadObj loop :2645-3000, LoadMtl :2125-2480).  Written from the behaviour, not copied.
#include "obj_loader.hpp"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <set>
#include <sstream>

#include "parallel.hpp"

namespace ptb {

static inline bool is_digit(char c) { return c >= '0' && c <= '9'; }
static inline bool is_space(char c) { return c == ' ' || c == '\t'; }
static inline bool is_eol(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// Decimal -> double exactly the way tinyobj does it: the integer part accumulates as
// mantissa*10+d, fraction digits add d * 10^-k (table for k < 8, pow() beyond), and a decimal
// exponent e is applied as ldexp(mantissa * 5^e, e).  This is not correctly rounded, so it
// must be reproduced rather than replaced by strtod.
bool tinyobj_parse_double(const char* s, const char* s_end, double* result) {
    if (s >= s_end) return false;
    double mantissa = 0.0;
    int exponent = 0;
    char sign = '+', exp_sign = '+';
    const char* curr = s;
    int read = 0;
    bool more = false, leading_dot = false;

    if (*curr == '+' || *curr == '-') {
        sign = *curr++;
        if (curr != s_end && *curr == '.') leading_dot = true;
    } else if (is_digit(*curr)) {
    } else if (*curr == '.') {
        leading_dot = true;
    } else {
        return false;
    }
    more = (curr != s_end);
    if (!leading_dot) {
        while (more && is_digit(*curr)) {
            mantissa *= 10;
            mantissa += (int)(*curr - '0');
            curr++; read++;
            more = (curr != s_end);
        }
        if (read == 0) return false;
    }
    if (more) {
        bool want_exp = false;
        if (*curr == '.') {
            static const double lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
            curr++; read = 1;
            more = (curr != s_end);
            while (more && is_digit(*curr)) {
                mantissa += (int)(*curr - '0') * (read < 8 ? lut[read] : std::pow(10.0, -read));
                read++; curr++;
                more = (curr != s_end);
            }
            want_exp = more;
        } else if (*curr == 'e' || *curr == 'E') {
            want_exp = true;
        }
        if (want_exp && (*curr == 'e' || *curr == 'E')) {
            curr++;
            more = (curr != s_end);
            if (more && (*curr == '+' || *curr == '-')) exp_sign = *curr++;
            else if (more && is_digit(*curr)) {}
            else return false;
            read = 0;
            more = (curr != s_end);
            while (more && is_digit(*curr)) {
                if (exponent > 2147483647 / 10) return false;
                exponent = exponent * 10 + (int)(*curr - '0');
                curr++; read++;
                more = (curr != s_end);
            }
            exponent *= (exp_sign == '+' ? 1 : -1);
            if (read == 0) return false;
        }
    }
    *result = (sign == '+' ? 1 : -1) *
              (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
    return true;
}

static float parse_real(const char** tok, double dflt = 0.0) {
    *tok += strspn(*tok, " \t");
    const char* end = *tok + strcspn(*tok, " \t\r");
    double v = dflt;
    tinyobj_parse_double(*tok, end, &v);
    *tok = end;
    return (float)v;
}
static int parse_int(const char** tok) {
    *tok += strspn(*tok, " \t");
    int i = atoi(*tok);
    *tok += strcspn(*tok, " \t\r");
    return i;
}
static std::string parse_string(const char** tok) {
    *tok += strspn(*tok, " \t");
    size_t e = strcspn(*tok, " \t\r");
    std::string s(*tok, *tok + e);
    *tok += e;
    return s;
}

static bool fix_index(int idx, int n, int* ret, bool allow_zero) {
    if (idx > 0) { *ret = idx - 1; return true; }
    if (idx == 0) { *ret = -1; return allow_zero; }
    *ret = n + idx;
    return *ret >= 0;
}

static bool parse_triple(const char** tok, int vsize, int vnsize, int vtsize, ObjIndex* out) {
    ObjIndex vi;
    if (!fix_index(atoi(*tok), vsize, &vi.v, false)) return false;
    *tok += strcspn(*tok, "/ \t\r");
    if ((*tok)[0] != '/') { *out = vi; return true; }
    (*tok)++;
    if ((*tok)[0] == '/') {   // i//k
        (*tok)++;
        if (!fix_index(atoi(*tok), vnsize, &vi.vn, true)) return false;
        *tok += strcspn(*tok, "/ \t\r");
        *out = vi;
        return true;
    }
    if (!fix_index(atoi(*tok), vtsize, &vi.vt, true)) return false;   // i/j
    *tok += strcspn(*tok, "/ \t\r");
    if ((*tok)[0] != '/') { *out = vi; return true; }
    (*tok)++;                                                          // i/j/k
    if (!fix_index(atoi(*tok), vnsize, &vi.vn, true)) return false;
    *tok += strcspn(*tok, "/ \t\r");
    *out = vi;
    return true;
}

static void getline_any(std::istream& is, std::string& t) {   // handles \n, \r\n and bare \r
    t.clear();
    std::streambuf* sb = is.rdbuf();
    for (;;) {
        int c = sb->sbumpc();
        if (c == '\n') return;
        if (c == '\r') { if (sb->sgetc() == '\n') sb->sbumpc(); return; }
        if (c == EOF) { if (t.empty()) is.setstate(std::ios::eofbit); return; }
        t += (char)c;
    }
}

static std::string strip_bom(const std::string& s) {
    if (s.size() >= 3 && (unsigned char)s[0] == 0xEF && (unsigned char)s[1] == 0xBB && (unsigned char)s[2] == 0xBF)
        return s.substr(3);
    return s;
}

// "-opt args... name": skip the MTL texture options tinyobj recognises, keep the name.
static std::string parse_texname(const char* tok) {
    std::string name;
    while (!is_eol(*tok)) {
        tok += strspn(tok, " \t");
        if (is_eol(*tok)) break;
        if (*tok == '-') {
            std::string opt = parse_string(&tok);
            int nargs = 1;
            if (opt == "-o" || opt == "-s" || opt == "-t") nargs = 3;
            else if (opt == "-mm") nargs = 2;
            for (int i = 0; i < nargs; i++) {
                const char* save = tok;
                save += strspn(save, " \t");
                if (nargs == 3 && i > 0 && !(is_digit(*save) || *save == '-' || *save == '+' || *save == '.')) break;
                parse_string(&tok);
            }
        } else {
            name = parse_string(&tok);   // tinyobj keeps the last bare token
        }
    }
    return name;
}

static void load_mtl(std::istream& in, std::vector<MtlMaterial>& materials, std::map<std::string, int>& material_map) {
    MtlMaterial m;
    bool has_d = false, has_kd = false;
    size_t line_no = 0;
    std::string line;
    while (in.peek() != -1) {
        getline_any(in, line);
        line_no++;
        if (!line.empty()) line = line.substr(0, line.find_last_not_of(" \t") + 1);
        if (line.empty()) continue;
        if (line_no == 1) line = strip_bom(line);
        const char* tok = line.c_str();
        tok += strspn(tok, " \t");
        if (tok[0] == '\0' || tok[0] == '#') continue;

        if (!strncmp(tok, "newmtl", 6) && is_space(tok[6])) {
            if (!m.name.empty()) {
                material_map.insert(std::make_pair(m.name, (int)materials.size()));
                materials.push_back(m);
            }
            m = MtlMaterial();
            has_d = has_kd = false;
            tok += 7;
            m.name = parse_string(&tok);
            continue;
        }
        auto real3 = [&](float* dst) { tok += 2; dst[0] = parse_real(&tok); dst[1] = parse_real(&tok); dst[2] = parse_real(&tok); };
        if (tok[0] == 'K' && tok[1] == 'd' && is_space(tok[2])) { real3(m.diffuse); has_kd = true; continue; }
        if (tok[0] == 'K' && tok[1] == 's' && is_space(tok[2])) { real3(m.specular); continue; }
        if (tok[0] == 'K' && tok[1] == 'a' && is_space(tok[2])) { float t[3]; real3(t); continue; }
        if ((tok[0] == 'K' && tok[1] == 't' && is_space(tok[2])) || (tok[0] == 'T' && tok[1] == 'f' && is_space(tok[2]))) { real3(m.transmittance); continue; }
        if (tok[0] == 'K' && tok[1] == 'e' && is_space(tok[2])) { real3(m.emission); continue; }
        if (tok[0] == 'N' && tok[1] == 'i' && is_space(tok[2])) { tok += 2; m.ior = parse_real(&tok); continue; }
        if (tok[0] == 'N' && tok[1] == 's' && is_space(tok[2])) { tok += 2; m.shininess = parse_real(&tok); continue; }
        if (!strncmp(tok, "illum", 5) && is_space(tok[5])) { tok += 6; m.illum = parse_int(&tok); continue; }
        if (tok[0] == 'd' && is_space(tok[1])) { tok += 1; m.dissolve = parse_real(&tok); has_d = true; continue; }
        if (tok[0] == 'T' && tok[1] == 'r' && is_space(tok[2])) {
            tok += 2;
            if (!has_d) m.dissolve = 1.0f - parse_real(&tok);
            continue;
        }
        if (!strncmp(tok, "map_Kd", 6) && is_space(tok[6])) {
            m.diffuse_texname = parse_texname(tok + 7);
            if (!has_kd) { m.diffuse[0] = m.diffuse[1] = m.diffuse[2] = 0.6f; }
            continue;
        }
        if ((!strncmp(tok, "map_bump", 8) || !strncmp(tok, "map_Bump", 8)) && is_space(tok[8])) { m.bump_texname = parse_texname(tok + 9); continue; }
        if (!strncmp(tok, "bump", 4) && is_space(tok[4])) { m.bump_texname = parse_texname(tok + 5); continue; }
        if (!strncmp(tok, "map_Pr", 6) && is_space(tok[6])) { m.roughness_texname = parse_texname(tok + 7); continue; }
        // every other statement is irrelevant to Scene::loadObjects
    }
    material_map.insert(std::make_pair(m.name, (int)materials.size()));   // tinyobj flushes the last one unconditionally
    materials.push_back(m);
}

// ---- OBJ body: chunked, multi-threaded, same result as a line-by-line scan --------------------------------------
// The only order dependences of the statement stream are (a) relative (negative) face indices, which need the number
// of v / vn / vt records seen so far, (b) the material in effect (`usemtl`, resolved through `mtllib`) and (c)
// tinyobj's flush points (material change, `g`, `o`, end of file), which matter because a quad is dropped when one
// of its vertex indices is not yet defined *at the flush that exports it* (exportGroupsToShape).  So:
//   pass 1 (parallel)  cut the file into chunks at line ends; count v / vn / vt records and lines per chunk
//   prefix sums        give every chunk its record bases
//   pass 2 (parallel)  parse: reals go straight to their final position, faces and statement events stay per chunk
//   replay (serial)    walk the few events in file order: load MTLs, resolve materials, number the flushes
//   pass 3 (parallel)  triangulate per chunk (the quad split needs the positions), then copy out by prefix offsets
namespace {

struct ObjEvent {
    enum Kind { USEMTL, MTLLIB, GROUP } kind;
    std::string arg;
    size_t face_pos;     // faces of this chunk parsed before the statement
    size_t v_count;      // v records of the whole file seen before the statement
};

struct FaceRun {         // faces [f0, f1) of a chunk that share a material and a flush
    size_t f0, f1;
    int material;
    size_t flush;
};

struct ObjChunk {
    const char* begin = nullptr;
    const char* end = nullptr;
    size_t nv = 0, nvn = 0, nvt = 0, nlines = 0;     // pass 1
    size_t v0 = 0, vn0 = 0, vt0 = 0, line0 = 0;      // prefix sums
    std::vector<ObjIndex> corners;                    // pass 2: all face corners, faces back to back
    std::vector<uint32_t> face_start;                 //         first corner of each face (+ sentinel)
    std::vector<ObjEvent> events;
    int greatest_v = -1, greatest_vn = -1, greatest_vt = -1;
    std::string err;
    std::vector<FaceRun> runs;                        // replay
    std::vector<ObjIndex> tri;                        // pass 3
    std::vector<int> tri_material;
    size_t tri0 = 0;
};

// One text line [p, e) -> next line start.  Line ends: \n, \r\n, bare \r (getline_any above).
inline const char* line_end(const char* p, const char* end) {
    while (p < end && *p != '\n' && *p != '\r') p++;
    return p;
}
inline const char* advance_line(const char* e, const char* end) {
    if (e >= end) return end;
    if (*e == '\n') return e + 1;
    /* '\r' */
    if (e + 1 < end && e[1] == '\n') return e + 2;
    return e + 1;
}

}  // namespace

bool load_obj(const std::string& filename, const std::string& base_dir, ObjData& out, std::string& err) {
    std::string text;
    {
        std::ifstream ifs(filename.c_str(), std::ios::binary);
        if (!ifs) { err = "Cannot open file [" + filename + "]"; return false; }
        ifs.seekg(0, std::ios::end);
        std::streamoff size = ifs.tellg();
        ifs.seekg(0, std::ios::beg);
        text.resize((size_t)std::max<std::streamoff>(size, 0));
        if (size > 0) ifs.read(&text[0], size);
    }
    std::string dir = base_dir;
    if (!dir.empty() && dir[dir.size() - 1] != '/') dir += '/';
    const char* file_begin = text.data();
    const char* file_end = text.data() + text.size();
    if (text.size() >= 3 && (unsigned char)text[0] == 0xEF && (unsigned char)text[1] == 0xBB && (unsigned char)text[2] == 0xBF)
        file_begin += 3;   // BOM on line 1

    const int threads = loader_threads();
    // ---- at most 4 chunks per thread, of at least 1 MiB, cut at line ends
    size_t chunk_bytes = 1u << 20;
    if (const char* e = getenv("PTB_LOADER_CHUNK_BYTES")) chunk_bytes = (size_t)std::max(1, atoi(e));   // tests: many chunks on small files
    size_t want = std::max<size_t>(1, std::min<size_t>((size_t)threads * 4, (size_t)(file_end - file_begin) / chunk_bytes));
    std::vector<ObjChunk> chunks;
    {
        const char* p = file_begin;
        for (size_t c = 0; c < want && p < file_end; c++) {
            const char* q = (c + 1 == want) ? file_end : file_begin + (size_t)(file_end - file_begin) * (c + 1) / want;
            if (q < p) q = p;
            if (q < file_end) q = advance_line(line_end(q, file_end), file_end);
            ObjChunk ch;
            ch.begin = p;
            ch.end = q;
            chunks.push_back(std::move(ch));
            p = q;
        }
        if (chunks.empty()) { ObjChunk ch; ch.begin = ch.end = file_begin; chunks.push_back(std::move(ch)); }
    }

    // ---- pass 1: record counts
    parallel_tasks(chunks.size(), threads, [&](size_t c) {
        ObjChunk& ch = chunks[c];
        for (const char* p = ch.begin; p < ch.end;) {
            const char* e = line_end(p, ch.end);
            const char* t = p;
            while (t < e && (*t == ' ' || *t == '\t')) t++;
            if (t + 1 < e && t[0] == 'v') {
                if (is_space(t[1])) ch.nv++;
                else if (t + 2 < e && t[1] == 'n' && is_space(t[2])) ch.nvn++;
                else if (t + 2 < e && t[1] == 't' && is_space(t[2])) ch.nvt++;
            }
            ch.nlines++;
            p = advance_line(e, ch.end);
        }
    });
    size_t tv = 0, tvn = 0, tvt = 0, tl = 0;
    for (ObjChunk& ch : chunks) {
        ch.v0 = tv; ch.vn0 = tvn; ch.vt0 = tvt; ch.line0 = tl;
        tv += ch.nv; tvn += ch.nvn; tvt += ch.nvt; tl += ch.nlines;
    }
    out.v.assign(tv * 3, 0.0f);
    out.vn.assign(tvn * 3, 0.0f);
    out.vt.assign(tvt * 2, 0.0f);

    // ---- pass 2: parse
    parallel_tasks(chunks.size(), threads, [&](size_t c) {
        ObjChunk& ch = chunks[c];
        size_t iv = ch.v0, ivn = ch.vn0, ivt = ch.vt0, line_no = ch.line0;
        std::string line;
        ch.corners.reserve((size_t)(ch.end - ch.begin) / 12);
        for (const char* p = ch.begin; p < ch.end;) {
            const char* e = line_end(p, ch.end);
            line.assign(p, e);
            p = advance_line(e, ch.end);
            line_no++;
            if (line.empty()) continue;
            const char* tok = line.c_str();
            tok += strspn(tok, " \t");
            if (tok[0] == '\0' || tok[0] == '#') continue;

            if (tok[0] == 'v' && is_space(tok[1])) {
                tok += 2;
                float x = parse_real(&tok), y = parse_real(&tok), z = parse_real(&tok);
                out.v[3 * iv] = x; out.v[3 * iv + 1] = y; out.v[3 * iv + 2] = z;
                iv++;
                continue;
            }
            if (tok[0] == 'v' && tok[1] == 'n' && is_space(tok[2])) {
                tok += 3;
                float x = parse_real(&tok), y = parse_real(&tok), z = parse_real(&tok);
                out.vn[3 * ivn] = x; out.vn[3 * ivn + 1] = y; out.vn[3 * ivn + 2] = z;
                ivn++;
                continue;
            }
            if (tok[0] == 'v' && tok[1] == 't' && is_space(tok[2])) {
                tok += 3;
                float x = parse_real(&tok), y = parse_real(&tok);
                out.vt[2 * ivt] = x; out.vt[2 * ivt + 1] = y;
                ivt++;
                continue;
            }
            if (tok[0] == 'f' && is_space(tok[1])) {
                tok += 2;
                tok += strspn(tok, " \t");
                ch.face_start.push_back((uint32_t)ch.corners.size());
                while (!is_eol(tok[0]) && tok[0] != '#') {
                    ObjIndex vi;
                    if (!parse_triple(&tok, (int)iv, (int)ivn, (int)ivt, &vi)) {
                        if (ch.err.empty()) ch.err = "Failed to parse `f' line (zero or invalid relative vertex index), line " + std::to_string(line_no);
                        return;
                    }
                    if (vi.v > ch.greatest_v) ch.greatest_v = vi.v;
                    if (vi.vn > ch.greatest_vn) ch.greatest_vn = vi.vn;
                    if (vi.vt > ch.greatest_vt) ch.greatest_vt = vi.vt;
                    ch.corners.push_back(vi);
                    tok += strspn(tok, " \t\r");
                }
                continue;
            }
            if (!strncmp(tok, "usemtl", 6)) {
                tok += 6;
                ch.events.push_back(ObjEvent{ObjEvent::USEMTL, parse_string(&tok), ch.face_start.size(), iv});
                continue;
            }
            if (!strncmp(tok, "mtllib", 6) && is_space(tok[6])) {
                ch.events.push_back(ObjEvent{ObjEvent::MTLLIB, std::string(tok + 7), ch.face_start.size(), iv});
                continue;
            }
            if ((tok[0] == 'g' || tok[0] == 'o') && is_space(tok[1])) {
                ch.events.push_back(ObjEvent{ObjEvent::GROUP, std::string(), ch.face_start.size(), iv});
                continue;
            }
            // s, l, p, vw, t and unknown statements do not produce triangles
        }
        ch.face_start.push_back((uint32_t)ch.corners.size());   // sentinel
    });
    for (const ObjChunk& ch : chunks)
        if (!ch.err.empty()) { err = ch.err; return false; }

    // ---- replay the statement events in file order
    std::map<std::string, int> material_map;
    std::set<std::string> material_files;
    std::vector<size_t> flush_v_count;   // v records defined when flush k happened
    int material = -1;
    for (ObjChunk& ch : chunks) {
        const size_t nfaces = ch.face_start.size() - 1;
        size_t f0 = 0;
        auto close_run = [&](size_t f1) {
            if (f1 > f0) ch.runs.push_back(FaceRun{f0, f1, material, flush_v_count.size()});
            f0 = f1;
        };
        for (const ObjEvent& ev : ch.events) {
            if (ev.kind == ObjEvent::MTLLIB) {
                std::stringstream ss(ev.arg);
                std::string fn;
                bool found = false;
                while (ss >> fn) {
                    if (material_files.count(fn)) { found = true; continue; }
                    std::ifstream m((dir + fn).c_str(), std::ios::binary);
                    if (m) {
                        load_mtl(m, out.materials, material_map);
                        material_files.insert(fn);
                        found = true;
                        break;
                    }
                }
                if (!found) out.warn += "Failed to load material file(s). Use default material.\n";
            } else if (ev.kind == ObjEvent::USEMTL) {
                int id = -1;
                auto it = material_map.find(ev.arg);
                if (it != material_map.end()) id = it->second;
                else out.warn += "material [ '" + ev.arg + "' ] not found in .mtl\n";
                if (id != material) {
                    close_run(ev.face_pos);
                    flush_v_count.push_back(ev.v_count);
                    material = id;
                }
            } else {   // g / o: new shape, order unchanged
                close_run(ev.face_pos);
                flush_v_count.push_back(ev.v_count);
            }
        }
        close_run(nfaces);
    }
    flush_v_count.push_back(tv);   // end of file

    // ---- pass 3: triangulate (tinyobj exportGroupsToShape with triangulate = true, tiny_obj_loader.h:1555-1663)
    const std::vector<float>& v = out.v;
    parallel_tasks(chunks.size(), threads, [&](size_t c) {
        ObjChunk& ch = chunks[c];
        ch.tri.reserve(ch.corners.size());
        for (const FaceRun& run : ch.runs) {
            const size_t v_defined = flush_v_count[run.flush] * 3;
            for (size_t f = run.f0; f < run.f1; f++) {
                const ObjIndex* fc = &ch.corners[ch.face_start[f]];
                const size_t n = ch.face_start[f + 1] - ch.face_start[f];
                if (n < 3) continue;                       // "Degenerated face found"
                if (n == 3) {
                    ch.tri.insert(ch.tri.end(), fc, fc + 3);
                    ch.tri_material.push_back(run.material);
                } else if (n == 4) {
                    size_t vi[4];
                    bool bad = false;
                    for (int k = 0; k < 4; k++) { vi[k] = (size_t)fc[k].v; if (3 * vi[k] + 2 >= v_defined) bad = true; }
                    if (bad) continue;                     // "Face with invalid vertex index found"
                    float e02x = v[vi[2] * 3 + 0] - v[vi[0] * 3 + 0], e02y = v[vi[2] * 3 + 1] - v[vi[0] * 3 + 1], e02z = v[vi[2] * 3 + 2] - v[vi[0] * 3 + 2];
                    float e13x = v[vi[3] * 3 + 0] - v[vi[1] * 3 + 0], e13y = v[vi[3] * 3 + 1] - v[vi[1] * 3 + 1], e13z = v[vi[3] * 3 + 2] - v[vi[1] * 3 + 2];
                    float sqr02 = e02x * e02x + e02y * e02y + e02z * e02z;
                    float sqr13 = e13x * e13x + e13y * e13y + e13z * e13z;
                    const int a[2][3] = {{0, 1, 2}, {0, 2, 3}}, b[2][3] = {{0, 1, 3}, {1, 2, 3}};
                    const int (*sel)[3] = (sqr02 < sqr13) ? a : b;
                    for (int t = 0; t < 2; t++) {
                        for (int k = 0; k < 3; k++) ch.tri.push_back(fc[sel[t][k]]);
                        ch.tri_material.push_back(run.material);
                    }
                } else {
                    if (ch.err.empty()) ch.err = "polygon with " + std::to_string(n) + " vertices: only triangles and quads are supported";
                    return;
                }
            }
        }
    });
    size_t ntri = 0;
    int greatest_v = -1, greatest_vn = -1, greatest_vt = -1;
    for (ObjChunk& ch : chunks) {
        if (!ch.err.empty()) { err = ch.err; return false; }
        ch.tri0 = ntri;
        ntri += ch.tri_material.size();
        greatest_v = std::max(greatest_v, ch.greatest_v);
        greatest_vn = std::max(greatest_vn, ch.greatest_vn);
        greatest_vt = std::max(greatest_vt, ch.greatest_vt);
    }
    out.tri_indices.resize(ntri * 3);
    out.tri_material.resize(ntri);
    parallel_tasks(chunks.size(), threads, [&](size_t c) {
        const ObjChunk& ch = chunks[c];
        std::copy(ch.tri.begin(), ch.tri.end(), out.tri_indices.begin() + ch.tri0 * 3);
        std::copy(ch.tri_material.begin(), ch.tri_material.end(), out.tri_material.begin() + ch.tri0);
    });
    if (greatest_v >= (int)tv || greatest_vn >= (int)tvn || greatest_vt >= (int)tvt) {
        err = "face index out of range in " + filename;
        return false;
    }
    return true;
}

}  // namespace ptb
