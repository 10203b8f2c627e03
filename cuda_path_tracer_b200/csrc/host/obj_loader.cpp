// obj_loader.cpp -- see obj_loader.hpp.  Behavioural restatement of the tinyobjloader 2.0.0
// paths Scene::loadObjects exercises (external/include/tiny_obj_loader.h: number parsing
// :913-1050, index fix-up :835-866, triple parsing :1218-1269, quad split :1555-1663,
// LoadObj loop :2645-3000, LoadMtl :2125-2480).  Written from the behaviour, not copied.
#include "obj_loader.hpp"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <set>
#include <sstream>

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

// One flush of tinyobj's prim_group into the running shape (exportGroupsToShape): faces keep
// file order and all take the material that was current when the flush happened.
static bool flush_faces(std::vector<std::vector<ObjIndex>>& faces, int material, const std::vector<float>& v,
                        ObjData& out, std::string& err) {
    for (const auto& f : faces) {
        size_t n = f.size();
        if (n < 3) continue;                       // "Degenerated face found"
        if (n == 3) {
            out.tri_indices.insert(out.tri_indices.end(), f.begin(), f.end());
            out.tri_material.push_back(material);
        } else if (n == 4) {
            size_t vi[4];
            bool bad = false;
            for (int k = 0; k < 4; k++) { vi[k] = (size_t)f[k].v; if (3 * vi[k] + 2 >= v.size()) bad = true; }
            if (bad) continue;                     // "Face with invalid vertex index found"
            float e02x = v[vi[2] * 3 + 0] - v[vi[0] * 3 + 0], e02y = v[vi[2] * 3 + 1] - v[vi[0] * 3 + 1], e02z = v[vi[2] * 3 + 2] - v[vi[0] * 3 + 2];
            float e13x = v[vi[3] * 3 + 0] - v[vi[1] * 3 + 0], e13y = v[vi[3] * 3 + 1] - v[vi[1] * 3 + 1], e13z = v[vi[3] * 3 + 2] - v[vi[1] * 3 + 2];
            float sqr02 = e02x * e02x + e02y * e02y + e02z * e02z;
            float sqr13 = e13x * e13x + e13y * e13y + e13z * e13z;
            const int a[2][3] = {{0, 1, 2}, {0, 2, 3}}, b[2][3] = {{0, 1, 3}, {1, 2, 3}};
            const int (*sel)[3] = (sqr02 < sqr13) ? a : b;
            for (int t = 0; t < 2; t++) {
                for (int k = 0; k < 3; k++) out.tri_indices.push_back(f[sel[t][k]]);
                out.tri_material.push_back(material);
            }
        } else {
            err = "polygon with " + std::to_string(n) + " vertices: only triangles and quads are supported";
            return false;
        }
    }
    faces.clear();
    return true;
}

bool load_obj(const std::string& filename, const std::string& base_dir, ObjData& out, std::string& err) {
    std::ifstream ifs(filename.c_str(), std::ios::binary);
    if (!ifs) { err = "Cannot open file [" + filename + "]"; return false; }
    std::string dir = base_dir;
    if (!dir.empty() && dir[dir.size() - 1] != '/') dir += '/';

    std::vector<std::vector<ObjIndex>> faces;       // tinyobj's prim_group.faceGroup
    std::map<std::string, int> material_map;
    std::set<std::string> material_files;
    int material = -1;
    int greatest_v = -1, greatest_vn = -1, greatest_vt = -1;
    size_t line_no = 0;
    std::string line;
    while (ifs.peek() != -1) {
        getline_any(ifs, line);
        line_no++;
        if (line.empty()) continue;
        if (line_no == 1) line = strip_bom(line);
        const char* tok = line.c_str();
        tok += strspn(tok, " \t");
        if (tok[0] == '\0' || tok[0] == '#') continue;

        if (tok[0] == 'v' && is_space(tok[1])) {
            tok += 2;
            float x = parse_real(&tok), y = parse_real(&tok), z = parse_real(&tok);
            out.v.push_back(x); out.v.push_back(y); out.v.push_back(z);
            continue;
        }
        if (tok[0] == 'v' && tok[1] == 'n' && is_space(tok[2])) {
            tok += 3;
            float x = parse_real(&tok), y = parse_real(&tok), z = parse_real(&tok);
            out.vn.push_back(x); out.vn.push_back(y); out.vn.push_back(z);
            continue;
        }
        if (tok[0] == 'v' && tok[1] == 't' && is_space(tok[2])) {
            tok += 3;
            float x = parse_real(&tok), y = parse_real(&tok);
            out.vt.push_back(x); out.vt.push_back(y);
            continue;
        }
        if (tok[0] == 'f' && is_space(tok[1])) {
            tok += 2;
            tok += strspn(tok, " \t");
            std::vector<ObjIndex> face;
            face.reserve(3);
            while (!is_eol(tok[0]) && tok[0] != '#') {
                ObjIndex vi;
                if (!parse_triple(&tok, (int)(out.v.size() / 3), (int)(out.vn.size() / 3), (int)(out.vt.size() / 2), &vi)) {
                    err = "Failed to parse `f' line (zero or invalid relative vertex index), line " + std::to_string(line_no);
                    return false;
                }
                if (vi.v > greatest_v) greatest_v = vi.v;
                if (vi.vn > greatest_vn) greatest_vn = vi.vn;
                if (vi.vt > greatest_vt) greatest_vt = vi.vt;
                face.push_back(vi);
                tok += strspn(tok, " \t\r");
            }
            faces.push_back(std::move(face));
            continue;
        }
        if (!strncmp(tok, "usemtl", 6)) {
            tok += 6;
            std::string name = parse_string(&tok);
            int id = -1;
            auto it = material_map.find(name);
            if (it != material_map.end()) id = it->second;
            else out.warn += "material [ '" + name + "' ] not found in .mtl\n";
            if (id != material) {
                if (!flush_faces(faces, material, out.v, out, err)) return false;
                material = id;
            }
            continue;
        }
        if (!strncmp(tok, "mtllib", 6) && is_space(tok[6])) {
            tok += 7;
            std::stringstream ss(tok);
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
            continue;
        }
        if ((tok[0] == 'g' || tok[0] == 'o') && is_space(tok[1])) {
            if (!flush_faces(faces, material, out.v, out, err)) return false;   // new shape; order unchanged
            continue;
        }
        // s, l, p, vw, t and unknown statements do not produce triangles
    }
    if (!flush_faces(faces, material, out.v, out, err)) return false;
    if (greatest_v >= (int)(out.v.size() / 3) || greatest_vn >= (int)(out.vn.size() / 3) ||
        greatest_vt >= (int)(out.vt.size() / 2)) {
        err = "face index out of range in " + filename;
        return false;
    }
    return true;
}

}  // namespace ptb
