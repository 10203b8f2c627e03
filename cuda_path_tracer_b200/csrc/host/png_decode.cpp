// png_decode.cpp -- PNG (zlib / deflate + unfiltering) and run-length TGA decoders for the scene loader.
// The reference reads textures through stb_image 2.06 (Scene::loadTexture, src/scene.cpp:130-181: stbi_load with the file's
// own channel count); a drop-in loader has to return the same pixels and the same channel count for the formats real assets
// use.  Written from the format specifications (RFC 1950 / 1951, PNG 1.2, Truevision TGA 2.0); the conventions that are
// stb_image's rather than the formats' are marked "stb".  Pinned against the reference's stb_image by
// tests/test_host.py::test_texture_decoders_match_stb.
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace ptb {

// ------------------------------------------------------------------------------------------------ inflate (RFC 1951)
namespace {

struct Bits {
    const uint8_t* p; size_t n, pos = 0; uint32_t buf = 0; int cnt = 0; bool eof = false;
    Bits(const uint8_t* p_, size_t n_) : p(p_), n(n_) {}
    uint32_t get(int k) {                       // k <= 16 bits, least significant first
        while (cnt < k) {
            if (pos >= n) { eof = true; return 0; }
            buf |= (uint32_t)p[pos++] << cnt; cnt += 8;
        }
        uint32_t v = buf & ((1u << k) - 1u);
        buf >>= k; cnt -= k;
        return v;
    }
    void align() { buf = 0; cnt = 0; }
};

struct Huffman {                                // canonical code: symbols ordered by (length, value)
    uint16_t count[16]; uint16_t symbol[320];
    bool build(const uint8_t* len, int n) {
        memset(count, 0, sizeof(count));
        for (int i = 0; i < n; i++) count[len[i]]++;
        count[0] = 0;
        int left = 1;
        for (int l = 1; l < 16; l++) { left = (left << 1) - count[l]; if (left < 0) return false; }   // over-subscribed
        uint16_t offs[16]; offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
        for (int i = 0; i < n; i++) if (len[i]) symbol[offs[len[i]]++] = (uint16_t)i;
        return true;
    }
    int decode(Bits& b) const {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l < 16; l++) {
            code |= (int)b.get(1);
            if (b.eof) return -1;
            int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

bool inflate_block(Bits& b, const Huffman& lit, const Huffman& dist, std::vector<uint8_t>& out) {
    for (;;) {
        int s = lit.decode(b);
        if (s < 0) return false;
        if (s < 256) { out.push_back((uint8_t)s); continue; }
        if (s == 256) return true;
        s -= 257;
        if (s >= 29) return false;
        size_t len = kLenBase[s] + b.get(kLenExtra[s]);
        int d = dist.decode(b);
        if (d < 0 || d >= 30) return false;
        size_t back = kDistBase[d] + b.get(kDistExtra[d]);
        if (b.eof || back > out.size()) return false;
        size_t from = out.size() - back;
        for (size_t i = 0; i < len; i++) out.push_back(out[from + i]);       // may overlap: byte by byte
    }
}

bool inflate_zlib(const uint8_t* data, size_t n, std::vector<uint8_t>& out, std::string& err) {
    if (n < 2 || (data[0] & 15) != 8 || ((data[0] << 8) | data[1]) % 31 != 0 || (data[1] & 32)) { err = "bad zlib header"; return false; }
    Bits b(data + 2, n - 2);
    for (;;) {
        uint32_t last = b.get(1), type = b.get(2);
        if (b.eof) { err = "deflate stream truncated"; return false; }
        if (type == 0) {
            b.align();
            if (b.pos + 4 > b.n) { err = "deflate stream truncated"; return false; }
            uint32_t len = b.p[b.pos] | (b.p[b.pos + 1] << 8), nlen = b.p[b.pos + 2] | (b.p[b.pos + 3] << 8);
            b.pos += 4;
            if ((len ^ 0xffffu) != nlen || b.pos + len > b.n) { err = "bad stored block"; return false; }
            out.insert(out.end(), b.p + b.pos, b.p + b.pos + len);
            b.pos += len;
        } else if (type == 1 || type == 2) {
            Huffman lit, dist;
            uint8_t len[320];
            if (type == 1) {
                for (int i = 0; i < 288; i++) len[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
                lit.build(len, 288);
                for (int i = 0; i < 30; i++) len[i] = 5;
                dist.build(len, 30);
            } else {
                int nlit = (int)b.get(5) + 257, ndist = (int)b.get(5) + 1, ncode = (int)b.get(4) + 4;
                static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                uint8_t cl[19] = {0};
                for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)b.get(3);
                Huffman code;
                if (nlit > 286 || ndist > 30 || !code.build(cl, 19)) { err = "bad dynamic block header"; return false; }
                int i = 0;
                while (i < nlit + ndist) {
                    int s = code.decode(b);
                    if (s < 0) { err = "bad code lengths"; return false; }
                    if (s < 16) { len[i++] = (uint8_t)s; continue; }
                    int rep, val = 0;
                    if (s == 16) { if (i == 0) { err = "bad code lengths"; return false; } val = len[i - 1]; rep = 3 + (int)b.get(2); }
                    else if (s == 17) rep = 3 + (int)b.get(3);
                    else rep = 11 + (int)b.get(7);
                    if (i + rep > nlit + ndist) { err = "bad code lengths"; return false; }
                    while (rep--) len[i++] = (uint8_t)val;
                }
                if (!lit.build(len, nlit) || !dist.build(len + nlit, ndist)) { err = "bad Huffman tables"; return false; }
            }
            if (!inflate_block(b, lit, dist, out)) { err = "corrupt deflate data"; return false; }
        } else { err = "bad deflate block type"; return false; }
        if (last) return true;
    }
}

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | (p[1] << 16) | (p[2] << 8) | p[3]; }
inline int paeth(int a, int b, int c) {
    int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

// ------------------------------------------------------------------------------------------------ PNG
// Channel count as stbi_load(..., 0) reports it: grey 1, grey + alpha 2, RGB 3, RGBA 4; a palette expands to RGB (RGBA with
// tRNS); tRNS on grey / RGB adds an alpha channel (0 where the pixel equals the key, else 255); 1/2/4-bit grey is scaled to
// 0..255 (stb).  16-bit samples and Adam7 interlacing are rejected (stb 2.06 rejects the former too).
bool decode_png(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (d.size() < 8 + 25 || memcmp(d.data(), sig, 8) != 0) { err = "not a PNG file"; return false; }
    size_t p = 8;
    int depth = 0, color = -1, interlace = 0;
    std::vector<uint8_t> idat, palette;
    uint8_t pal_alpha[256]; memset(pal_alpha, 255, sizeof(pal_alpha));
    bool has_trans = false, pal_trans = false;
    uint8_t key[3] = {0, 0, 0};
    static const uint8_t depth_scale[9] = {0, 0xff, 0x55, 0, 0x11, 0, 0, 0, 0x01};
    bool end = false;
    while (!end && p + 12 <= d.size()) {
        uint32_t len = be32(&d[p]);
        const uint8_t* type = &d[p + 4];
        const uint8_t* body = &d[p + 8];
        if (p + 12 + (size_t)len > d.size()) { err = "PNG chunk truncated"; return false; }
        if (!memcmp(type, "IHDR", 4)) {
            if (len != 13) { err = "bad IHDR"; return false; }
            w = (int)be32(body); h = (int)be32(body + 4);
            depth = body[8]; color = body[9]; interlace = body[12];
            if (w <= 0 || h <= 0 || w > (1 << 24) || h > (1 << 24)) { err = "bad PNG size"; return false; }
            if (depth == 16) { err = "unsupported: 16-bit PNG"; return false; }
            if (depth != 1 && depth != 2 && depth != 4 && depth != 8) { err = "bad PNG bit depth"; return false; }
            if (color != 0 && color != 2 && color != 3 && color != 4 && color != 6) { err = "bad PNG colour type"; return false; }
            if (body[10] != 0 || body[11] != 0) { err = "bad PNG compression / filter method"; return false; }
            if (interlace) { err = "unsupported: interlaced PNG"; return false; }
        } else if (!memcmp(type, "PLTE", 4)) {
            if (len > 768 || len % 3) { err = "bad PLTE"; return false; }
            palette.assign(body, body + len);
        } else if (!memcmp(type, "tRNS", 4)) {
            if (color == 3) {
                if (len > 256) { err = "bad tRNS"; return false; }
                memcpy(pal_alpha, body, len); pal_trans = true;
            } else if (color == 0 || color == 2) {
                int n = color == 0 ? 1 : 3;
                if (len != (uint32_t)(2 * n)) { err = "bad tRNS"; return false; }
                for (int k = 0; k < n; k++) key[k] = (uint8_t)(body[2 * k + 1] * depth_scale[depth]);      // low byte, scaled (stb)
                has_trans = true;
            }
        } else if (!memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!memcmp(type, "IEND", 4)) {
            end = true;
        }
        p += 12 + (size_t)len;
    }
    if (color < 0 || idat.empty()) { err = "PNG without IHDR / IDAT"; return false; }
    if (color == 3 && palette.empty()) { err = "paletted PNG without PLTE"; return false; }
    const int samples = color == 0 ? 1 : color == 2 ? 3 : color == 3 ? 1 : color == 4 ? 2 : 4;      // per pixel, in the file
    const size_t row_bytes = ((size_t)w * samples * depth + 7) / 8;
    const int bpp = (samples * depth + 7) / 8;                                                         // filter distance, >= 1
    std::vector<uint8_t> raw;
    raw.reserve((row_bytes + 1) * h);
    if (!inflate_zlib(idat.data(), idat.size(), raw, err)) { err = "PNG: " + err; return false; }
    if (raw.size() < (row_bytes + 1) * (size_t)h) { err = "PNG: not enough pixel data"; return false; }
    // unfilter in place (PNG 1.2 section 6); `prev` is the reconstructed row above
    std::vector<uint8_t> zero(row_bytes, 0);
    for (int y = 0; y < h; y++) {
        uint8_t* row = &raw[(row_bytes + 1) * (size_t)y];
        const int filter = row[0];
        uint8_t* cur = row + 1;
        const uint8_t* prev = y ? row - row_bytes : zero.data();
        if (filter > 4) { err = "PNG: bad filter type"; return false; }
        for (size_t i = 0; i < row_bytes; i++) {
            const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= (size_t)bpp ? prev[i - bpp] : 0;
            int x = cur[i];
            switch (filter) {
                case 1: x += a; break;
                case 2: x += b; break;
                case 3: x += (a + b) >> 1; break;
                case 4: x += paeth(a, b, c); break;
                default: break;
            }
            cur[i] = (uint8_t)x;
        }
    }
    // expand to 8-bit samples, apply palette / colour key
    const int base = color == 3 ? 3 : samples;
    ch = color == 3 ? (pal_trans ? 4 : 3) : samples + (has_trans ? 1 : 0);
    px.resize((size_t)w * h * ch);
    const int mask = (1 << depth) - 1;
    for (int y = 0; y < h; y++) {
        const uint8_t* cur = &raw[(row_bytes + 1) * (size_t)y + 1];
        uint8_t* o = &px[(size_t)y * w * ch];
        for (int x = 0; x < w; x++, o += ch) {
            uint8_t s[4];
            for (int k = 0; k < samples; k++) {
                const size_t bit = ((size_t)x * samples + k) * depth;
                int v = depth == 8 ? cur[bit >> 3] : (cur[bit >> 3] >> (8 - depth - (int)(bit & 7))) & mask;
                s[k] = (uint8_t)(color == 3 || depth == 8 ? v : v * depth_scale[depth]);
            }
            if (color == 3) {
                if ((size_t)s[0] * 3 + 2 >= palette.size()) { o[0] = o[1] = o[2] = 0; if (ch == 4) o[3] = 255; continue; }
                o[0] = palette[s[0] * 3]; o[1] = palette[s[0] * 3 + 1]; o[2] = palette[s[0] * 3 + 2];
                if (ch == 4) o[3] = pal_alpha[s[0]];
            } else {
                for (int k = 0; k < base; k++) o[k] = s[k];
                if (has_trans) {
                    bool same = true;
                    for (int k = 0; k < base; k++) same = same && s[k] == key[k];
                    o[base] = same ? 0 : 255;
                }
            }
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------------ TGA, run-length (types 10, 11)
// Packets of 1..128 pixels: header bit 7 set = one pixel value repeated, clear = that many literal pixels; packets may cross
// scanlines.  Output order and BGR(A) -> RGB(A) swap as the uncompressed decoder (image_io.cpp).
bool decode_tga_rle(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err) {
    if (d.size() < 18) { err = "TGA truncated"; return false; }
    const int idlen = d[0], cmap = d[1], type = d[2], bpp = d[16], desc = d[17];
    if (cmap != 0 || (type != 10 && type != 11)) { err = "unsupported: colour-mapped TGA"; return false; }
    w = d[12] | (d[13] << 8); h = d[14] | (d[15] << 8);
    const int bytes = bpp / 8;
    if ((type == 10 && bytes != 3 && bytes != 4) || (type == 11 && bytes != 1)) { err = "unsupported: TGA pixel depth"; return false; }
    ch = bytes;
    const size_t total = (size_t)w * h;
    std::vector<uint8_t> lin(total * bytes);
    size_t p = 18 + (size_t)idlen, i = 0;
    while (i < total) {
        if (p >= d.size()) { err = "TGA truncated"; return false; }
        const int head = d[p++];
        size_t count = (size_t)(head & 127) + 1;
        if (i + count > total) count = total - i;
        if (head & 128) {
            if (p + bytes > d.size()) { err = "TGA truncated"; return false; }
            for (size_t k = 0; k < count; k++) memcpy(&lin[(i + k) * bytes], &d[p], bytes);
            p += bytes;
        } else {
            if (p + count * bytes > d.size()) { err = "TGA truncated"; return false; }
            memcpy(&lin[i * bytes], &d[p], count * bytes);
            p += count * bytes;
        }
        i += count;
    }
    const bool top_down = (desc & 0x20) != 0;
    px.resize(total * ch);
    for (int y = 0; y < h; y++) {
        const uint8_t* row = &lin[(size_t)(top_down ? y : (h - 1 - y)) * w * bytes];
        for (int x = 0; x < w; x++) {
            const uint8_t* s = row + (size_t)x * bytes;
            uint8_t* o = &px[((size_t)y * w + x) * ch];
            if (bytes == 1) o[0] = s[0];
            else { o[0] = s[2]; o[1] = s[1]; o[2] = s[0]; if (bytes == 4) o[3] = s[3]; }
        }
    }
    return true;
}

}  // namespace ptb
