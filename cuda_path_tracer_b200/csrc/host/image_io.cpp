// image_io.cpp -- decoders for the texture / environment formats of the procedural scenes.
// The reference decodes through stb_image 2.06 (src/scene.cpp:140 stbi_load, :187 stbi_loadf).
// Of the formats stb reads, the headless generators only emit the uncompressed ones, so this
// file restates just those: binary PNM (P5/P6, maxval 255), BMP (24/32 bpp BI_RGB), TGA
// (type 2/3, 8/24/32 bpp) and Radiance RGBE .hdr (flat and new-style RLE scanlines).
// RGBE -> float follows stb: value = mantissa * 2^(e - 136), all-zero when e == 0.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cctype>
#include <cstring>
#include <string>
#include <vector>

#include "scene.hpp"

namespace ptb {

static bool read_file(const std::string& path, std::vector<uint8_t>& data) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    data.resize(n > 0 ? (size_t)n : 0);
    size_t got = n > 0 ? fread(data.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == data.size();
}

static bool pnm_token(const std::vector<uint8_t>& d, size_t& p, int& value) {
    for (;;) {
        while (p < d.size() && (d[p] == ' ' || d[p] == '\t' || d[p] == '\n' || d[p] == '\r')) p++;
        if (p < d.size() && d[p] == '#') { while (p < d.size() && d[p] != '\n' && d[p] != '\r') p++; continue; }
        break;
    }
    if (p >= d.size() || d[p] < '0' || d[p] > '9') return false;
    value = 0;
    while (p < d.size() && d[p] >= '0' && d[p] <= '9') value = value * 10 + (d[p++] - '0');
    return true;
}

static bool decode_pnm(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err) {
    ch = (d[1] == '5') ? 1 : 3;
    size_t p = 2;
    int maxv = 0;
    if (!pnm_token(d, p, w) || !pnm_token(d, p, h) || !pnm_token(d, p, maxv)) { err = "bad PNM header"; return false; }
    if (maxv > 255) { err = "PNM max value > 255 not supported"; return false; }
    p++;  // single whitespace after maxval
    size_t need = (size_t)w * h * ch;
    if (p + need > d.size()) { err = "PNM truncated"; return false; }
    px.assign(d.begin() + p, d.begin() + p + need);
    return true;
}

static uint32_t le32(const uint8_t* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint16_t le16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

static bool decode_bmp(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err) {
    if (d.size() < 54) { err = "BMP truncated"; return false; }
    uint32_t off = le32(&d[10]);
    uint32_t hsz = le32(&d[14]);
    if (hsz < 40) { err = "BMP core header not supported"; return false; }
    int32_t bw = (int32_t)le32(&d[18]), bh = (int32_t)le32(&d[22]);
    int bpp = le16(&d[28]);
    uint32_t comp = le32(&d[30]);
    if ((bpp != 24 && bpp != 32) || (comp != 0 && comp != 3)) { err = "BMP: only 24/32 bpp uncompressed"; return false; }
    bool flip = bh > 0;
    w = bw; h = bh > 0 ? bh : -bh;
    ch = (bpp == 32) ? 4 : 3;
    size_t stride = ((size_t)w * (bpp / 8) + 3) & ~(size_t)3;
    if (off + stride * h > d.size()) { err = "BMP truncated"; return false; }
    px.resize((size_t)w * h * ch);
    for (int y = 0; y < h; y++) {
        const uint8_t* row = &d[off + stride * (flip ? (h - 1 - y) : y)];
        for (int x = 0; x < w; x++) {
            const uint8_t* s = row + x * (bpp / 8);
            uint8_t* o = &px[((size_t)y * w + x) * ch];
            o[0] = s[2]; o[1] = s[1]; o[2] = s[0];
            if (ch == 4) o[3] = s[3];
        }
    }
    return true;
}

static bool decode_tga(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err) {
    if (d.size() < 18) { err = "TGA truncated"; return false; }
    int idlen = d[0], cmap = d[1], type = d[2], bpp = d[16], desc = d[17];
    if (cmap != 0 || (type != 2 && type != 3)) { err = "TGA: only uncompressed true-colour / grey"; return false; }
    w = le16(&d[12]); h = le16(&d[14]);
    int bytes = bpp / 8;
    if (bytes != 1 && bytes != 3 && bytes != 4) { err = "TGA: unsupported depth"; return false; }
    ch = bytes;
    size_t off = 18 + (size_t)idlen;
    if (off + (size_t)w * h * bytes > d.size()) { err = "TGA truncated"; return false; }
    bool top_down = (desc & 0x20) != 0;
    px.resize((size_t)w * h * ch);
    for (int y = 0; y < h; y++) {
        const uint8_t* row = &d[off + (size_t)(top_down ? y : (h - 1 - y)) * w * bytes];
        for (int x = 0; x < w; x++) {
            const uint8_t* s = row + x * bytes;
            uint8_t* o = &px[((size_t)y * w + x) * ch];
            if (bytes == 1) o[0] = s[0];
            else { o[0] = s[2]; o[1] = s[1]; o[2] = s[0]; if (bytes == 4) o[3] = s[3]; }
        }
    }
    return true;
}

// png_decode.cpp
bool decode_png(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err);
bool decode_tga_rle(const std::vector<uint8_t>& d, int& w, int& h, int& ch, std::vector<uint8_t>& px, std::string& err);

// What stbi_load reads and this loader reads too: PNG (8-bit and below, non-interlaced), BMP 24/32, TGA true-colour / grey
// (plain and run-length), PNM P5/P6.  `err` starts with "cannot read" when the file itself is missing (the reference then
// prints a message and renders untextured, src/scene.cpp:141-146) and with "unsupported" when the file is there but in a
// format or variant this loader does not decode (JPEG, GIF, PSD, 16-bit / interlaced PNG, colour-mapped TGA): the scene
// loader turns that into PTB_ERR_UNSUPPORTED instead of silently rendering something else than the reference would.
bool load_image_u8(const std::string& path, int& w, int& h, int& channels, std::vector<uint8_t>& px, std::string& err) {
    std::vector<uint8_t> d;
    if (!read_file(path, d) || d.size() < 4) { err = "cannot read " + path; return false; }
    bool ok;
    if (d.size() >= 8 && d[0] == 0x89 && d[1] == 'P' && d[2] == 'N' && d[3] == 'G') ok = decode_png(d, w, h, channels, px, err);
    else if (d[0] == 'P' && (d[1] == '5' || d[1] == '6')) ok = decode_pnm(d, w, h, channels, px, err);
    else if (d[0] == 'B' && d[1] == 'M') ok = decode_bmp(d, w, h, channels, px, err);
    else {
        size_t dot = path.find_last_of('.');
        std::string ext = dot == std::string::npos ? "" : path.substr(dot);
        for (char& c : ext) c = (char)tolower((unsigned char)c);
        if (ext == ".tga") ok = (d.size() >= 3 && (d[2] == 10 || d[2] == 11)) ? decode_tga_rle(d, w, h, channels, px, err) : decode_tga(d, w, h, channels, px, err);
        else { err = "unsupported: image format of " + path + " (PNG, BMP, TGA and PNM are decoded; JPEG, GIF, PSD are not)"; return false; }
    }
    if (!ok && err.compare(0, 11, "unsupported") != 0) err = "unsupported: " + err + " (" + path + ")";
    return ok;
}

static void rgbe_to_float(const uint8_t* in, float* out) {
    if (in[3] != 0) {
        float f1 = (float)std::ldexp(1.0f, in[3] - (int)(128 + 8));
        out[0] = in[0] * f1; out[1] = in[1] * f1; out[2] = in[2] * f1;
    } else {
        out[0] = out[1] = out[2] = 0.0f;
    }
}

bool load_image_hdr(const std::string& path, int& w, int& h, std::vector<float>& rgb, std::string& err) {
    std::vector<uint8_t> d;
    if (!read_file(path, d)) { err = "cannot read " + path; return false; }
    size_t p = 0;
    auto next_line = [&](std::string& s) {
        s.clear();
        while (p < d.size() && d[p] != '\n') s += (char)d[p++];
        if (p < d.size()) p++;
    };
    std::string line;
    next_line(line);
    if (line != "#?RADIANCE" && line != "#?RGBE") { err = "not a Radiance HDR file"; return false; }
    bool fmt_ok = false;
    for (;;) {
        next_line(line);
        if (line.empty()) break;
        if (line == "FORMAT=32-bit_rle_rgbe") fmt_ok = true;
        if (p >= d.size()) break;
    }
    if (!fmt_ok) { err = "HDR: unsupported format"; return false; }
    next_line(line);
    if (sscanf(line.c_str(), "-Y %d +X %d", &h, &w) != 2) { err = "HDR: unsupported data layout"; return false; }
    rgb.resize((size_t)w * h * 3);
    std::vector<uint8_t> scan((size_t)w * 4);
    for (int y = 0; y < h; y++) {
        bool rle = false;
        if (w >= 8 && w < 32768 && p + 4 <= d.size() && d[p] == 2 && d[p + 1] == 2 && !(d[p + 2] & 0x80)) {
            int len = (d[p + 2] << 8) | d[p + 3];
            if (len == w) rle = true;
        }
        if (!rle) {
            if (p + (size_t)w * 4 > d.size()) { err = "HDR truncated"; return false; }
            // flat file: stb switches to flat decoding for the whole image when the first
            // scanline is not RLE; mixed files are not produced by the generators
            for (int x = 0; x < w; x++) rgbe_to_float(&d[p + (size_t)x * 4], &rgb[((size_t)y * w + x) * 3]);
            p += (size_t)w * 4;
            continue;
        }
        p += 4;
        for (int k = 0; k < 4; k++) {
            int i = 0;
            while (i < w) {
                if (p >= d.size()) { err = "HDR truncated"; return false; }
                int count = d[p++];
                if (count > 128) {
                    count -= 128;
                    if (p >= d.size() || i + count > w) { err = "HDR corrupt"; return false; }
                    uint8_t v = d[p++];
                    for (int z = 0; z < count; z++) scan[(size_t)(i++) * 4 + k] = v;
                } else {
                    if (p + count > d.size() || i + count > w) { err = "HDR corrupt"; return false; }
                    for (int z = 0; z < count; z++) scan[(size_t)(i++) * 4 + k] = d[p++];
                }
            }
        }
        for (int x = 0; x < w; x++) rgbe_to_float(&scan[(size_t)x * 4], &rgb[((size_t)y * w + x) * 3]);
    }
    return true;
}

}  // namespace ptb

// ---- writers (the reference's Image::savePNG / saveHDR, src/image.cpp:22-49, via stb_image_write) ---------------
namespace ptb {

static uint32_t crc32_of(const uint8_t* p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        ready = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}

static void be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }

static void png_chunk(std::vector<uint8_t>& out, const char* type, const std::vector<uint8_t>& data) {
    be32(out, (uint32_t)data.size());
    size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    be32(out, crc32_of(out.data() + start, out.size() - start));
}

// 8-bit RGB PNG, one filter-0 scanline per row, zlib stream of stored (uncompressed) deflate blocks: valid for every
// PNG reader and needs no compression library.
bool save_png_rgb8(const std::string& path, const uint8_t* rgb, int w, int h, std::string& err) {
    if (w <= 0 || h <= 0) { err = "empty image"; return false; }
    std::vector<uint8_t> raw;
    raw.reserve((size_t)h * (3 * (size_t)w + 1));
    for (int y = 0; y < h; y++) {
        raw.push_back(0);
        raw.insert(raw.end(), rgb + 3 * (size_t)w * y, rgb + 3 * (size_t)w * (y + 1));
    }
    std::vector<uint8_t> z;
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;   // adler32
    size_t pos = 0;
    while (pos < raw.size()) {
        size_t n = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back(n & 0xff); z.push_back(n >> 8); z.push_back(~n & 0xff); z.push_back((~n >> 8) & 0xff);
        for (size_t i = 0; i < n; i++) { a = (a + raw[pos + i]) % 65521; b = (b + a) % 65521; }
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        pos += n;
    }
    be32(z, (b << 16) | a);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    png_chunk(out, "IHDR", ihdr);
    png_chunk(out, "IDAT", z);
    png_chunk(out, "IEND", {});
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { err = "cannot open " + path; return false; }
    bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
    fclose(f);
    if (!ok) err = "short write to " + path;
    return ok;
}

// Radiance .hdr, flat RGBE scanlines (stb_image_write's float -> RGBE rule: frexp of the largest channel).
bool save_hdr_rgb32f(const std::string& path, const float* rgb, int w, int h, std::string& err) {
    if (w <= 0 || h <= 0) { err = "empty image"; return false; }
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { err = "cannot open " + path; return false; }
    fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
    std::vector<uint8_t> row(4 * (size_t)w);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            const float* p = rgb + 3 * ((size_t)y * w + x);
            float m = std::max(p[0], std::max(p[1], p[2]));
            uint8_t* o = &row[4 * (size_t)x];
            if (!(m >= 1e-32f)) { o[0] = o[1] = o[2] = o[3] = 0; continue; }
            int e;
            float scale = std::frexp(m, &e) * 256.0f / m;
            o[0] = (uint8_t)(p[0] * scale); o[1] = (uint8_t)(p[1] * scale); o[2] = (uint8_t)(p[2] * scale);
            o[3] = (uint8_t)(e + 128);
        }
        if (fwrite(row.data(), 1, row.size(), f) != row.size()) { fclose(f); err = "short write to " + path; return false; }
    }
    fclose(f);
    return true;
}

}  // namespace ptb
