// image_io.cpp -- texture files for `map_Kd`: what the reference reads through stb_image
// (src/Model.cpp:81-128: stbi_load(..., STBI_rgb_alpha), then mirrored along y) -- PNG (all colour types, 8 / 16 bit,
// non-interlaced; zlib inflates the IDAT stream), TGA (types 2 / 3 / 10 / 11, 8 / 24 / 32 bit) and binary PPM.
// Output: RGBA8, BOTTOM row first (the reference's mirror), as orb_scene_add_mesh takes it.
// Own code: stb_image is a third-party library the reference vendors; only its observable result is matched
// (grey -> (g,g,g,255), palette + tRNS -> RGBA, 16-bit samples keep their high byte).
#include <zlib.h>

#include <cctype>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "scene_io.h"

namespace orbhost {
namespace {

bool slurp(const std::string& path, std::string& out) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::ostringstream ss;
    ss << f.rdbuf();
    out = ss.str();
    return true;
}

uint32_t be32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// rows arrive top-down; the caller wants the bottom row first
void put_flipped(std::vector<uint8_t>& rgba, int w, int h, int y_top, const uint8_t* row_rgba) {
    memcpy(&rgba[4 * (size_t)(h - 1 - y_top) * w], row_rgba, 4 * (size_t)w);
}

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool load_png(const std::string& s, std::vector<uint8_t>& rgba, int& w, int& h) {
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (s.size() < 33 || memcmp(s.data(), sig, 8) != 0) return false;
    const unsigned char* d = (const unsigned char*)s.data();
    size_t p = 8;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    bool have_hdr = false;
    while (p + 12 <= s.size()) {
        const uint32_t len = be32(d + p);
        const char* type = (const char*)d + p + 4;
        if (p + 12 + (size_t)len > s.size()) return false;
        const unsigned char* body = d + p + 8;
        if (!memcmp(type, "IHDR", 4) && len >= 13) {
            w = (int)be32(body);
            h = (int)be32(body + 4);
            depth = body[8];
            ctype = body[9];
            interlace = body[12];
            have_hdr = true;
        } else if (!memcmp(type, "PLTE", 4)) {
            plte.assign(body, body + len);
        } else if (!memcmp(type, "tRNS", 4)) {
            trns.assign(body, body + len);
        } else if (!memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!memcmp(type, "IEND", 4)) {
            break;
        }
        p += 12 + (size_t)len;
    }
    if (!have_hdr || w <= 0 || h <= 0 || (size_t)w * h > ((size_t)1 << 28) || interlace != 0) return false;
    if (depth != 8 && depth != 16 && !(ctype == 3 || ctype == 0)) return false;
    if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels || (ctype == 3 && plte.size() < 3)) return false;
    const int bpp_bits = channels * depth;
    const size_t stride = ((size_t)w * bpp_bits + 7) / 8, bpp = (size_t)(bpp_bits + 7) / 8;
    std::vector<unsigned char> raw((stride + 1) * (size_t)h);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size()) return false;
    rgba.assign((size_t)4 * w * h, 255);
    std::vector<unsigned char> prev(stride, 0), cur(stride);
    std::vector<uint8_t> out_row((size_t)4 * w);
    for (int y = 0; y < h; y++) {
        const unsigned char* in = &raw[(stride + 1) * (size_t)y];
        const int filter = in[0];
        for (size_t i = 0; i < stride; i++) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int v = in[1 + i];
            switch (filter) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: return false;
            }
            cur[i] = (unsigned char)v;
        }
        for (int x = 0; x < w; x++) {
            uint8_t* o = &out_row[4 * (size_t)x];
            auto sample = [&](int ch) -> int {  // 8-bit value of channel ch of pixel x (16-bit: the high byte)
                if (depth == 8) return cur[(size_t)x * channels + ch];
                if (depth == 16) return cur[((size_t)x * channels + ch) * 2];
                const int bit = x * depth, v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
                return ctype == 3 ? v : v * 255 / ((1 << depth) - 1);
            };
            if (ctype == 3) {
                const size_t k = (size_t)sample(0);
                if (3 * k + 2 >= plte.size()) return false;
                o[0] = plte[3 * k], o[1] = plte[3 * k + 1], o[2] = plte[3 * k + 2];
                o[3] = k < trns.size() ? trns[k] : 255;
            } else if (ctype == 0 || ctype == 4) {
                o[0] = o[1] = o[2] = (uint8_t)sample(0);
                o[3] = ctype == 4 ? (uint8_t)sample(1) : 255;
                if (ctype == 0 && trns.size() >= 2 && depth == 8 && o[0] == trns[1]) o[3] = 0;
            } else {
                o[0] = (uint8_t)sample(0), o[1] = (uint8_t)sample(1), o[2] = (uint8_t)sample(2);
                o[3] = ctype == 6 ? (uint8_t)sample(3) : 255;
                if (ctype == 2 && trns.size() >= 6 && depth == 8 && o[0] == trns[1] && o[1] == trns[3] && o[2] == trns[5]) o[3] = 0;
            }
        }
        put_flipped(rgba, w, h, y, out_row.data());
        prev.swap(cur);
    }
    return true;
}

bool load_tga(const std::string& s, std::vector<uint8_t>& rgba, int& w, int& h) {
    if (s.size() < 18) return false;
    const unsigned char* d = (const unsigned char*)s.data();
    const int id_len = d[0], cmap = d[1], type = d[2], bits = d[16], desc = d[17];
    w = d[12] | (d[13] << 8);
    h = d[14] | (d[15] << 8);
    if (cmap != 0 || w <= 0 || h <= 0) return false;
    const bool rle = type == 10 || type == 11, grey = type == 3 || type == 11;
    if (!(type == 2 || type == 3 || rle)) return false;
    if (!((grey && bits == 8) || (!grey && (bits == 24 || bits == 32)))) return false;
    const int bpp = bits / 8;
    size_t p = 18 + (size_t)id_len;
    std::vector<unsigned char> px((size_t)w * h * bpp);
    if (!rle) {
        if (s.size() < p + px.size()) return false;
        memcpy(px.data(), d + p, px.size());
    } else {
        size_t o = 0;
        while (o < px.size()) {
            if (p >= s.size()) return false;
            const int c = d[p++], n = (c & 127) + 1;
            if (c & 128) {
                if (p + bpp > s.size()) return false;
                for (int k = 0; k < n && o < px.size(); k++, o += bpp) memcpy(&px[o], d + p, bpp);
                p += bpp;
            } else {
                if (p + (size_t)n * bpp > s.size() || o + (size_t)n * bpp > px.size()) return false;
                memcpy(&px[o], d + p, (size_t)n * bpp);
                p += (size_t)n * bpp;
                o += (size_t)n * bpp;
            }
        }
    }
    const bool top_down = (desc & 0x20) != 0;
    rgba.assign((size_t)4 * w * h, 255);
    std::vector<uint8_t> row((size_t)4 * w);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            const unsigned char* q = &px[((size_t)y * w + x) * bpp];
            uint8_t* o = &row[4 * (size_t)x];
            if (grey) { o[0] = o[1] = o[2] = q[0]; o[3] = 255; }
            else { o[0] = q[2]; o[1] = q[1]; o[2] = q[0]; o[3] = bpp == 4 ? q[3] : 255; }  // stored BGR(A)
        }
        put_flipped(rgba, w, h, top_down ? y : h - 1 - y, row.data());
    }
    return true;
}

}  // namespace

// RGBA8, bottom row first.  Formats by content (PNG signature, "P6"), TGA by its extension (it has no signature).
bool load_image_rgba8(const std::string& path, std::vector<uint8_t>& rgba, int& w, int& h) {
    std::string s;
    if (!slurp(path, s) || s.size() < 4) return false;
    if ((unsigned char)s[0] == 0x89 && s[1] == 'P') return load_png(s, rgba, w, h);
    if (s[0] == 'P' && s[1] == '6') return load_ppm_rgba8(path, rgba, w, h);
    std::string ext = path.size() >= 4 ? path.substr(path.size() - 4) : "";
    for (auto& c : ext) c = (char)tolower((unsigned char)c);
    if (ext == ".tga") return load_tga(s, rgba, w, h);
    return false;
}

}  // namespace orbhost
