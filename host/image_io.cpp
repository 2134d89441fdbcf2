// image_io.cpp -- texture files for `map_Kd`: what the reference reads through stb_image
// (src/Model.cpp:81-128: stbi_load(..., STBI_rgb_alpha), then mirrored along y) -- PNG (all colour types and bit depths,
// Adam7 interlace; zlib inflates the IDAT stream), JPEG (baseline and progressive Huffman), TGA, BMP and binary PPM.
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
    if (!have_hdr || w <= 0 || h <= 0 || (size_t)w * h > ((size_t)1 << 28) || interlace > 1) return false;
    if (depth != 8 && depth != 16 && !(ctype == 3 || ctype == 0)) return false;
    if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels || (ctype == 3 && plte.size() < 3)) return false;
    const int bpp_bits = channels * depth;
    const size_t bpp = (size_t)(bpp_bits + 7) / 8;
    // one pass (the whole image) or the seven passes of Adam7: pass k holds the pixels (x0 + i dx, y0 + j dy)
    static const int ax0[7] = {0, 4, 0, 2, 0, 1, 0}, ay0[7] = {0, 0, 4, 0, 2, 0, 1};
    static const int adx[7] = {8, 8, 4, 4, 2, 2, 1}, ady[7] = {8, 8, 8, 4, 4, 2, 2};
    const int n_pass = interlace ? 7 : 1;
    int pw[7], ph[7];
    size_t total = 0;
    for (int k = 0; k < n_pass; k++) {
        pw[k] = interlace ? (w - ax0[k] + adx[k] - 1) / adx[k] : w;
        ph[k] = interlace ? (h - ay0[k] + ady[k] - 1) / ady[k] : h;
        if (pw[k] > 0 && ph[k] > 0) total += (((size_t)pw[k] * bpp_bits + 7) / 8 + 1) * (size_t)ph[k];
    }
    std::vector<unsigned char> raw(total);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size()) return false;
    rgba.assign((size_t)4 * w * h, 255);
    size_t off = 0;
    for (int k = 0; k < n_pass; k++) {
        if (pw[k] <= 0 || ph[k] <= 0) continue;
        const size_t stride = ((size_t)pw[k] * bpp_bits + 7) / 8;
        const int x0 = interlace ? ax0[k] : 0, y0 = interlace ? ay0[k] : 0, dx = interlace ? adx[k] : 1, dy = interlace ? ady[k] : 1;
        std::vector<unsigned char> prev(stride, 0), cur(stride);
        for (int y = 0; y < ph[k]; y++) {
            const unsigned char* in = &raw[off];
            off += stride + 1;
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
            const int Y = y0 + y * dy;
            for (int x = 0; x < pw[k]; x++) {
                uint8_t* o = &rgba[4 * ((size_t)(h - 1 - Y) * w + (size_t)(x0 + x * dx))];  // bottom row first
                auto sample = [&](int ch) -> int {  // 8-bit value of channel ch of pixel x (16-bit: the high byte)
                    if (depth == 8) return cur[(size_t)x * channels + ch];
                    if (depth == 16) return cur[((size_t)x * channels + ch) * 2];
                    const int bit = x * depth, v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
                    return ctype == 3 ? v : v * 255 / ((1 << depth) - 1);
                };
                if (ctype == 3) {
                    const size_t q = (size_t)sample(0);
                    if (3 * q + 2 >= plte.size()) return false;
                    o[0] = plte[3 * q], o[1] = plte[3 * q + 1], o[2] = plte[3 * q + 2];
                    o[3] = q < trns.size() ? trns[q] : 255;
                } else if (ctype == 0 || ctype == 4) {
                    o[0] = o[1] = o[2] = (uint8_t)sample(0);
                    o[3] = ctype == 4 ? (uint8_t)sample(1) : 255;
                    if (ctype == 0 && trns.size() >= 2) {  // the one transparent grey level: compared at full depth
                        if (depth == 16) {
                            if (cur[(size_t)x * 2] == trns[0] && cur[(size_t)x * 2 + 1] == trns[1]) o[3] = 0;
                        } else if (o[0] == (uint8_t)(trns[1] * (depth == 8 ? 1 : 255 / ((1 << depth) - 1)))) {
                            o[3] = 0;
                        }
                    }
                } else {
                    o[0] = (uint8_t)sample(0), o[1] = (uint8_t)sample(1), o[2] = (uint8_t)sample(2);
                    o[3] = ctype == 6 ? (uint8_t)sample(3) : 255;
                    if (ctype == 2 && trns.size() >= 6) {  // the one transparent colour
                        if (depth == 16) {
                            if (!memcmp(&cur[(size_t)x * 6], trns.data(), 6)) o[3] = 0;
                        } else if (o[0] == trns[1] && o[1] == trns[3] && o[2] == trns[5]) {
                            o[3] = 0;
                        }
                    }
                }
            }
            prev.swap(cur);
        }
    }
    return true;
}

// TGA: true colour (15 / 16 / 24 / 32 bit), grey (8 bit, or 16 = grey + alpha) and colour-mapped (8- or 16-bit indices into
// a 15 / 16 / 24 / 32-bit palette), raw or run-length encoded, bottom-up or top-down.  16-bit colours are 5-5-5
// (c * 255 / 31 per channel); an index past the palette reads entry 0.
bool load_tga(const std::string& s, std::vector<uint8_t>& rgba, int& w, int& h) {
    if (s.size() < 18) return false;
    const unsigned char* d = (const unsigned char*)s.data();
    const int id_len = d[0], indexed = d[1], bits = d[16], desc = d[17];
    int type = d[2];
    const int pal_start = d[3] | (d[4] << 8), pal_len = d[5] | (d[6] << 8), pal_bits = d[7];
    w = d[12] | (d[13] << 8);
    h = d[14] | (d[15] << 8);
    if (w <= 0 || h <= 0 || indexed > 1) return false;
    const bool rle = type >= 8;
    if (rle) type -= 8;
    if (!(type == 1 || type == 2 || type == 3) || (type == 1) != (indexed == 1)) return false;
    // components of a colour of `b` bits: 1 grey, 2 grey + alpha, 3 / 4 colour; rgb16 = packed 5-5-5
    bool rgb16 = false;
    auto comps = [&](int b, bool grey) {
        if (b == 8) return 1;
        if (b == 16 && grey) return 2;
        if (b == 15 || b == 16) { rgb16 = true; return 3; }
        return (b == 24 || b == 32) ? b / 8 : 0;
    };
    if (indexed && bits != 8 && bits != 16) return false;
    const int comp = indexed ? comps(pal_bits, false) : comps(bits, type == 3);
    if (!comp) return false;
    size_t p = 18 + (size_t)id_len;
    auto read16 = [&](uint8_t* o) {  // 5-5-5, already in R G B order
        const int px = d[p] | (d[p + 1] << 8);
        p += 2;
        o[0] = (uint8_t)((((px >> 10) & 31) * 255) / 31);
        o[1] = (uint8_t)((((px >> 5) & 31) * 255) / 31);
        o[2] = (uint8_t)(((px & 31) * 255) / 31);
    };
    std::vector<uint8_t> pal;
    if (indexed) {
        if (pal_len == 0) return false;
        p += (size_t)pal_start;
        pal.resize((size_t)pal_len * comp);
        const size_t entry = rgb16 ? 2 : (size_t)comp;
        if (p + (size_t)pal_len * entry > s.size()) return false;
        if (rgb16) {
            for (int i = 0; i < pal_len; i++) read16(&pal[(size_t)i * comp]);
        } else {
            memcpy(pal.data(), d + p, pal.size());
            p += pal.size();
        }
    }
    const size_t in_bytes = indexed ? (size_t)bits / 8 : (rgb16 ? 2 : (size_t)comp);
    std::vector<uint8_t> px((size_t)w * h * comp);
    uint8_t cur[4] = {0, 0, 0, 0};
    int run = 0;
    bool repeat = false, fresh = true;
    for (size_t i = 0; i < (size_t)w * h; i++) {
        if (rle) {
            if (run == 0) {
                if (p >= s.size()) return false;
                const int c = d[p++];
                run = (c & 127) + 1;
                repeat = (c & 128) != 0;
                fresh = true;
            } else if (!repeat) {
                fresh = true;
            }
        } else {
            fresh = true;
        }
        if (fresh) {
            if (p + in_bytes > s.size()) return false;
            if (indexed) {
                size_t k = bits == 8 ? d[p] : (size_t)(d[p] | (d[p + 1] << 8));
                p += in_bytes;
                if (k >= (size_t)pal_len) k = 0;
                memcpy(cur, &pal[k * comp], (size_t)comp);
            } else if (rgb16) {
                read16(cur);
            } else {
                memcpy(cur, d + p, (size_t)comp);
                p += (size_t)comp;
            }
            fresh = false;
        }
        memcpy(&px[i * comp], cur, (size_t)comp);
        --run;
    }
    const bool top_down = (desc & 0x20) != 0;
    rgba.assign((size_t)4 * w * h, 255);
    std::vector<uint8_t> row((size_t)4 * w);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            const uint8_t* q = &px[((size_t)y * w + x) * comp];
            uint8_t* o = &row[4 * (size_t)x];
            if (comp <= 2) { o[0] = o[1] = o[2] = q[0]; o[3] = comp == 2 ? q[1] : 255; }
            else if (rgb16) { o[0] = q[0]; o[1] = q[1]; o[2] = q[2]; o[3] = 255; }
            else { o[0] = q[2]; o[1] = q[1]; o[2] = q[0]; o[3] = comp == 4 ? q[3] : 255; }  // stored B G R (A)
        }
        put_flipped(rgba, w, h, top_down ? y : h - 1 - y, row.data());
    }
    return true;
}

// BMP: uncompressed 4 / 8-bit palette, 24-bit, and 16 / 32-bit with the default or explicit channel masks (info headers
// of 12, 40, 56, 108 and 124 bytes); run-length encoded and 1-bit files are refused, as the reference's decoder does.
// Its conventions, which decide texels: a channel of fewer than 8 bits is widened by repeating its bits (v + v >> n
// + ...) after an ARITHMETIC shift of the masked value taken as a signed int; a 32-bit file whose alpha bytes are all
// zero is opaque; a positive height means bottom-up rows.
bool load_bmp(const std::string& s, std::vector<uint8_t>& rgba, int& w, int& h) {
    const unsigned char* d = (const unsigned char*)s.data();
    const size_t n = s.size();
    size_t p = 0;
    auto g8 = [&]() -> unsigned { return p < n ? d[p++] : (p++, 0u); };  // past the end reads as zero
    auto g16 = [&]() -> unsigned { const unsigned a = g8(); return a | (g8() << 8); };
    auto g32 = [&]() -> unsigned { const unsigned a = g16(); return a | (g16() << 16); };
    if (g8() != 'B' || g8() != 'M') return false;
    g32(); g16(); g16();
    const int offset = (int)g32(), hsz = (int)g32();
    if (hsz != 12 && hsz != 40 && hsz != 56 && hsz != 108 && hsz != 124) return false;
    int hh;
    if (hsz == 12) { w = (int)g16(); hh = (int)g16(); }
    else { w = (int)g32(); hh = (int)g32(); }
    if (g16() != 1) return false;
    const int bpp = (int)g16();
    unsigned mr = 0, mg = 0, mb = 0, ma = 0, all_a = 255;
    if (bpp == 1) return false;
    if (hsz != 12) {
        const unsigned compress = g32();
        if (compress == 1 || compress == 2) return false;
        for (int i = 0; i < 5; i++) g32();  // image size, resolutions, colours used / important
        if (hsz == 40 || hsz == 56) {
            if (hsz == 56) for (int i = 0; i < 4; i++) g32();
            if (bpp == 16 || bpp == 32) {
                if (compress == 0) {
                    if (bpp == 32) { mr = 0xffu << 16; mg = 0xffu << 8; mb = 0xffu; ma = 0xffu << 24; all_a = 0; }
                    else { mr = 31u << 10; mg = 31u << 5; mb = 31u; }
                } else if (compress == 3) {
                    mr = g32(); mg = g32(); mb = g32();
                    if (mr == mg && mg == mb) return false;
                } else {
                    return false;
                }
            }
        } else {
            mr = g32(); mg = g32(); mb = g32(); ma = g32();
            for (int i = 0; i < 13; i++) g32();  // colour space
            if (hsz == 124) for (int i = 0; i < 4; i++) g32();
        }
    }
    const bool bottom_up = hh > 0;
    h = hh < 0 ? -hh : hh;
    if (w <= 0 || h <= 0 || (size_t)w * h > ((size_t)1 << 28)) return false;
    int psize = 0;
    if (hsz == 12) { if (bpp < 24) psize = (offset - 14 - 24) / 3; }
    else if (bpp < 16) psize = (offset - 14 - hsz) >> 2;
    std::vector<uint8_t> img((size_t)4 * w * h);  // top-down as stored, flipped below
    size_t z = 0;
    if (bpp < 16) {
        if (psize <= 0 || psize > 256 || (bpp != 4 && bpp != 8)) return false;
        uint8_t pal[256][3] = {};
        for (int i = 0; i < psize; i++) {
            pal[i][2] = (uint8_t)g8(); pal[i][1] = (uint8_t)g8(); pal[i][0] = (uint8_t)g8();
            if (hsz != 12) g8();
        }
        p += (size_t)(offset - 14 - hsz - psize * (hsz == 12 ? 3 : 4));
        const int width = bpp == 4 ? (w + 1) >> 1 : w, pad = (-width) & 3;
        for (int j = 0; j < h; j++) {
            for (int i = 0; i < w; i += 2) {
                unsigned v = g8(), v2 = 0;
                if (bpp == 4) { v2 = v & 15; v >>= 4; }
                memcpy(&img[z], pal[v], 3); img[z + 3] = 255; z += 4;
                if (i + 1 == w) break;
                v = bpp == 8 ? g8() : v2;
                memcpy(&img[z], pal[v], 3); img[z + 3] = 255; z += 4;
            }
            p += (size_t)pad;
        }
    } else {
        if (bpp != 16 && bpp != 24 && bpp != 32) return false;
        p += (size_t)(offset - 14 - hsz);
        const int width = bpp == 24 ? 3 * w : (bpp == 16 ? 2 * w : 0), pad = (-width) & 3;
        const int easy = bpp == 24 ? 1 : ((bpp == 32 && mb == 0xff && mg == 0xff00 && mr == 0x00ff0000 && ma == 0xff000000) ? 2 : 0);
        auto high_bit = [](unsigned m) { int k = -1; while (m) { k++; m >>= 1; } return k; };
        auto bit_count = [](unsigned m) { int k = 0; while (m) { k += m & 1; m >>= 1; } return k; };
        auto widen = [](unsigned masked, int shift, int bits) -> int {  // to 8 bits, the reference decoder's way
            long long v = (int)masked;  // signed: a mask that reaches bit 31 shifts arithmetically
            v = shift < 0 ? (long long)((unsigned long long)v << -shift) : v >> shift;
            v = (int)v;
            long long r = v;
            for (int k = bits; k < 8 && bits > 0; k += bits) r += v >> k;
            return (int)r;
        };
        int rs = 0, gs = 0, bs = 0, as = 0, rc = 0, gc = 0, bc = 0, ac = 0;
        if (!easy) {
            if (!mr || !mg || !mb) return false;
            rs = high_bit(mr) - 7; rc = bit_count(mr);
            gs = high_bit(mg) - 7; gc = bit_count(mg);
            bs = high_bit(mb) - 7; bc = bit_count(mb);
            as = high_bit(ma) - 7; ac = bit_count(ma);
        }
        for (int j = 0; j < h; j++) {
            for (int i = 0; i < w; i++, z += 4) {
                if (easy) {
                    img[z + 2] = (uint8_t)g8(); img[z + 1] = (uint8_t)g8(); img[z] = (uint8_t)g8();
                    const unsigned a = easy == 2 ? g8() : 255u;
                    all_a |= a;
                    img[z + 3] = (uint8_t)a;
                } else {
                    const unsigned v = bpp == 16 ? g16() : g32();
                    img[z] = (uint8_t)widen(v & mr, rs, rc);
                    img[z + 1] = (uint8_t)widen(v & mg, gs, gc);
                    img[z + 2] = (uint8_t)widen(v & mb, bs, bc);
                    const int a = ma ? widen(v & ma, as, ac) : 255;
                    all_a |= (unsigned)a;
                    img[z + 3] = (uint8_t)a;
                }
            }
            p += (size_t)pad;
        }
        if (!ma) for (size_t i = 3; i < img.size(); i += 4) img[i] = 255;  // three components: opaque
    }
    if (all_a == 0) for (size_t i = 3; i < img.size(); i += 4) img[i] = 255;
    rgba.assign((size_t)4 * w * h, 255);
    for (int y = 0; y < h; y++)  // stored row y is image row (bottom_up ? h-1-y : y) counted from the top
        put_flipped(rgba, w, h, bottom_up ? h - 1 - y : y, &img[(size_t)4 * w * y]);
    return true;
}

// ---- JPEG: baseline / extended sequential (SOF0 / SOF1) and progressive (SOF2) Huffman, 8-bit, grey or YCbCr (RGB when
// an Adobe marker says so), any sampling factors, interleaved or one scan per component, restart intervals.
// Arithmetic-coded, lossless and hierarchical files are refused (the caller prints what is read).  A JPEG's texels are not defined bit for bit by the format: what the reference gets is what
// ITS decoder makes of the file, so the three places where decoders differ follow stb_image's choices -- a 12-bit
// fixed-point "islow" inverse DCT (two extra bits kept between the passes), triangle-filtered chroma upsampling for
// 2x1 / 1x2 / 2x2 sampling (replication otherwise), and a 20-bit fixed-point YCbCr -> RGB.
struct JpegHuff {
    uint8_t size[257];
    uint16_t code[256];
    uint8_t value[256];
    int maxcode[18];  // left-aligned to 16 bits, exclusive
    int delta[17];
    int n = 0;
    bool build(const unsigned char* counts, const unsigned char* vals) {
        int k = 0;
        for (int len = 1; len <= 16; len++)
            for (int i = 0; i < counts[len - 1]; i++) {
                if (k >= 256) return false;
                size[k++] = (uint8_t)len;
            }
        size[k] = 0;
        n = k;
        memcpy(value, vals, (size_t)k);
        int c = 0;
        k = 0;
        for (int len = 1; len <= 16; len++) {
            delta[len] = k - c;
            if (size[k] == len) {
                while (size[k] == len) code[k++] = (uint16_t)c++;
                if (c - 1 >= (1 << len)) return false;
            }
            maxcode[len] = c << (16 - len);
            c <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        return true;
    }
};

struct JpegBits {
    const unsigned char* d;
    size_t p, end;
    uint32_t acc = 0;
    int n = 0;
    bool hit_marker = false;
    void fill() {
        while (n <= 24) {
            int b = 0;
            if (!hit_marker && p < end) {
                b = d[p];
                if (b == 0xff) {
                    const int m = p + 1 < end ? d[p + 1] : 0xd9;
                    if (m == 0) p += 2;           // stuffed zero
                    else { hit_marker = true; b = 0; }  // a marker ends the entropy segment: feed zeros
                } else {
                    p++;
                }
            }
            acc |= (uint32_t)b << (24 - n);
            n += 8;
        }
    }
    int get(int k) {  // k bits, MSB first
        if (k == 0) return 0;
        if (n < k) fill();
        const int v = (int)(acc >> (32 - k));
        acc <<= k;
        n -= k;
        return v;
    }
    int decode(const JpegHuff& h) {
        if (n < 16) fill();
        const int top = (int)(acc >> 16);
        int len = 1;
        while (len <= 16 && top >= h.maxcode[len]) len++;
        if (len > 16) return -1;
        const int c = (top >> (16 - len)) + h.delta[len];
        if (c < 0 || c >= h.n) return -1;
        acc <<= len;
        n -= len;
        return h.value[c];
    }
    static int extend(int v, int bits) { return bits && v < (1 << (bits - 1)) ? v - (1 << bits) + 1 : v; }
    void reset() { acc = 0; n = 0; hit_marker = false; }
};

// one dimension of the inverse DCT on eight values `s`, 12-bit constants; even part in x[0..3], odd part in t[0..3]
// (64-bit intermediates: a corrupt file's coefficients may be anything; valid data never leaves 32 bits)
typedef long long idct_t;
inline void jpeg_idct_1d(const idct_t* s, int stride, idct_t x[4], idct_t t[4]) {
    auto F = [](double v) { return (idct_t)(int)(v * 4096 + 0.5); };
    idct_t p2 = s[2 * stride], p3 = s[6 * stride];
    idct_t p1 = (p2 + p3) * F(0.5411961);
    const idct_t e2 = p1 + p3 * F(-1.847759065), e3 = p1 + p2 * F(0.765366865);
    p2 = s[0];
    p3 = s[4 * stride];
    const idct_t e0 = (p2 + p3) * 4096, e1 = (p2 - p3) * 4096;
    x[0] = e0 + e3;
    x[3] = e0 - e3;
    x[1] = e1 + e2;
    x[2] = e1 - e2;
    idct_t t0 = s[7 * stride], t1 = s[5 * stride], t2 = s[3 * stride], t3 = s[1 * stride];
    p3 = t0 + t2;
    idct_t p4 = t1 + t3;
    p1 = t0 + t3;
    p2 = t1 + t2;
    const idct_t p5 = (p3 + p4) * F(1.175875602);
    t0 *= F(0.298631336);
    t1 *= F(2.053119869);
    t2 *= F(3.072711026);
    t3 *= F(1.501321110);
    p1 = p5 + p1 * F(-0.899976223);
    p2 = p5 + p2 * F(-2.562915447);
    p3 *= F(-1.961570560);
    p4 *= F(-0.390180644);
    t[3] = t3 + p1 + p4;
    t[2] = t2 + p2 + p3;
    t[1] = t1 + p2 + p4;
    t[0] = t0 + p1 + p3;
}

inline uint8_t clamp8(long long v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

void jpeg_idct_block(const short* coef, uint8_t* out, int out_stride) {
    idct_t mid[64];
    for (int c = 0; c < 8; c++) {  // columns
        const short* d = coef + c;
        idct_t col[8];
        bool ac = false;
        for (int r = 0; r < 8; r++) {
            col[r] = d[8 * r];
            ac = ac || (r && col[r]);
        }
        if (!ac) {
            for (int r = 0; r < 8; r++) mid[8 * r + c] = col[0] * 4;
            continue;
        }
        idct_t x[4], t[4];
        jpeg_idct_1d(col, 1, x, t);
        for (int k = 0; k < 4; k++) {
            const idct_t e = x[k] + 512;
            mid[8 * k + c] = (e + t[3 - k]) >> 10;
            mid[8 * (7 - k) + c] = (e - t[3 - k]) >> 10;
        }
    }
    for (int r = 0; r < 8; r++) {  // rows; +128 and the rounding of the final >> 17 folded into the even part
        idct_t x[4], t[4];
        jpeg_idct_1d(mid + 8 * r, 1, x, t);
        uint8_t* o = out + (size_t)r * out_stride;
        for (int k = 0; k < 4; k++) {
            const idct_t e = x[k] + 65536 + (128 << 17);
            o[k] = clamp8((e + t[3 - k]) >> 17);
            o[7 - k] = clamp8((e - t[3 - k]) >> 17);
        }
    }
}

// One JPEG file.  Every scan decodes into per-component coefficient planes (64 shorts per 8x8 block); dequantisation
// and the inverse DCT run once, at the end -- which is what lets progressive files (SOF2: spectral selection and
// successive approximation over many scans) and baseline files with one scan per component share the code.
struct JpegDecoder {
    struct Comp {
        int id = 0, hs = 1, vs = 1, tq = 0, td = 0, ta = 0, dc = 0;
        int bw = 0, bh = 0;  // blocks across / down, padded to whole MCUs
        int cw = 0, ch = 0;  // true size in samples
        std::vector<short> coef;
        std::vector<uint8_t> px;
    } comp[3];
    uint16_t qt[4][64] = {};
    bool have_qt[4] = {};
    JpegHuff hdc[4], hac[4];
    bool have_dc[4] = {}, have_ac[4] = {};
    int ncomp = 0, restart = 0, adobe_transform = -1;
    int hmax = 1, vmax = 1, mcux = 0, mcuy = 0;
    int w = 0, h = 0;
    bool have_frame = false, progressive = false;
    // the current scan
    int scan_n = 0, order[3] = {0, 0, 0}, ss = 0, se = 63, ah = 0, al = 0, eob_run = 0;

    static const unsigned char* zigzag() {
        static const unsigned char z[64 + 15] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                                 41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                                 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
                                                 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};  // a corrupt run stays inside
        return z;
    }

    bool block_sequential(JpegBits& br, Comp& c, short* blk) {  // a whole block of a baseline scan
        const int t = br.decode(hdc[c.td]);
        if (t < 0 || t > 15) return false;
        c.dc += JpegBits::extend(br.get(t), t);
        blk[0] = (short)c.dc;
        const unsigned char* z = zigzag();
        for (int k = 1; k < 64;) {
            const int rs = br.decode(hac[c.ta]);
            if (rs < 0) return false;
            const int run = rs >> 4, sz = rs & 15;
            if (sz == 0) {
                if (run != 15) break;  // end of block
                k += 16;
                continue;
            }
            k += run;
            if (k > 63) return false;
            blk[z[k++]] = (short)JpegBits::extend(br.get(sz), sz);
        }
        return true;
    }

    bool block_dc_progressive(JpegBits& br, Comp& c, short* blk) {
        if (ah == 0) {  // first pass: the difference, scaled by the point transform
            const int t = br.decode(hdc[c.td]);
            if (t < 0 || t > 15) return false;
            c.dc += JpegBits::extend(br.get(t), t);
            blk[0] = (short)(c.dc * (1 << al));
        } else if (br.get(1)) {  // refinement: one more bit
            blk[0] = (short)(blk[0] + (1 << al));
        }
        return true;
    }

    bool block_ac_progressive(JpegBits& br, Comp& c, short* blk) {
        const unsigned char* z = zigzag();
        if (ah == 0) {  // first pass of a band of coefficients
            if (eob_run) {
                --eob_run;
                return true;
            }
            int k = ss;
            do {
                const int rs = br.decode(hac[c.ta]);
                if (rs < 0) return false;
                const int sz = rs & 15, r = rs >> 4;
                if (sz == 0) {
                    if (r < 15) {
                        eob_run = 1 << r;
                        if (r) eob_run += br.get(r);
                        --eob_run;
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    if (k > 63) return false;
                    blk[z[k++]] = (short)(JpegBits::extend(br.get(sz), sz) * (1 << al));
                }
            } while (k <= se);
            return true;
        }
        const short bit = (short)(1 << al);  // refinement: one more bit of the non-zero coefficients, new +-1s between them
        auto refine = [&](short* q) {
            if (br.get(1) && (*q & bit) == 0) *q = (short)(*q > 0 ? *q + bit : *q - bit);
        };
        if (eob_run) {
            --eob_run;
            for (int k = ss; k <= se; k++) {
                short* q = &blk[z[k]];
                if (*q != 0) refine(q);
            }
            return true;
        }
        int k = ss;
        do {
            const int rs = br.decode(hac[c.ta]);
            if (rs < 0) return false;
            int sz = rs & 15, r = rs >> 4;
            if (sz == 0) {
                if (r < 15) {
                    eob_run = (1 << r) - 1;
                    if (r) eob_run += br.get(r);
                    r = 64;  // nothing new in the rest of this block
                }
            } else {
                if (sz != 1) return false;
                sz = br.get(1) ? bit : -bit;
            }
            while (k <= se) {
                short* q = &blk[z[k++]];
                if (*q != 0) {
                    refine(q);
                } else {
                    if (r == 0) {
                        *q = (short)sz;
                        break;
                    }
                    --r;
                }
            }
        } while (k <= se);
        return true;
    }

    bool block(JpegBits& br, Comp& c, int bx, int by) {
        short* blk = &c.coef[((size_t)by * c.bw + bx) * 64];
        if (!progressive) return block_sequential(br, c, blk);
        return ss == 0 ? block_dc_progressive(br, c, blk) : block_ac_progressive(br, c, blk);
    }

    // the entropy-coded segment after an SOS header that ended at `start`; returns the offset of the next marker
    bool scan(const unsigned char* d, size_t n, size_t start, size_t& next) {
        JpegBits br{d, start, n};
        for (int i = 0; i < scan_n; i++) comp[order[i]].dc = 0;
        eob_run = 0;
        int todo = restart;
        auto at_restart = [&](bool last) {
            if (!restart || --todo != 0 || last) return true;
            size_t q = br.p;  // the next marker must be RSTn: byte-align, skip it, reset the predictors
            while (q + 1 < n && !(d[q] == 0xff && d[q + 1] >= 0xd0 && d[q + 1] <= 0xd7)) {
                if (d[q] == 0xff && d[q + 1] != 0 && d[q + 1] != 0xff) return false;  // another marker: the scan is short
                q++;
            }
            if (q + 1 >= n) return false;
            br.p = q + 2;
            br.reset();
            for (int i = 0; i < scan_n; i++) comp[order[i]].dc = 0;
            eob_run = 0;
            todo = restart;
            return true;
        };
        if (scan_n == 1) {  // one component: its blocks in raster order, only those that hold image samples
            Comp& c = comp[order[0]];
            const int nx = (c.cw + 7) >> 3, ny = (c.ch + 7) >> 3;
            for (int by = 0; by < ny; by++)
                for (int bx = 0; bx < nx; bx++) {
                    if (!block(br, c, bx, by)) return false;
                    if (!at_restart(by == ny - 1 && bx == nx - 1)) return false;
                }
        } else {  // interleaved: MCU by MCU, hs x vs blocks of every component of the scan
            if (progressive && ss != 0) return false;
            for (int my = 0; my < mcuy; my++)
                for (int mx = 0; mx < mcux; mx++) {
                    for (int i = 0; i < scan_n; i++) {
                        Comp& c = comp[order[i]];
                        for (int y = 0; y < c.vs; y++)
                            for (int x = 0; x < c.hs; x++)
                                if (!block(br, c, mx * c.hs + x, my * c.vs + y)) return false;
                    }
                    if (!at_restart(my == mcuy - 1 && mx == mcux - 1)) return false;
                }
        }
        size_t q = br.p;  // the marker that ended the segment (the bit reader never reads past one)
        while (q + 1 < n && !(d[q] == 0xff && d[q + 1] != 0 && d[q + 1] != 0xff && !(d[q + 1] >= 0xd0 && d[q + 1] <= 0xd7))) q++;
        next = q;
        return true;
    }

    bool parse(const unsigned char* d, size_t n) {
        if (n < 4 || d[0] != 0xff || d[1] != 0xd8) return false;
        const unsigned char* z = zigzag();
        size_t p = 2;
        bool any_scan = false;
        while (p + 2 <= n) {
            if (d[p] != 0xff) return false;
            while (p < n && d[p] == 0xff) p++;  // fill bytes
            if (p >= n) break;
            const int m = d[p++];
            if (m == 0xd9) break;
            if (m == 0x01 || (m >= 0xd0 && m <= 0xd7)) continue;
            if (p + 2 > n) return false;
            const size_t len = ((size_t)d[p] << 8) | d[p + 1];
            if (len < 2 || p + len > n) return false;
            const unsigned char* b = d + p + 2;
            const size_t bl = len - 2;
            if (m == 0xdb) {  // DQT
                size_t o = 0;
                while (o < bl) {
                    const int pq = b[o] >> 4, t = b[o] & 15;
                    o++;
                    if (t > 3 || o + (pq ? 128 : 64) > bl) return false;
                    for (int i = 0; i < 64; i++, o += pq ? 2 : 1) qt[t][z[i]] = pq ? (uint16_t)((b[o] << 8) | b[o + 1]) : b[o];
                    have_qt[t] = true;
                }
            } else if (m == 0xc4) {  // DHT
                size_t o = 0;
                while (o + 17 <= bl) {
                    const int tc = b[o] >> 4, t = b[o] & 15;
                    if (tc > 1 || t > 3) return false;
                    int total = 0;
                    for (int i = 0; i < 16; i++) total += b[o + 1 + i];
                    if (total > 256 || o + 17 + (size_t)total > bl) return false;
                    JpegHuff& hf = tc ? hac[t] : hdc[t];
                    if (!hf.build(b + o + 1, b + o + 17)) return false;
                    (tc ? have_ac : have_dc)[t] = true;
                    o += 17 + (size_t)total;
                }
            } else if (m == 0xc0 || m == 0xc1 || m == 0xc2) {  // SOF0 / SOF1 / SOF2
                if (have_frame || bl < 6 || b[0] != 8) return false;
                progressive = m == 0xc2;
                h = (b[1] << 8) | b[2];
                w = (b[3] << 8) | b[4];
                ncomp = b[5];
                if ((ncomp != 1 && ncomp != 3) || w <= 0 || h <= 0 || bl < 6 + 3 * (size_t)ncomp) return false;
                for (int i = 0; i < ncomp; i++) {
                    comp[i].id = b[6 + 3 * i];
                    comp[i].hs = b[7 + 3 * i] >> 4;
                    comp[i].vs = b[7 + 3 * i] & 15;
                    comp[i].tq = b[8 + 3 * i];
                    if (comp[i].hs < 1 || comp[i].hs > 4 || comp[i].vs < 1 || comp[i].vs > 4 || comp[i].tq > 3) return false;
                }
                if (ncomp == 1) comp[0].hs = comp[0].vs = 1;  // a lone component is never interleaved: plain 8x8 blocks
                for (int i = 0; i < ncomp; i++) {
                    hmax = comp[i].hs > hmax ? comp[i].hs : hmax;
                    vmax = comp[i].vs > vmax ? comp[i].vs : vmax;
                }
                mcux = (w + 8 * hmax - 1) / (8 * hmax);
                mcuy = (h + 8 * vmax - 1) / (8 * vmax);
                for (int i = 0; i < ncomp; i++) {
                    Comp& c = comp[i];
                    if (hmax % c.hs || vmax % c.vs) return false;  // whole-number upsampling factors (every real file)
                    c.bw = mcux * c.hs;
                    c.bh = mcuy * c.vs;
                    c.cw = (w * c.hs + hmax - 1) / hmax;
                    c.ch = (h * c.vs + vmax - 1) / vmax;
                    c.coef.assign((size_t)64 * c.bw * c.bh, 0);
                }
                have_frame = true;
            } else if (m >= 0xc3 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) {
                return false;  // lossless, hierarchical, arithmetic coding
            } else if (m == 0xdd) {
                if (bl < 2) return false;
                restart = (b[0] << 8) | b[1];
            } else if (m == 0xee && bl >= 12 && !memcmp(b, "Adobe", 5)) {
                adobe_transform = b[11];
            } else if (m == 0xda) {  // SOS
                if (!have_frame || bl < 1) return false;
                scan_n = b[0];
                if (scan_n < 1 || scan_n > ncomp || bl < 1 + 2 * (size_t)scan_n + 3) return false;
                for (int i = 0; i < scan_n; i++) {
                    int k = 0;
                    while (k < ncomp && comp[k].id != b[1 + 2 * i]) k++;
                    if (k == ncomp) return false;
                    order[i] = k;
                    comp[k].td = b[2 + 2 * i] >> 4;
                    comp[k].ta = b[2 + 2 * i] & 15;
                    if (comp[k].td > 3 || comp[k].ta > 3) return false;
                }
                ss = b[1 + 2 * scan_n];
                se = b[2 + 2 * scan_n];
                ah = b[3 + 2 * scan_n] >> 4;
                al = b[3 + 2 * scan_n] & 15;
                if (progressive) {
                    if (ss > 63 || se > 63 || ss > se || ah > 13 || al > 13) return false;
                } else {
                    if (ss != 0 || ah != 0 || al != 0) return false;
                    se = 63;
                }
                for (int i = 0; i < scan_n; i++) {
                    const Comp& c = comp[order[i]];
                    const bool need_dc = !progressive || (ss == 0 && ah == 0), need_ac = !progressive || ss != 0;
                    if ((need_dc && !have_dc[c.td]) || (need_ac && !have_ac[c.ta])) return false;
                }
                size_t next = 0;
                if (!scan(d, n, p + len, next)) return false;
                any_scan = true;
                p = next;
                continue;
            }
            p += len;
        }
        return have_frame && any_scan;
    }

    // coefficient planes -> sample planes (dequantise in 16 bits, as the coefficients are held)
    bool finish() {
        for (int i = 0; i < ncomp; i++) {
            Comp& c = comp[i];
            if (!have_qt[c.tq]) return false;
            c.px.assign((size_t)64 * c.bw * c.bh, 0);
            const size_t stride = (size_t)8 * c.bw;
            const int nx = progressive || ncomp == 1 ? (c.cw + 7) >> 3 : c.bw, ny = progressive || ncomp == 1 ? (c.ch + 7) >> 3 : c.bh;
            for (int by = 0; by < ny; by++)
                for (int bx = 0; bx < nx; bx++) {
                    short* blk = &c.coef[((size_t)by * c.bw + bx) * 64];
                    for (int k = 0; k < 64; k++) blk[k] = (short)(blk[k] * qt[c.tq][k]);
                    jpeg_idct_block(blk, &c.px[((size_t)by * 8) * stride + (size_t)bx * 8], (int)stride);
                }
        }
        return true;
    }
};

bool load_jpeg(const std::string& s, std::vector<uint8_t>& rgba, int& w, int& h) {
    JpegDecoder J;
    if (!J.parse((const unsigned char*)s.data(), s.size()) || !J.finish()) return false;
    w = J.w;
    h = J.h;
    const int ncomp = J.ncomp, hmax = J.hmax, vmax = J.vmax;
    // ---- samples -> RGBA rows (top-down), written bottom row first
    rgba.assign((size_t)4 * w * h, 255);
    std::vector<uint8_t> line[3], row((size_t)4 * w);
    for (int i = 0; i < ncomp; i++) line[i].assign((size_t)(J.comp[i].bw * 8) * hmax + 16, 0);
    const bool to_rgb = ncomp == 3 && J.adobe_transform != 0;
    for (int y = 0; y < h; y++) {
        const uint8_t* src[3] = {nullptr, nullptr, nullptr};
        for (int i = 0; i < ncomp; i++) {
            const JpegDecoder::Comp& c = J.comp[i];
            const size_t stride = (size_t)8 * c.bw;
            const int fh = hmax / c.hs, fv = vmax / c.vs;  // upsampling factors
            const int wl = (w + fh - 1) / fh;               // low-resolution samples across
            int near = y / fv, far = near;
            if (fv == 2) far = (y & 1) ? near + 1 : near - 1;
            near = near < 0 ? 0 : (near > c.ch - 1 ? c.ch - 1 : near);
            far = far < 0 ? 0 : (far > c.ch - 1 ? c.ch - 1 : far);
            const uint8_t* a = &c.px[(size_t)near * stride];
            const uint8_t* f = &c.px[(size_t)far * stride];
            uint8_t* o = line[i].data();
            if (fh == 1 && fv == 1) {
                src[i] = a;
                continue;
            } else if (fh == 1 && fv == 2) {
                for (int x = 0; x < wl; x++) o[x] = (uint8_t)((3 * a[x] + f[x] + 2) >> 2);
            } else if (fh == 2 && fv == 1) {
                if (wl == 1) {
                    o[0] = o[1] = a[0];
                } else {
                    o[0] = a[0];
                    o[1] = (uint8_t)((a[0] * 3 + a[1] + 2) >> 2);
                    int x = 1;
                    for (; x < wl - 1; x++) {
                        const int t = 3 * a[x] + 2;
                        o[2 * x] = (uint8_t)((t + a[x - 1]) >> 2);
                        o[2 * x + 1] = (uint8_t)((t + a[x + 1]) >> 2);
                    }
                    o[2 * x] = (uint8_t)((a[wl - 2] * 3 + a[wl - 1] + 2) >> 2);  // the reference decoder's last column, as it is
                    o[2 * x + 1] = a[wl - 1];
                }
            } else if (fh == 2 && fv == 2) {
                int t1 = 3 * a[0] + f[0];
                if (wl == 1) {
                    o[0] = o[1] = (uint8_t)((t1 + 2) >> 2);
                } else {
                    o[0] = (uint8_t)((t1 + 2) >> 2);
                    for (int x = 1; x < wl; x++) {
                        const int t0 = t1;
                        t1 = 3 * a[x] + f[x];
                        o[2 * x - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
                        o[2 * x] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
                    }
                    o[2 * wl - 1] = (uint8_t)((t1 + 2) >> 2);
                }
            } else {  // any other ratio: nearest sample
                for (int x = 0; x < wl; x++)
                    for (int k = 0; k < fh; k++) o[x * fh + k] = a[x];
            }
            src[i] = o;
        }
        for (int x = 0; x < w; x++) {
            uint8_t* o = &row[4 * (size_t)x];
            if (ncomp == 1) {
                o[0] = o[1] = o[2] = src[0][x];
            } else if (!to_rgb) {
                o[0] = src[0][x], o[1] = src[1][x], o[2] = src[2][x];
            } else {
                auto FX = [](double v) { return ((int)(v * 4096.0 + 0.5)) << 8; };
                const int yf = (src[0][x] << 20) + (1 << 19), cb = src[1][x] - 128, cr = src[2][x] - 128;
                const int r = yf + cr * FX(1.40200);
                const int g = yf + cr * -FX(0.71414) + (int)((unsigned)(cb * -FX(0.34414)) & 0xffff0000u);
                const int bl2 = yf + cb * FX(1.77200);
                o[0] = clamp8(r >> 20), o[1] = clamp8(g >> 20), o[2] = clamp8(bl2 >> 20);
            }
            o[3] = 255;
        }
        put_flipped(rgba, w, h, y, row.data());
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
    if (s[0] == 'B' && s[1] == 'M') return load_bmp(s, rgba, w, h);
    if ((unsigned char)s[0] == 0xff && (unsigned char)s[1] == 0xd8) return load_jpeg(s, rgba, w, h);
    std::string ext = path.size() >= 4 ? path.substr(path.size() - 4) : "";
    for (auto& c : ext) c = (char)tolower((unsigned char)c);
    if (ext == ".tga") return load_tga(s, rgba, w, h);
    return false;
}

}  // namespace orbhost
