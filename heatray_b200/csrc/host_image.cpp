// host_image.cpp — image decoding for textures and image encoding for screenshots (host side, C++).
//
// Reference: Source/Utility/TextureLoader.cpp:17-159 decodes every non-.exr/.tiff texture with stb_image
// (stbi_load / stbi_loadf, vertical flip on) and HeatrayRenderer::saveScreenshot (HeatrayRenderer.cpp:1619-1662) writes
// through FreeImage. Neither library is used here: this file carries its own PNG (zlib inflate, all colour types and bit
// depths, Adam7, tRNS), JPEG (baseline + progressive Huffman, 8-bit) and Radiance RGBE decoders, and PNG / RGBE / PFM
// writers. tests/test_images.py pins the decoders bit for bit against stb_image as vendored by the reference. To be bit-identical
// the JPEG decoder has to follow stb_image's arithmetic, not just the standard: the inverse DCT below is the 12-bit fixed-point
// factorisation of 3rdParty/stb/stb_image.h:2270-2327 (stbi__idct_block: same constants, same rounding offsets, same pass order),
// and the chroma upsampling / YCbCr conversion use stb_image's fixed-point weights (stbi__resample_row_v_2 / h_2 / hv_2 and stbi__YCbCr_to_RGB_row, stb_image.h:3239-3310, 3434-3460).
#include "host_api.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <new>
#include <stdexcept>
#include <vector>

namespace {

thread_local std::string g_imgErr;
int imgFail(const char* what) { g_imgErr = what; return HRC_ERR_INVALID_VALUE; }

bool readFile(const char* path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return false; }
    out.resize((size_t)n);
    const size_t got = n ? std::fread(out.data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == (size_t)n;
}

// ---------------------------------------------------------------------------------------
// Checksums used by PNG / zlib
// ---------------------------------------------------------------------------------------
uint32_t crc32(const uint8_t* p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256]; static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; }
        init = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; ++i) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}
uint32_t adler32(const uint8_t* p, size_t n) {
    uint32_t a = 1, b = 0;
    while (n) {
        const size_t k = std::min<size_t>(n, 5552);
        for (size_t i = 0; i < k; ++i) { a += p[i]; b += a; }
        a %= 65521u; b %= 65521u; p += k; n -= k;
    }
    return (b << 16) | a;
}

// ---------------------------------------------------------------------------------------
// zlib inflate (RFC 1950/1951): stored, fixed and dynamic Huffman blocks
// ---------------------------------------------------------------------------------------
struct BitReader {
    const uint8_t* p; size_t n, pos = 0; uint32_t buf = 0; int cnt = 0; bool overrun = false;
    BitReader(const uint8_t* p_, size_t n_) : p(p_), n(n_) {}
    uint32_t bits(int k) {
        while (cnt < k) { uint32_t b = 0; if (pos < n) b = p[pos++]; else overrun = true; buf |= b << cnt; cnt += 8; }
        const uint32_t v = k ? (buf & ((1u << k) - 1u)) : 0u;
        buf >>= k; cnt -= k;
        return v;
    }
    void alignByte() { buf = 0; cnt = 0; }
};

struct Huffman {
    uint16_t count[16]; uint16_t symbol[320];
    bool build(const uint8_t* lengths, int n) {
        std::memset(count, 0, sizeof(count));
        for (int i = 0; i < n; ++i) count[lengths[i]]++;
        count[0] = 0;
        int left = 1;
        for (int len = 1; len < 16; ++len) { left <<= 1; left -= count[len]; if (left < 0) return false; }
        uint16_t offs[16]; offs[1] = 0;
        for (int len = 1; len < 15; ++len) offs[len + 1] = offs[len] + count[len];
        for (int i = 0; i < n; ++i) if (lengths[i]) symbol[offs[lengths[i]]++] = (uint16_t)i;
        return true;
    }
    int decode(BitReader& br) const {
        int code = 0, first = 0, index = 0;
        for (int len = 1; len < 16; ++len) {
            code |= (int)br.bits(1);
            const int c = count[len];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

bool inflateRaw(BitReader& br, std::vector<uint8_t>& out) {
    static const uint16_t lbase[29] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
    static const uint16_t lext[29] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
    static const uint16_t dbase[30] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
    static const uint16_t dext[30] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
    int last;
    do {
        last = (int)br.bits(1);
        const int type = (int)br.bits(2);
        if (type == 0) {
            br.alignByte();
            if (br.pos + 4 > br.n) return false;
            const uint32_t len = br.p[br.pos] | (br.p[br.pos + 1] << 8), nlen = br.p[br.pos + 2] | (br.p[br.pos + 3] << 8);
            br.pos += 4;
            if ((len ^ 0xffffu) != nlen || br.pos + len > br.n) return false;
            out.insert(out.end(), br.p + br.pos, br.p + br.pos + len);
            br.pos += len;
        } else if (type == 1 || type == 2) {
            Huffman lit, dist;
            uint8_t lengths[320];
            if (type == 1) {
                int i = 0;
                for (; i < 144; ++i) lengths[i] = 8;
                for (; i < 256; ++i) lengths[i] = 9;
                for (; i < 280; ++i) lengths[i] = 7;
                for (; i < 288; ++i) lengths[i] = 8;
                lit.build(lengths, 288);
                for (i = 0; i < 30; ++i) lengths[i] = 5;
                dist.build(lengths, 30);
            } else {
                static const uint8_t order[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};
                const int nlen = (int)br.bits(5) + 257, ndist = (int)br.bits(5) + 1, ncode = (int)br.bits(4) + 4;
                if (nlen > 286 || ndist > 30) return false;
                uint8_t cl[19] = {0};
                for (int i = 0; i < ncode; ++i) cl[order[i]] = (uint8_t)br.bits(3);
                Huffman clh;
                if (!clh.build(cl, 19)) return false;
                int idx = 0;
                while (idx < nlen + ndist) {
                    const int sym = clh.decode(br);
                    if (sym < 0) return false;
                    if (sym < 16) lengths[idx++] = (uint8_t)sym;
                    else {
                        int rep, val = 0;
                        if (sym == 16) { if (idx == 0) return false; val = lengths[idx - 1]; rep = 3 + (int)br.bits(2); }
                        else if (sym == 17) rep = 3 + (int)br.bits(3);
                        else rep = 11 + (int)br.bits(7);
                        if (idx + rep > nlen + ndist) return false;
                        while (rep--) lengths[idx++] = (uint8_t)val;
                    }
                }
                if (!lit.build(lengths, nlen)) return false;
                if (!dist.build(lengths + nlen, ndist)) { /* incomplete distance codes are legal when only one is used */ }
            }
            for (;;) {
                const int sym = lit.decode(br);
                if (sym < 0 || br.overrun) return false;
                if (sym < 256) out.push_back((uint8_t)sym);
                else if (sym == 256) break;
                else {
                    const int li = sym - 257;
                    if (li >= 29) return false;
                    const int len = lbase[li] + (int)br.bits(lext[li]);
                    const int ds = dist.decode(br);
                    if (ds < 0 || ds >= 30) return false;
                    const size_t d = dbase[ds] + br.bits(dext[ds]);
                    if (d > out.size()) return false;
                    const size_t start = out.size() - d;
                    for (int i = 0; i < len; ++i) out.push_back(out[start + i]);
                }
            }
        } else return false;
    } while (!last);
    return true;
}

bool zlibInflate(const uint8_t* p, size_t n, std::vector<uint8_t>& out) {
    if (n < 6) return false;
    if ((p[0] & 0x0f) != 8 || ((p[0] << 8) | p[1]) % 31 != 0 || (p[1] & 0x20)) return false;
    BitReader br(p + 2, n - 2);
    return inflateRaw(br, out);
}

} // namespace

// ---------------------------------------------------------------------------------------
// PNG decoder (ISO/IEC 15948): every colour type and bit depth, Adam7 interlace, tRNS.
// Output conventions follow what the reference's loader hands on (stbi_load, TextureLoader.cpp:101): 8 bits per channel,
// 16-bit samples keep their high byte, sub-byte greys are scaled to 0..255, a tRNS chunk adds an alpha channel.
// ---------------------------------------------------------------------------------------
namespace {

struct Decoded { int w = 0, h = 0, comp = 0; bool isFloat = false; std::vector<uint8_t> u8; std::vector<float> f32; };

// Header sanity shared by the decoders (stb_image: STBI_MAX_DIMENSIONS = 1 << 24 per side): a corrupt header must not turn into a giant allocation.
inline bool dimsOk(int64_t w, int64_t h) { return w > 0 && h > 0 && w <= (1 << 24) && h <= (1 << 24) && w * h <= (int64_t(1) << 28); }
inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// Undo the scanline filters of one (sub)image: `raw` holds h rows of (1 + stride) bytes; result is h*stride bytes.
bool pngUnfilter(const uint8_t* raw, size_t rawSize, int h, size_t stride, int bpp, std::vector<uint8_t>& out) {
    if (rawSize < (size_t)h * (stride + 1)) return false;
    out.assign((size_t)h * stride, 0);
    std::vector<uint8_t> zero(stride, 0);
    for (int y = 0; y < h; ++y) {
        const uint8_t* src = raw + (size_t)y * (stride + 1);
        const int ft = src[0]; ++src;
        uint8_t* cur = out.data() + (size_t)y * stride;
        const uint8_t* up = y ? cur - stride : zero.data();
        if (ft > 4) return false;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = up[i], c = i >= (size_t)bpp ? up[i - bpp] : 0;
            int v = src[i];
            switch (ft) {
            case 1: v += a; break;
            case 2: v += b; break;
            case 3: v += (a + b) >> 1; break;
            case 4: v += paeth(a, b, c); break;
            default: break;
            }
            cur[i] = (uint8_t)v;
        }
    }
    return true;
}

bool decodePNG(const uint8_t* p, size_t n, Decoded& img) {
    static const uint8_t sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
    if (n < 8 || std::memcmp(p, sig, 8) != 0) return false;
    size_t pos = 8;
    int w = 0, h = 0, depth = 0, color = 0, interlace = 0;
    std::vector<uint8_t> idat, palette; uint8_t palAlpha[256]; int palAlphaCount = 0;
    bool hasTrns = false; uint16_t trns[3] = {0, 0, 0};
    bool gotHdr = false, done = false;
    while (!done && pos + 12 <= n) {
        const uint32_t len = be32(p + pos); const uint8_t* type = p + pos + 4; const uint8_t* data = p + pos + 8;
        if (pos + 12 + (size_t)len > n) return false;
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13) return false;
            w = (int)be32(data); h = (int)be32(data + 4); depth = data[8]; color = data[9]; interlace = data[12];
            if (!dimsOk(w, h) || data[10] != 0 || data[11] != 0 || interlace > 1) return false;
            if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
            if (color != 0 && color != 2 && color != 3 && color != 4 && color != 6) return false;
            if (color == 3 && depth == 16) return false;
            if ((color == 2 || color == 4 || color == 6) && depth < 8) return false;
            gotHdr = true;
        } else if (!std::memcmp(type, "PLTE", 4)) {
            if (len > 768 || len % 3) return false;
            palette.assign(data, data + len);
        } else if (!std::memcmp(type, "tRNS", 4)) {
            if (!gotHdr) return false;
            if (color == 3) { palAlphaCount = (int)std::min<uint32_t>(len, 256); std::memcpy(palAlpha, data, palAlphaCount); hasTrns = true; }
            else if (color == 0 && len >= 2) { trns[0] = (uint16_t)((data[0] << 8) | data[1]); hasTrns = true; }
            else if (color == 2 && len >= 6) { for (int k = 0; k < 3; ++k) trns[k] = (uint16_t)((data[2 * k] << 8) | data[2 * k + 1]); hasTrns = true; }
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), data, data + len);
        } else if (!std::memcmp(type, "IEND", 4)) done = true;
        pos += 12 + (size_t)len;
    }
    if (!gotHdr || idat.empty()) return false;
    if (color == 3 && palette.empty()) return false;
    std::vector<uint8_t> raw;
    if (!zlibInflate(idat.data(), idat.size(), raw)) return false;

    const int fileComp = color == 0 ? 1 : color == 2 ? 3 : color == 3 ? 1 : color == 4 ? 2 : 4;     // samples per pixel in the file
    const int bitsPerPixel = fileComp * depth;
    const int bpp = std::max(1, bitsPerPixel / 8);                                                   // filter distance in bytes
    // samples[y][x][c] as 16-bit values straight from the file (palette index for colour type 3)
    std::vector<uint16_t> samples((size_t)w * h * fileComp);
    auto unpack = [&](const std::vector<uint8_t>& rows, int sw, int sh, int x0, int y0, int dx, int dy) {
        const size_t stride = ((size_t)sw * bitsPerPixel + 7) / 8;
        for (int y = 0; y < sh; ++y) {
            const uint8_t* r = rows.data() + (size_t)y * stride;
            for (int x = 0; x < sw; ++x) for (int c = 0; c < fileComp; ++c) {
                const size_t s = (size_t)x * fileComp + c;
                uint16_t v;
                if (depth == 8) v = r[s];
                else if (depth == 16) v = (uint16_t)((r[2 * s] << 8) | r[2 * s + 1]);
                else { const size_t bit = s * depth; v = (uint16_t)((r[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1)); }
                samples[(((size_t)(y0 + y * dy)) * w + (x0 + x * dx)) * fileComp + c] = v;
            }
        }
    };
    if (!interlace) {
        std::vector<uint8_t> rows;
        if (!pngUnfilter(raw.data(), raw.size(), h, ((size_t)w * bitsPerPixel + 7) / 8, bpp, rows)) return false;
        unpack(rows, w, h, 0, 0, 1, 1);
    } else {
        static const int xo[7] = {0, 4, 0, 2, 0, 1, 0}, yo[7] = {0, 0, 4, 0, 2, 0, 1}, xs[7] = {8, 8, 4, 4, 2, 2, 1}, ys[7] = {8, 8, 8, 4, 4, 2, 2};
        size_t off = 0;
        for (int pass = 0; pass < 7; ++pass) {
            const int sw = (w - xo[pass] + xs[pass] - 1) / xs[pass], sh = (h - yo[pass] + ys[pass] - 1) / ys[pass];
            if (sw <= 0 || sh <= 0) continue;
            const size_t stride = ((size_t)sw * bitsPerPixel + 7) / 8;
            std::vector<uint8_t> rows;
            if (off > raw.size() || !pngUnfilter(raw.data() + off, raw.size() - off, sh, stride, bpp, rows)) return false;
            off += (size_t)sh * (stride + 1);
            unpack(rows, sw, sh, xo[pass], yo[pass], xs[pass], ys[pass]);
        }
    }

    // samples -> 8-bit output
    const int scale = depth == 1 ? 255 : depth == 2 ? 85 : depth == 4 ? 17 : 1;
    int outComp;
    if (color == 3) outComp = hasTrns ? 4 : 3;
    else outComp = fileComp + ((hasTrns && (color == 0 || color == 2)) ? 1 : 0);
    img.w = w; img.h = h; img.comp = outComp; img.isFloat = false;
    img.u8.resize((size_t)w * h * outComp);
    const size_t px = (size_t)w * h;
    for (size_t i = 0; i < px; ++i) {
        const uint16_t* s = &samples[i * fileComp];
        uint8_t* o = &img.u8[i * outComp];
        if (color == 3) {
            const size_t idx = s[0];
            if (idx * 3 + 2 < palette.size()) { o[0] = palette[idx * 3]; o[1] = palette[idx * 3 + 1]; o[2] = palette[idx * 3 + 2]; }
            else { o[0] = o[1] = o[2] = 0; }
            if (hasTrns) o[3] = idx < (size_t)palAlphaCount ? palAlpha[idx] : 255;
            continue;
        }
        for (int c = 0; c < fileComp; ++c) o[c] = depth == 16 ? (uint8_t)(s[c] >> 8) : (uint8_t)(s[c] * scale);
        if (hasTrns && color == 0) o[1] = (s[0] == (depth == 16 ? trns[0] : (uint16_t)(trns[0] & ((1 << depth) - 1)))) ? 0 : 255;
        if (hasTrns && color == 2) {
            bool same = true;
            for (int c = 0; c < 3; ++c) same = same && (s[c] == (depth == 16 ? trns[c] : (uint16_t)(trns[c] & 255)));
            o[3] = same ? 0 : 255;
        }
    }
    return true;
}

} // namespace

// ---------------------------------------------------------------------------------------
// TIFF (the FreeImage path of util::loadTexture, TextureLoader.cpp:44-87; the reference ships its multiscatter LUT as
// Resources/multiscatter_lut.tiff: 128 x 128, one 32-bit float sample, LZW strips). Baseline subset: II / MM, strips, chunky planar
// configuration, compression none / LZW / PackBits, predictor 1 / 2, 32-bit float with 1 or 3 samples (FIT_FLOAT / FIT_RGBF) and
// 8-bit RGBA (FIT_BITMAP). Rows are returned top-down like the other decoders; the texture loader flips them (FreeImage bitmaps are
// bottom-up). Quirk kept: FreeImage holds 8-bit pixels as B,G,R,A on little-endian hosts and the reference uploads those bytes as RL_RGBA.
// ---------------------------------------------------------------------------------------
namespace {

struct TiffReader {
    const uint8_t* p; size_t n; bool le;
    bool ok(size_t off, size_t len) const { return off <= n && len <= n - off; }
    uint32_t u16(size_t o) const { return le ? (uint32_t)(p[o] | (p[o + 1] << 8)) : (uint32_t)((p[o] << 8) | p[o + 1]); }
    uint32_t u32(size_t o) const { return le ? ((uint32_t)p[o] | ((uint32_t)p[o + 1] << 8) | ((uint32_t)p[o + 2] << 16) | ((uint32_t)p[o + 3] << 24)) : be32(p + o); }
    // value(s) of an IFD entry of type SHORT or LONG
    bool values(size_t entry, std::vector<uint32_t>& out) const {
        const uint32_t type = u16(entry + 2), count = u32(entry + 4);
        const size_t sz = type == 3 ? 2 : type == 4 ? 4 : type == 1 ? 1 : 0;
        if (!sz || count == 0 || count > (1u << 24)) return false;
        size_t at = entry + 8;
        if (sz * count > 4) { at = u32(entry + 8); if (!ok(at, sz * count)) return false; }
        out.resize(count);
        for (uint32_t i = 0; i < count; ++i) out[i] = sz == 2 ? u16(at + 2 * i) : sz == 4 ? u32(at + 4 * i) : p[at + i];
        return true;
    }
};

// TIFF flavour of LZW (TIFF 6.0 section 13): MSB-first codes of 9..12 bits, ClearCode 256, EoiCode 257, code width grows one code early.
bool tiffLZW(const uint8_t* src, size_t n, std::vector<uint8_t>& out, size_t expected) {
    std::vector<uint16_t> prefix(4096); std::vector<uint8_t> suffix(4096), first(4096); std::vector<uint16_t> length(4096);
    for (int i = 0; i < 256; ++i) { prefix[i] = 0xffff; suffix[i] = first[i] = (uint8_t)i; length[i] = 1; }
    int bits = 9, next = 258, old = -1; uint32_t acc = 0; int have = 0; size_t pos = 0;
    out.clear(); out.reserve(expected);
    auto emit = [&](int code) {
        const size_t len = length[code], at = out.size();
        out.resize(at + len);
        for (size_t k = len; k-- > 0;) { out[at + k] = suffix[code]; code = prefix[code]; }
    };
    for (;;) {
        while (have < bits) { if (pos >= n) return out.size() >= expected; acc = (acc << 8) | src[pos++]; have += 8; }
        const int code = (int)((acc >> (have - bits)) & ((1u << bits) - 1)); have -= bits;
        if (code == 257) break;
        if (code == 256) { bits = 9; next = 258; old = -1; continue; }
        if (old < 0) { if (code >= 256) return false; emit(code); old = code; continue; }
        if (code < next) {
            emit(code);
            if (next < 4096) { prefix[next] = (uint16_t)old; suffix[next] = first[code]; first[next] = first[old]; length[next] = (uint16_t)(length[old] + 1); ++next; }
        } else if (code == next && next < 4096) {
            prefix[next] = (uint16_t)old; suffix[next] = first[old]; first[next] = first[old]; length[next] = (uint16_t)(length[old] + 1); ++next;
            emit(code);
        } else return false;
        old = code;
        if (next + 1 >= (1 << bits) && bits < 12) ++bits;
        if (out.size() > expected + 4096) return false;
    }
    return out.size() >= expected;
}

bool tiffPackBits(const uint8_t* src, size_t n, std::vector<uint8_t>& out, size_t expected) {
    out.clear(); size_t pos = 0;
    while (pos < n && out.size() < expected) {
        const int8_t c = (int8_t)src[pos++];
        if (c >= 0) { const size_t len = (size_t)c + 1; if (pos + len > n) return false; out.insert(out.end(), src + pos, src + pos + len); pos += len; }
        else if (c != -128) { if (pos >= n) return false; out.insert(out.end(), (size_t)(1 - c), src[pos++]); }
    }
    return out.size() >= expected;
}

bool isTIFF(const uint8_t* p, size_t n) { return n >= 8 && ((p[0] == 'I' && p[1] == 'I' && p[2] == 42 && p[3] == 0) || (p[0] == 'M' && p[1] == 'M' && p[2] == 0 && p[3] == 42)); }

bool decodeTIFF(const uint8_t* p, size_t n, Decoded& img) {
    if (!isTIFF(p, n)) return false;
    TiffReader r{p, n, p[0] == 'I'};
    const size_t ifd = r.u32(4);
    if (!r.ok(ifd, 2)) return false;
    const uint32_t entries = r.u16(ifd);
    if (!r.ok(ifd + 2, (size_t)entries * 12)) return false;
    uint32_t w = 0, h = 0, compression = 1, spp = 1, rowsPerStrip = 0xffffffffu, planar = 1, predictor = 1, photometric = 1;
    std::vector<uint32_t> bitsPerSample{1}, sampleFormat{1}, offsets, counts, v;
    for (uint32_t i = 0; i < entries; ++i) {
        const size_t e = ifd + 2 + (size_t)i * 12;
        const uint32_t tag = r.u16(e);
        const bool known = tag == 256 || tag == 257 || tag == 258 || tag == 259 || tag == 262 || tag == 273 || tag == 277 || tag == 278 || tag == 279 || tag == 284 || tag == 317 || tag == 339;
        if (!known) continue;
        if (!r.values(e, v)) return false;
        switch (tag) {
        case 256: w = v[0]; break;            case 257: h = v[0]; break;
        case 258: bitsPerSample = v; break;   case 259: compression = v[0]; break;
        case 262: photometric = v[0]; break;  case 273: offsets = v; break;
        case 277: spp = v[0]; break;          case 278: rowsPerStrip = v[0]; break;
        case 279: counts = v; break;          case 284: planar = v[0]; break;
        case 317: predictor = v[0]; break;    case 339: sampleFormat = v; break;
        }
    }
    (void)photometric;
    if (!dimsOk(w, h) || planar != 1 || spp < 1 || spp > 4 || offsets.empty() || offsets.size() != counts.size() || rowsPerStrip == 0) return false;
    const uint32_t bps = bitsPerSample[0], fmt = sampleFormat[0];
    for (uint32_t b : bitsPerSample) if (b != bps) return false;
    const bool isFloat = bps == 32 && fmt == 3;
    if (isFloat) { if (spp != 1 && spp != 3) return false; }                // FIT_FLOAT, FIT_RGBF; the reference asserts on anything else (:60-75)
    else if (!(bps == 8 && fmt == 1 && spp == 4)) return false;             // FIT_BITMAP is uploaded as RL_RGBA (:65-68): only 32 bpp is consistent with that
    if (compression != 1 && compression != 5 && compression != 32773) return false;
    if (predictor != 1 && !(predictor == 2 && !isFloat)) return false;
    const size_t bytesPerSample = bps / 8, rowBytes = (size_t)w * spp * bytesPerSample;
    if (rowsPerStrip > h) rowsPerStrip = h;
    const size_t strips = ((size_t)h + rowsPerStrip - 1) / rowsPerStrip;
    if (offsets.size() < strips) return false;
    std::vector<uint8_t> raw((size_t)h * rowBytes), strip;
    for (size_t s = 0; s < strips; ++s) {
        const size_t rows = std::min<size_t>(rowsPerStrip, h - s * rowsPerStrip), want = rows * rowBytes;
        if (!r.ok(offsets[s], counts[s])) return false;
        const uint8_t* src = p + offsets[s];
        if (compression == 1) { if (counts[s] < want) return false; std::memcpy(raw.data() + s * rowsPerStrip * rowBytes, src, want); continue; }
        if (!(compression == 5 ? tiffLZW(src, counts[s], strip, want) : tiffPackBits(src, counts[s], strip, want))) return false;
        std::memcpy(raw.data() + s * rowsPerStrip * rowBytes, strip.data(), want);
    }
    if (predictor == 2) for (size_t y = 0; y < h; ++y) { uint8_t* row = raw.data() + y * rowBytes; for (size_t i = spp; i < rowBytes; ++i) row[i] = (uint8_t)(row[i] + row[i - spp]); }
    img.w = (int)w; img.h = (int)h; img.comp = (int)spp; img.isFloat = isFloat;
    if (isFloat) {
        img.f32.resize((size_t)w * h * spp);
        for (size_t i = 0; i < img.f32.size(); ++i) { const uint32_t bitsv = (r.le ? ((uint32_t)raw[4 * i] | ((uint32_t)raw[4 * i + 1] << 8) | ((uint32_t)raw[4 * i + 2] << 16) | ((uint32_t)raw[4 * i + 3] << 24)) : be32(raw.data() + 4 * i)); std::memcpy(&img.f32[i], &bitsv, 4); }
    } else {
        img.u8 = raw;
        for (size_t i = 0; i + 3 < img.u8.size(); i += 4) std::swap(img.u8[i], img.u8[i + 2]);   // FreeImage's B,G,R,A byte order, handed to RL_RGBA as is
    }
    return true;
}

} // namespace

// ---------------------------------------------------------------------------------------
// Radiance .hdr (RGBE): header "#?RADIANCE" / "#?RGBE", FORMAT=32-bit_rle_rgbe, "-Y h +X w"; flat or new-style RLE scanlines.
// Texel conversion as the reference's loader sees it (stbi_loadf, TextureLoader.cpp:95): mantissa * 2^(e - 136), no +0.5 bias,
// exponent 0 -> black. Three channels.
// ---------------------------------------------------------------------------------------
namespace {

bool hdrToken(const uint8_t* p, size_t n, size_t& pos, std::string& line) {
    line.clear();
    while (pos < n && p[pos] != '\n') { if (line.size() < 1023) line.push_back((char)p[pos]); ++pos; }
    if (pos < n) ++pos;
    return true;
}
inline void rgbeToFloat(const uint8_t* rgbe, float* out) {
    if (rgbe[3] != 0) {
        const float f = std::ldexp(1.0f, (int)rgbe[3] - (128 + 8));
        out[0] = rgbe[0] * f; out[1] = rgbe[1] * f; out[2] = rgbe[2] * f;
    } else out[0] = out[1] = out[2] = 0.0f;
}

bool isHDR(const uint8_t* p, size_t n) {
    return (n >= 11 && !std::memcmp(p, "#?RADIANCE\n", 11)) || (n >= 7 && !std::memcmp(p, "#?RGBE\n", 7));
}

bool decodeHDR(const uint8_t* p, size_t n, Decoded& img) {
    if (!isHDR(p, n)) return false;
    size_t pos = 0; std::string line; bool format = false;
    hdrToken(p, n, pos, line);
    for (;;) {
        if (pos >= n) return false;
        hdrToken(p, n, pos, line);
        if (line.empty()) break;
        if (line == "FORMAT=32-bit_rle_rgbe") format = true;
    }
    if (!format) return false;
    hdrToken(p, n, pos, line);
    int w = 0, h = 0;
    if (std::sscanf(line.c_str(), "-Y %d +X %d", &h, &w) != 2 || !dimsOk(w, h)) return false;
    img.w = w; img.h = h; img.comp = 3; img.isFloat = true;
    img.f32.assign((size_t)w * h * 3, 0.0f);
    std::vector<uint8_t> scan((size_t)w * 4);
    if (w < 8 || w >= 32768) {                                                           // flat data
        if (pos + (size_t)w * h * 4 > n) return false;
        for (size_t i = 0; i < (size_t)w * h; ++i) rgbeToFloat(p + pos + i * 4, &img.f32[i * 3]);
        return true;
    }
    for (int y = 0; y < h; ++y) {
        if (pos + 4 > n) return false;
        const uint8_t c1 = p[pos], c2 = p[pos + 1], len = p[pos + 2];
        if (c1 != 2 || c2 != 2 || (len & 0x80)) {                                       // not RLE: the rest of the file is flat
            if (y != 0) return false;
            if (pos + (size_t)w * h * 4 > n) return false;
            for (size_t i = 0; i < (size_t)w * h; ++i) rgbeToFloat(p + pos + i * 4, &img.f32[i * 3]);
            return true;
        }
        if ((((int)len << 8) | p[pos + 3]) != w) return false;
        pos += 4;
        for (int k = 0; k < 4; ++k) {
            int i = 0;
            while (i < w) {
                if (pos >= n) return false;
                int count = p[pos++];
                if (count > 128) {                                                      // run
                    count -= 128;
                    if (pos >= n || count > w - i) return false;
                    const uint8_t v = p[pos++];
                    for (int z = 0; z < count; ++z) scan[(size_t)(i++) * 4 + k] = v;
                } else {                                                                // dump
                    if (count == 0 || count > w - i || pos + count > n) return false;
                    for (int z = 0; z < count; ++z) scan[(size_t)(i++) * 4 + k] = p[pos++];
                }
            }
        }
        for (int x = 0; x < w; ++x) rgbeToFloat(&scan[(size_t)x * 4], &img.f32[((size_t)y * w + x) * 3]);
    }
    return true;
}

void flipRows(void* data, int h, size_t rowBytes) {
    std::vector<uint8_t> tmp(rowBytes);
    uint8_t* b = (uint8_t*)data;
    for (int y = 0; y < h / 2; ++y) {
        uint8_t* r0 = b + (size_t)y * rowBytes; uint8_t* r1 = b + (size_t)(h - 1 - y) * rowBytes;
        std::memcpy(tmp.data(), r0, rowBytes); std::memcpy(r0, r1, rowBytes); std::memcpy(r1, tmp.data(), rowBytes);
    }
}

} // namespace

// ---------------------------------------------------------------------------------------
// JPEG decoder (ITU-T T.81): 8-bit baseline / extended sequential / progressive Huffman, 1 or 3 components, restart
// intervals, any sampling factors. The reconstruction arithmetic is the one the reference's loader applies (stbi_load):
// the 12-bit fixed-point "islow" inverse DCT with 2 extra bits kept between the passes, triangle-filter chroma upsampling
// for 2x factors (nearest otherwise) and the 20-bit fixed-point YCbCr->RGB conversion — so decoded texels are bit-identical.
// ---------------------------------------------------------------------------------------
namespace {

const uint8_t kZigzag[64 + 15] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                  35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
                                  63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};   // tail: corrupt runs stay in bounds

struct JHuff {
    uint8_t bitsCount[17]; uint8_t values[256];
    int mincode[17], maxcode[18], valptr[17];
    bool present = false;
    void build() {
        int code = 0, k = 0;
        for (int len = 1; len <= 16; ++len) {
            valptr[len] = k; mincode[len] = code;
            code += bitsCount[len]; k += bitsCount[len];
            maxcode[len] = bitsCount[len] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        present = true;
    }
};

struct JComp {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
    int x = 0, y = 0;            // size in samples
    int w2 = 0, h2 = 0;          // MCU-padded plane size
    int dcPred = 0;
    std::vector<uint8_t> plane;  // reconstructed samples, w2 * h2
    std::vector<int16_t> coeff;  // progressive: (w2/8) * (h2/8) blocks of 64
};

struct JBits {
    const uint8_t* p; size_t n, pos; uint32_t buf = 0; int cnt = 0; int marker = -1;
    void reset() { buf = 0; cnt = 0; marker = -1; }
    void fill() {
        while (cnt <= 24) {
            uint32_t b = 0;
            if (marker < 0 && pos < n) {
                b = p[pos++];
                if (b == 0xff) {
                    uint32_t c = pos < n ? p[pos] : 0xd9u;
                    while (c == 0xff && pos + 1 < n) { ++pos; c = p[pos]; }
                    if (c == 0) ++pos;                       // stuffed zero
                    else { marker = (int)c; ++pos; b = 0; }  // a real marker ends the entropy-coded segment
                }
            }
            buf |= b << (24 - cnt); cnt += 8;
        }
    }
    int bits(int k) { if (k == 0) return 0; if (cnt < k) fill(); const int v = (int)(buf >> (32 - k)); buf <<= k; cnt -= k; return v; }
    int bit() { return bits(1); }
};

struct JDecoder {
    const uint8_t* p; size_t n, pos = 0;
    int w = 0, h = 0, ncomp = 0; bool progressive = false, jfif = false; int adobeTransform = -1;
    int hmax = 1, vmax = 1, mcuW = 0, mcuH = 0, mcuX = 0, mcuY = 0;
    uint16_t quant[4][64] = {}; bool quantPresent[4] = {false, false, false, false};
    JHuff dc[4], ac[4];
    JComp comp[4];
    int restartInterval = 0;
    int specStart = 0, specEnd = 63, succHigh = 0, succLow = 0, eobRun = 0;
    int scanN = 0, scanOrder[4];
    JBits br;

    int decodeHuff(const JHuff& hf) {
        if (!hf.present) return -1;                                  // the scan names a table no DHT segment defined
        int code = 0;
        for (int len = 1; len <= 16; ++len) {
            code = (code << 1) | br.bit();
            if (hf.maxcode[len] >= 0 && code <= hf.maxcode[len] && code >= hf.mincode[len]) {
                const int idx = hf.valptr[len] + code - hf.mincode[len];
                return (idx >= 0 && idx < 256) ? hf.values[idx] : -1;   // over-subscribed length counts
            }
        }
        return -1;
    }
    int extend(int s) {   // RECEIVE + EXTEND (T.81 F.2.2.1)
        if (s == 0) return 0;
        const int v = br.bits(s);
        return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
    }
};

// 1-D pass of the 12-bit fixed-point inverse DCT (the STBI__IDCT_1D step, stb_image.h:2270-2307); outputs the even part x0..x3 and the odd part t0..t3.
struct Idct1D { int x0, x1, x2, x3, t0, t1, t2, t3; };
inline int f2f(double x) { return (int)(x * 4096 + 0.5); }
inline Idct1D idct1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7) {
    static const int c0541 = f2f(0.5411961f), c1847 = f2f(-1.847759065f), c0765 = f2f(0.765366865f), c1175 = f2f(1.175875602f), c0298 = f2f(0.298631336f),
                     c2053 = f2f(2.053119869f), c3072 = f2f(3.072711026f), c1501 = f2f(1.501321110f), c0899 = f2f(-0.899976223f), c2562 = f2f(-2.562915447f),
                     c1961 = f2f(-1.961570560f), c0390 = f2f(-0.390180644f);
    Idct1D r;
    int p1 = (s2 + s6) * c0541;
    const int e2 = p1 + s6 * c1847, e3 = p1 + s2 * c0765;
    const int e0 = (s0 + s4) * 4096, e1 = (s0 - s4) * 4096;
    r.x0 = e0 + e3; r.x3 = e0 - e3; r.x1 = e1 + e2; r.x2 = e1 - e2;
    int t0 = s7, t1 = s5, t2 = s3, t3 = s1;
    int p3 = t0 + t2, p4 = t1 + t3; p1 = t0 + t3; int p2 = t1 + t2;
    const int p5 = (p3 + p4) * c1175;
    t0 *= c0298; t1 *= c2053; t2 *= c3072; t3 *= c1501;
    p1 = p5 + p1 * c0899; p2 = p5 + p2 * c2562; p3 *= c1961; p4 *= c0390;
    r.t3 = t3 + p1 + p4; r.t2 = t2 + p2 + p3; r.t1 = t1 + p2 + p4; r.t0 = t0 + p1 + p3;
    return r;
}
inline uint8_t clamp8(int x) { return (uint8_t)(x < 0 ? 0 : x > 255 ? 255 : x); }

void idctBlock(uint8_t* out, int stride, const int16_t* d) {
    int val[64];
    for (int i = 0; i < 8; ++i) {
        const int16_t* c = d + i; int* v = val + i;
        if (!c[8] && !c[16] && !c[24] && !c[32] && !c[40] && !c[48] && !c[56]) {
            const int dcterm = c[0] * 4;
            for (int k = 0; k < 8; ++k) v[k * 8] = dcterm;
        } else {
            Idct1D r = idct1d(c[0], c[8], c[16], c[24], c[32], c[40], c[48], c[56]);
            r.x0 += 512; r.x1 += 512; r.x2 += 512; r.x3 += 512;
            v[0] = (r.x0 + r.t3) >> 10; v[56] = (r.x0 - r.t3) >> 10; v[8] = (r.x1 + r.t2) >> 10; v[48] = (r.x1 - r.t2) >> 10;
            v[16] = (r.x2 + r.t1) >> 10; v[40] = (r.x2 - r.t1) >> 10; v[24] = (r.x3 + r.t0) >> 10; v[32] = (r.x3 - r.t0) >> 10;
        }
    }
    for (int i = 0; i < 8; ++i) {
        const int* v = val + i * 8; uint8_t* o = out + (size_t)i * stride;
        Idct1D r = idct1d(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
        const int bias = 65536 + (128 << 17);
        r.x0 += bias; r.x1 += bias; r.x2 += bias; r.x3 += bias;
        o[0] = clamp8((r.x0 + r.t3) >> 17); o[7] = clamp8((r.x0 - r.t3) >> 17); o[1] = clamp8((r.x1 + r.t2) >> 17); o[6] = clamp8((r.x1 - r.t2) >> 17);
        o[2] = clamp8((r.x2 + r.t1) >> 17); o[5] = clamp8((r.x2 - r.t1) >> 17); o[3] = clamp8((r.x3 + r.t0) >> 17); o[4] = clamp8((r.x3 - r.t0) >> 17);
    }
}

} // namespace

namespace {

// ---- entropy-coded block decoding (T.81 F.2.2 baseline, G.1.2 progressive) ----
bool blockBaseline(JDecoder& j, JComp& c, int16_t* data) {
    const uint16_t* q = j.quant[c.tq];
    std::memset(data, 0, 64 * sizeof(int16_t));
    const int t = j.decodeHuff(j.dc[c.td]);
    if (t < 0 || t > 15) return false;
    const int diff = t ? j.extend(t) : 0;
    c.dcPred += diff;
    data[0] = (int16_t)(c.dcPred * q[0]);
    int k = 1;
    do {
        const int rs = j.decodeHuff(j.ac[c.ta]);
        if (rs < 0) return false;
        const int s = rs & 15, r = rs >> 4;
        if (s == 0) { if (rs != 0xf0) break; k += 16; }
        else { k += r; const int zig = kZigzag[k++]; data[zig] = (int16_t)(j.extend(s) * q[zig]); }
    } while (k < 64);
    return true;
}

bool blockProgDC(JDecoder& j, JComp& c, int16_t* data) {
    if (j.specEnd != 0) return false;
    if (j.succHigh == 0) {
        std::memset(data, 0, 64 * sizeof(int16_t));
        const int t = j.decodeHuff(j.dc[c.td]);
        if (t < 0 || t > 15) return false;
        const int diff = t ? j.extend(t) : 0;
        c.dcPred += diff;
        data[0] = (int16_t)(c.dcPred * (1 << j.succLow));
    } else if (j.br.bit()) data[0] += (int16_t)(1 << j.succLow);
    return true;
}

bool blockProgAC(JDecoder& j, JComp& c, int16_t* data) {
    if (j.specStart == 0) return false;
    const JHuff& hf = j.ac[c.ta];
    if (j.succHigh == 0) {
        const int shift = j.succLow;
        if (j.eobRun) { --j.eobRun; return true; }
        int k = j.specStart;
        do {
            const int rs = j.decodeHuff(hf);
            if (rs < 0) return false;
            const int s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (r < 15) { j.eobRun = 1 << r; if (r) j.eobRun += j.br.bits(r); --j.eobRun; break; }
                k += 16;
            } else { k += r; const int zig = kZigzag[k++]; data[zig] = (int16_t)(j.extend(s) * (1 << shift)); }
        } while (k <= j.specEnd);
    } else {
        const int16_t bit = (int16_t)(1 << j.succLow);
        auto refine = [&](int16_t* p) { if (j.br.bit() && (*p & bit) == 0) { if (*p > 0) *p += bit; else *p -= bit; } };
        if (j.eobRun) {
            --j.eobRun;
            for (int k = j.specStart; k <= j.specEnd; ++k) { int16_t* p = &data[kZigzag[k]]; if (*p != 0) refine(p); }
        } else {
            int k = j.specStart;
            do {
                const int rs = j.decodeHuff(hf);
                if (rs < 0) return false;
                int s = rs & 15, r = rs >> 4;
                if (s == 0) {
                    if (r < 15) { j.eobRun = (1 << r) - 1; if (r) j.eobRun += j.br.bits(r); r = 64; }   // rest of the block only gets refinement bits
                } else {
                    if (s != 1) return false;
                    s = j.br.bit() ? bit : -bit;
                }
                while (k <= j.specEnd) {
                    int16_t* p = &data[kZigzag[k++]];
                    if (*p != 0) refine(p);
                    else { if (r == 0) { *p = (int16_t)s; break; } --r; }
                }
            } while (k <= j.specEnd);
        }
    }
    return true;
}

bool handleRestart(JDecoder& j, int& todo) {
    if (--todo > 0) return true;
    if (j.br.cnt < 24) j.br.fill();
    if (!(j.br.marker >= 0xd0 && j.br.marker <= 0xd7)) return false;     // no restart marker: the scan ends here
    j.br.reset();
    for (int k = 0; k < 4; ++k) j.comp[k].dcPred = 0;
    j.eobRun = 0;
    todo = j.restartInterval ? j.restartInterval : 0x7fffffff;
    return true;
}

bool decodeScan(JDecoder& j) {
    j.br.reset();
    for (int k = 0; k < 4; ++k) j.comp[k].dcPred = 0;
    j.eobRun = 0;
    int todo = j.restartInterval ? j.restartInterval : 0x7fffffff;
    int16_t block[64];
    auto one = [&](JComp& c, int bx, int by) -> bool {
        if (!j.progressive) {
            if (!blockBaseline(j, c, block)) return false;
            idctBlock(c.plane.data() + (size_t)by * 8 * c.w2 + (size_t)bx * 8, c.w2, block);
            return true;
        }
        int16_t* data = c.coeff.data() + 64 * ((size_t)by * (c.w2 / 8) + bx);
        return j.specStart == 0 ? blockProgDC(j, c, data) : blockProgAC(j, c, data);
    };
    if (j.scanN == 1) {                                                  // non-interleaved: the component's own block grid
        JComp& c = j.comp[j.scanOrder[0]];
        const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
        for (int by = 0; by < bh; ++by) for (int bx = 0; bx < bw; ++bx) {
            if (!one(c, bx, by)) return false;
            if (!handleRestart(j, todo)) return true;
        }
    } else {
        for (int my = 0; my < j.mcuY; ++my) for (int mx = 0; mx < j.mcuX; ++mx) {
            for (int k = 0; k < j.scanN; ++k) {
                JComp& c = j.comp[j.scanOrder[k]];
                for (int y = 0; y < c.v; ++y) for (int x = 0; x < c.h; ++x)
                    if (!one(c, mx * c.h + x, my * c.v + y)) return false;
            }
            if (!handleRestart(j, todo)) return true;
        }
    }
    return true;
}

void finishProgressive(JDecoder& j) {
    for (int n = 0; n < j.ncomp; ++n) {
        JComp& c = j.comp[n];
        const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
        for (int by = 0; by < bh; ++by) for (int bx = 0; bx < bw; ++bx) {
            int16_t* data = c.coeff.data() + 64 * ((size_t)by * (c.w2 / 8) + bx);
            for (int i = 0; i < 64; ++i) data[i] = (int16_t)(data[i] * j.quant[c.tq][i]);
            idctBlock(c.plane.data() + (size_t)by * 8 * c.w2 + (size_t)bx * 8, c.w2, data);
        }
    }
}

} // namespace

namespace {

inline int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

// ---- chroma upsampling: one output row from the nearer and the farther low-resolution row ----
inline uint8_t div4(int x) { return (uint8_t)(x >> 2); }
inline uint8_t div16(int x) { return (uint8_t)(x >> 4); }
const uint8_t* upsampleRow(uint8_t* out, const uint8_t* nearRow, const uint8_t* farRow, int w, int hs, int vs) {
    if (hs == 1 && vs == 1) return nearRow;
    if (hs == 1 && vs == 2) { for (int i = 0; i < w; ++i) out[i] = div4(3 * nearRow[i] + farRow[i] + 2); return out; }
    if (hs == 2 && vs == 1) {
        const uint8_t* in = nearRow;
        if (w == 1) { out[0] = out[1] = in[0]; return out; }
        out[0] = in[0]; out[1] = div4(in[0] * 3 + in[1] + 2);
        int i;
        for (i = 1; i < w - 1; ++i) { const int n = 3 * in[i] + 2; out[i * 2] = div4(n + in[i - 1]); out[i * 2 + 1] = div4(n + in[i + 1]); }
        out[i * 2] = div4(in[w - 2] * 3 + in[w - 1] + 2); out[i * 2 + 1] = in[w - 1];
        return out;
    }
    if (hs == 2 && vs == 2) {
        if (w == 1) { out[0] = out[1] = div4(3 * nearRow[0] + farRow[0] + 2); return out; }
        int t1 = 3 * nearRow[0] + farRow[0];
        out[0] = div4(t1 + 2);
        for (int i = 1; i < w; ++i) {
            const int t0 = t1; t1 = 3 * nearRow[i] + farRow[i];
            out[i * 2 - 1] = div16(3 * t0 + t1 + 8); out[i * 2] = div16(3 * t1 + t0 + 8);
        }
        out[w * 2 - 1] = div4(t1 + 2);
        return out;
    }
    for (int i = 0; i < w; ++i) for (int k = 0; k < hs; ++k) out[i * hs + k] = nearRow[i];     // other factors: nearest
    return out;
}

inline int fixed20(float x) { return ((int)(x * 4096.0f + 0.5f)) << 8; }

bool decodeJPEG(const uint8_t* p, size_t n, Decoded& img) {
    if (n < 4 || p[0] != 0xff || p[1] != 0xd8) return false;
    JDecoder j; j.p = p; j.n = n; j.pos = 2;
    bool gotFrame = false, sawScan = false; int rgbIds = 0;
    for (;;) {
        while (j.pos < n && p[j.pos] != 0xff) ++j.pos;                   // resynchronise on the next marker
        while (j.pos < n && p[j.pos] == 0xff) ++j.pos;
        if (j.pos >= n) break;
        const int m = p[j.pos++];
        if (m == 0xd9) break;
        if (m == 0x00 || m == 0x01 || (m >= 0xd0 && m <= 0xd7)) continue;
        if (j.pos + 2 > n) return false;
        const int len = be16(p + j.pos);
        if (len < 2 || j.pos + (size_t)len > n) return false;
        const uint8_t* d = p + j.pos + 2; const int dl = len - 2;
        if (m == 0xe0 && dl >= 5 && !std::memcmp(d, "JFIF\0", 5)) j.jfif = true;
        else if (m == 0xee && dl >= 12 && !std::memcmp(d, "Adobe\0", 6)) j.adobeTransform = d[11];
        else if (m == 0xdb) {                                            // DQT
            int o = 0;
            while (o < dl) {
                const int pq = d[o] >> 4, tq = d[o] & 15; ++o;
                if (pq > 1 || tq > 3 || o + (pq ? 128 : 64) > dl) return false;
                for (int i = 0; i < 64; ++i) { j.quant[tq][kZigzag[i]] = (uint16_t)(pq ? be16(d + o + 2 * i) : d[o + i]); }
                o += pq ? 128 : 64; j.quantPresent[tq] = true;
            }
        } else if (m == 0xc4) {                                          // DHT
            int o = 0;
            while (o + 17 <= dl) {
                const int tc = d[o] >> 4, th = d[o] & 15; ++o;
                if (tc > 1 || th > 3) return false;
                JHuff& hf = tc ? j.ac[th] : j.dc[th];
                int total = 0; hf.bitsCount[0] = 0;
                for (int i = 1; i <= 16; ++i) { hf.bitsCount[i] = d[o + i - 1]; total += hf.bitsCount[i]; }
                o += 16;
                if (total > 256 || o + total > dl) return false;
                std::memcpy(hf.values, d + o, total); o += total;
                hf.build();
            }
        } else if (m == 0xdd) { if (dl < 2) return false; j.restartInterval = be16(d); }
        else if (m == 0xc0 || m == 0xc1 || m == 0xc2) {                  // SOF0/1/2
            if (gotFrame || dl < 6 || d[0] != 8) return false;
            j.progressive = m == 0xc2;
            j.h = be16(d + 1); j.w = be16(d + 3); j.ncomp = d[5];
            if (!dimsOk(j.w, j.h) || (j.ncomp != 1 && j.ncomp != 3) || dl < 6 + 3 * j.ncomp) return false;
            static const char rgb[3] = {'R', 'G', 'B'};
            for (int i = 0; i < j.ncomp; ++i) {
                JComp& c = j.comp[i];
                c.id = d[6 + 3 * i]; c.h = d[7 + 3 * i] >> 4; c.v = d[7 + 3 * i] & 15; c.tq = d[8 + 3 * i];
                if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4 || c.tq > 3) return false;
                if (j.ncomp == 3 && c.id == rgb[i]) ++rgbIds;
                j.hmax = std::max(j.hmax, c.h); j.vmax = std::max(j.vmax, c.v);
            }
            j.mcuW = j.hmax * 8; j.mcuH = j.vmax * 8;
            j.mcuX = (j.w + j.mcuW - 1) / j.mcuW; j.mcuY = (j.h + j.mcuH - 1) / j.mcuH;
            for (int i = 0; i < j.ncomp; ++i) {
                JComp& c = j.comp[i];
                c.x = (j.w * c.h + j.hmax - 1) / j.hmax; c.y = (j.h * c.v + j.vmax - 1) / j.vmax;
                c.w2 = j.mcuX * c.h * 8; c.h2 = j.mcuY * c.v * 8;
                c.plane.assign((size_t)c.w2 * c.h2, 0);
                if (j.progressive) c.coeff.assign((size_t)c.w2 * c.h2, 0);
            }
            gotFrame = true;
        } else if (m == 0xc3 || (m >= 0xc5 && m <= 0xcf && m != 0xc8 && m != 0xcc)) return false;   // lossless / arithmetic / hierarchical: unsupported
        else if (m == 0xda) {                                            // SOS
            if (!gotFrame || dl < 1) return false;
            j.scanN = d[0];
            if (j.scanN < 1 || j.scanN > j.ncomp || dl < 1 + 2 * j.scanN + 3) return false;
            for (int i = 0; i < j.scanN; ++i) {
                const int id = d[1 + 2 * i]; int which = -1;
                for (int k = 0; k < j.ncomp; ++k) if (j.comp[k].id == id) which = k;
                if (which < 0) return false;
                j.comp[which].td = d[2 + 2 * i] >> 4; j.comp[which].ta = d[2 + 2 * i] & 15;
                if (j.comp[which].td > 3 || j.comp[which].ta > 3) return false;
                j.scanOrder[i] = which;
            }
            const uint8_t* t = d + 1 + 2 * j.scanN;
            j.specStart = t[0]; j.specEnd = t[1]; j.succHigh = t[2] >> 4; j.succLow = t[2] & 15;
            if (j.progressive) { if (j.specStart > 63 || j.specEnd > 63 || j.specStart > j.specEnd || j.succHigh > 13 || j.succLow > 13) return false; }
            else { if (j.specStart != 0 || j.succHigh != 0 || j.succLow != 0) return false; j.specEnd = 63; }
            j.br.p = p; j.br.n = n; j.br.pos = j.pos + (size_t)len;
            if (!decodeScan(j)) return false;
            sawScan = true;
            j.pos = j.br.pos;
            if (j.br.marker >= 0) { if (j.br.marker == 0xd9) break; j.pos -= 2; }       // re-read the marker that ended the scan
            continue;
        }
        j.pos += (size_t)len;
    }
    if (!gotFrame || !sawScan) return false;
    if (j.progressive) finishProgressive(j);

    // resample + colour conversion
    const bool isRgb = j.ncomp == 3 && (rgbIds == 3 || (j.adobeTransform == 0 && !j.jfif));
    img.w = j.w; img.h = j.h; img.comp = j.ncomp; img.isFloat = false;
    img.u8.assign((size_t)j.w * j.h * j.ncomp, 0);
    struct Res { int hs, vs, ystep, wLores, ypos; const uint8_t* line0; const uint8_t* line1; std::vector<uint8_t> buf; } res[3];
    for (int k = 0; k < j.ncomp; ++k) {
        Res& r = res[k]; const JComp& c = j.comp[k];
        r.hs = j.hmax / c.h; r.vs = j.vmax / c.v; r.ystep = r.vs >> 1; r.wLores = (j.w + r.hs - 1) / r.hs; r.ypos = 0;
        r.line0 = r.line1 = c.plane.data(); r.buf.assign((size_t)j.w + 8, 0);
    }
    const int kR = fixed20(1.40200f), kGr = fixed20(0.71414f), kGb = fixed20(0.34414f), kB = fixed20(1.77200f);
    for (int y = 0; y < j.h; ++y) {
        const uint8_t* row[3] = {nullptr, nullptr, nullptr};
        for (int k = 0; k < j.ncomp; ++k) {
            Res& r = res[k];
            const bool bot = r.ystep >= (r.vs >> 1);
            row[k] = upsampleRow(r.buf.data(), bot ? r.line1 : r.line0, bot ? r.line0 : r.line1, r.wLores, r.hs, r.vs);
            if (++r.ystep >= r.vs) { r.ystep = 0; r.line0 = r.line1; if (++r.ypos < j.comp[k].y) r.line1 += j.comp[k].w2; }
        }
        uint8_t* o = &img.u8[(size_t)y * j.w * j.ncomp];
        if (j.ncomp == 1) { std::memcpy(o, row[0], j.w); continue; }
        for (int x = 0; x < j.w; ++x, o += 3) {
            if (isRgb) { o[0] = row[0][x]; o[1] = row[1][x]; o[2] = row[2][x]; continue; }
            const int yf = (row[0][x] << 20) + (1 << 19), cr = row[2][x] - 128, cb = row[1][x] - 128;
            const int r = (yf + cr * kR) >> 20;
            const int g = (int)(yf + (cr * -kGr) + (int)((uint32_t)(cb * -kGb) & 0xffff0000u)) >> 20;
            const int b = (yf + cb * kB) >> 20;
            o[0] = clamp8(r); o[1] = clamp8(g); o[2] = clamp8(b);
        }
    }
    return true;
}

} // namespace

// ---------------------------------------------------------------------------------------
// Public entry points (include/heatray_host.h)
// ---------------------------------------------------------------------------------------
namespace {

int decodeAny(const uint8_t* p, size_t n, int flip, HrhImage* out) {
    if (!out) return imgFail("null output");
    std::memset(out, 0, sizeof(*out));
    Decoded d;
    bool ok;
    try {                                                                    // no exception may cross the C boundary
        if (n >= 8 && p[0] == 137 && p[1] == 'P') ok = decodePNG(p, n, d);
        else if (n >= 3 && p[0] == 0xff && p[1] == 0xd8) ok = decodeJPEG(p, n, d);
        else if (isHDR(p, n)) ok = decodeHDR(p, n, d);
        else if (isTIFF(p, n)) ok = decodeTIFF(p, n, d);
        else return imgFail("unknown image type (PNG, JPEG, Radiance HDR and baseline TIFF are decoded here)");
    } catch (const std::bad_alloc&) { g_imgErr = "out of memory"; return HRC_ERR_OUT_OF_MEMORY; }
      catch (const std::exception&) { return imgFail("corrupt or unsupported image data"); }
    if (!ok) return imgFail("corrupt or unsupported image data");
    const size_t elems = (size_t)d.w * d.h * d.comp, bytes = elems * (d.isFloat ? 4 : 1);
    void* mem = std::malloc(bytes ? bytes : 1);
    if (!mem) { g_imgErr = "out of memory"; return HRC_ERR_OUT_OF_MEMORY; }
    std::memcpy(mem, d.isFloat ? (const void*)d.f32.data() : (const void*)d.u8.data(), bytes);
    if (flip) flipRows(mem, d.h, (size_t)d.w * d.comp * (d.isFloat ? 4 : 1));
    out->width = d.w; out->height = d.h; out->channels = d.comp; out->isFloat = d.isFloat ? 1 : 0; out->data = mem;
    return HRC_OK;
}

bool writeAll(const char* path, const std::vector<uint8_t>& bytes) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    return std::fclose(f) == 0 && ok;
}
void put32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
void pngChunk(std::vector<uint8_t>& out, const char* type, const std::vector<uint8_t>& data) {
    put32(out, (uint32_t)data.size());
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put32(out, crc32(out.data() + start, out.size() - start));
}

} // namespace

extern "C" {

HRC_API const char* hrhImageLastError(void) { return g_imgErr.c_str(); }

HRC_API int hrhLoadImageFromMemory(const void* bytes, size_t size, int32_t flipVertically, HrhImage* out) {
    if (!bytes) return imgFail("null input");
    return decodeAny((const uint8_t*)bytes, size, flipVertically, out);
}
HRC_API int hrhLoadImage(const char* path, int32_t flipVertically, HrhImage* out) {
    std::vector<uint8_t> file;
    if (!path || !readFile(path, file)) return imgFail("cannot read image file");
    return decodeAny(file.data(), file.size(), flipVertically, out);
}
HRC_API void hrhFreeImage(HrhImage* img) { if (img && img->data) { std::free(img->data); img->data = nullptr; } }

// util::loadTextureInternal, TextureLoader.cpp:17-137
HRC_API int hrhLoadTexture(const char* path, int32_t generateMips, int32_t convertToLinear, HrhLoadedTexture* out) {
    if (!path || !out) return imgFail("null argument");
    std::memset(out, 0, sizeof(*out));
    std::string finalPath = path;                                           // :20-35: the file may be one directory back
    std::vector<uint8_t> file;
    if (!readFile(finalPath.c_str(), file)) {
        finalPath = "../" + finalPath;
        if (!readFile(finalPath.c_str(), file)) return imgFail("unable to find texture");
    }
    out->wrapS = out->wrapT = generateMips ? HRC_WRAP_REPEAT : HRC_WRAP_CLAMP;   // :37-42 vs the default sampler (Texture.h:43-55)
    const int rc = decodeAny(file.data(), file.size(), 1, &out->image);     // :89 stbi_set_flip_vertically_on_load(true)
    if (rc != HRC_OK) return rc;
    HrhImage& im = out->image;
    if (!im.isFloat && convertToLinear) {                                   // :109-133, alpha (channel 3) untouched
        uint8_t* px = (uint8_t*)im.data;
        const size_t count = (size_t)im.width * im.height;
        for (size_t i = 0; i < count; ++i) for (int ch = 0; ch < im.channels; ++ch) {
            if (ch == 3) continue;
            float v = float(px[i * im.channels + ch]) / 255.0f;
            if (v <= 0.04045f) v /= 12.92f; else v = powf((v + 0.055f) / (1.0f + 0.055f), 2.4f);
            px[i * im.channels + ch] = uint8_t(v * 255.0f);                 // truncation, like the reference
        }
    }
    switch (im.channels) {                                                  // :138-151
    case 1: out->format = im.isFloat ? HRC_TEX_R32F : HRC_TEX_RGBA8; break;
    case 3: out->format = im.isFloat ? HRC_TEX_RGB32F : HRC_TEX_RGBA8; break;
    case 4: out->format = im.isFloat ? HRC_TEX_RGBA32F : HRC_TEX_RGBA8; break;
    default: hrhFreeImage(&im); return imgFail("unsupported channel count (the reference skips such textures)");
    }
    return HRC_OK;
}

HRC_API int hrhExpandToRGBA8(const HrhImage* img, uint8_t* out) {
    if (!img || !img->data || !out || img->isFloat) return imgFail("expects a uint8 image");
    const uint8_t* s = (const uint8_t*)img->data; const size_t n = (size_t)img->width * img->height; const int c = img->channels;
    for (size_t i = 0; i < n; ++i) {
        if (c == 1) { out[4 * i] = out[4 * i + 1] = out[4 * i + 2] = s[i]; out[4 * i + 3] = 255; }            // RL_LUMINANCE: (l, l, l, 1)
        else if (c == 2) { out[4 * i] = out[4 * i + 1] = out[4 * i + 2] = s[2 * i]; out[4 * i + 3] = s[2 * i + 1]; }
        else if (c == 3) { out[4 * i] = s[3 * i]; out[4 * i + 1] = s[3 * i + 1]; out[4 * i + 2] = s[3 * i + 2]; out[4 * i + 3] = 255; }
        else std::memcpy(out + 4 * i, s + 4 * i, 4);
    }
    return HRC_OK;
}

// PNG writer: 8-bit grey / grey+alpha / RGB / RGBA, filter 0, zlib "stored" blocks (valid everywhere, no compressor needed).
HRC_API int hrhWritePNG(const char* path, const uint8_t* pixels, int32_t w, int32_t h, int32_t channels, int32_t rowsBottomUp) {
    if (!path || !pixels || w <= 0 || h <= 0 || channels < 1 || channels > 4) return imgFail("bad PNG arguments");
    std::vector<uint8_t> out = {137, 80, 78, 71, 13, 10, 26, 10}, hdr, raw, z;
    put32(hdr, (uint32_t)w); put32(hdr, (uint32_t)h);
    static const uint8_t ctype[5] = {0, 0, 4, 2, 6};
    hdr.push_back(8); hdr.push_back(ctype[channels]); hdr.push_back(0); hdr.push_back(0); hdr.push_back(0);
    pngChunk(out, "IHDR", hdr);
    const size_t stride = (size_t)w * channels;
    raw.reserve((stride + 1) * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* row = pixels + (size_t)(rowsBottomUp ? h - 1 - y : y) * stride;
        raw.push_back(0);
        raw.insert(raw.end(), row, row + stride);
    }
    z.push_back(0x78); z.push_back(0x01);
    size_t off = 0;
    while (off < raw.size() || off == 0) {
        const size_t k = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + k >= raw.size() ? 1 : 0);
        z.push_back(k & 0xff); z.push_back(k >> 8); z.push_back(~k & 0xff); z.push_back((~k >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + k);
        off += k;
        if (k == 0) break;
    }
    put32(z, adler32(raw.data(), raw.size()));
    pngChunk(out, "IDAT", z);
    pngChunk(out, "IEND", {});
    return writeAll(path, out) ? HRC_OK : imgFail("cannot write PNG file");
}

// Radiance RGBE writer (flat scanlines), rows top-down in the file.
HRC_API int hrhWriteHDR(const char* path, const float* rgb, int32_t w, int32_t h, int32_t rowsBottomUp) {
    if (!path || !rgb || w <= 0 || h <= 0) return imgFail("bad HDR arguments");
    char head[128];
    const int hl = std::snprintf(head, sizeof(head), "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
    std::vector<uint8_t> out(head, head + hl);
    out.reserve(out.size() + (size_t)w * h * 4);
    for (int y = 0; y < h; ++y) {
        const float* row = rgb + (size_t)(rowsBottomUp ? h - 1 - y : y) * w * 3;
        for (int x = 0; x < w; ++x) {
            const float r = row[3 * x], g = row[3 * x + 1], b = row[3 * x + 2];
            const float m = std::max(r, std::max(g, b));
            uint8_t e[4] = {0, 0, 0, 0};
            if (m >= 1e-32f && std::isfinite(m)) {
                int ex; const float f = std::frexp(m, &ex) * 256.0f / m;
                e[0] = (uint8_t)std::max(0.0f, r * f); e[1] = (uint8_t)std::max(0.0f, g * f); e[2] = (uint8_t)std::max(0.0f, b * f); e[3] = (uint8_t)(ex + 128);
            }
            out.insert(out.end(), e, e + 4);
        }
    }
    return writeAll(path, out) ? HRC_OK : imgFail("cannot write HDR file");
}

// PFM: lossless float RGB, little-endian, rows bottom-up in the file (the format's own convention).
HRC_API int hrhWritePFM(const char* path, const float* rgb, int32_t w, int32_t h, int32_t rowsBottomUp) {
    if (!path || !rgb || w <= 0 || h <= 0) return imgFail("bad PFM arguments");
    char head[64];
    const int hl = std::snprintf(head, sizeof(head), "PF\n%d %d\n-1.0\n", w, h);
    std::vector<uint8_t> out(head, head + hl);
    const size_t rowBytes = (size_t)w * 12;
    for (int y = 0; y < h; ++y) {
        const uint8_t* row = (const uint8_t*)rgb + (size_t)(rowsBottomUp ? y : h - 1 - y) * rowBytes;
        out.insert(out.end(), row, row + rowBytes);
    }
    return writeAll(path, out) ? HRC_OK : imgFail("cannot write PFM file");
}

} // extern "C"
