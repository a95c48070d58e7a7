// imageio.cpp -- image readers of the scene loader: what `new Bitmap(path)` + `Bitmap::convert(RGB, Float32, srgb_gamma = false)` do for the
// environment emitter (src/emitters/envmap.cpp:130-146) and for bitmap textures, for the formats the reference's own test scenes use.
//   * OpenEXR (src/core/bitmap.cpp:1100-1330 read_exr goes through the OpenEXR library, which is absent from the reference tree --
//     ext/openexr is an empty submodule -- so the format is decoded here from its published layout): single-part scan-line files, HALF /
//     FLOAT / UINT channels, NO_COMPRESSION and PIZ (wavelet + Huffman: the compression Bitmap::write_exr picks, bitmap.cpp:1350, hence what
//     resources/data/common/textures/museum.exr and every refs/*.exr use).  Tiled, multi-part, deep, sub-sampled and other compressions raise.
//   * PFM (bitmap.cpp:2213-2268 read_pfm): `PF` / `Pf`, bottom row first, sign of the scale = byte order.
// Everything is host code; pixels come back as linear RGB float32, top row first.
#include "host.h"
#include <algorithm>
#include <cstring>
#include <fstream>
#include <sstream>

namespace amvpt {
namespace {

struct Bytes {
    const uint8_t *p, *end; const std::string &path;
    [[noreturn]] void fail(const char *what) const { AM_THROW("io: \"%s\": %s", path.c_str(), what); }
    void need(size_t n) const { if ((size_t) (end - p) < n) fail("unexpected end of file"); }
    template <class T> T get() { need(sizeof(T)); T v; std::memcpy(&v, p, sizeof(T)); p += sizeof(T); return v; }
    std::string cstr() { const uint8_t *b = p; while (p < end && *p) ++p; if (p >= end) fail("unterminated string"); std::string s((const char *) b, p - b); ++p; return s; }
};

float half_to_float(uint16_t h) {
    uint32_t s = (uint32_t) (h >> 15) << 31, e = (h >> 10) & 31u, m = h & 1023u, u;
    if (e == 0) {
        if (m == 0) u = s;
        else { int sh = 0; while (!(m & 1024u)) { m <<= 1; ++sh; } u = s | ((uint32_t) (127 - 15 - sh + 1) << 23) | ((m & 1023u) << 13); }
    } else if (e == 31) u = s | 0x7f800000u | (m << 13);
    else u = s | ((e + 112u) << 23) | (m << 13);
    float f; std::memcpy(&f, &u, 4); return f;
}

// ---------------------------------------------------------------------------------------- PIZ: Huffman stage
// Canonical code over the 16-bit alphabet + one run-length symbol: code lengths packed 6 bits each with zero-run escapes; codes
// are assigned from the longest length down, symbols of equal length in increasing order; the bit stream is read MSB first.
struct BitReader {
    const uint8_t *p, *end; uint64_t c = 0; int lc = 0;
    uint32_t bits(int n) { while (lc < n) { c = (c << 8) | (p < end ? *p : 0); ++p; lc += 8; } lc -= n; return (uint32_t) ((c >> lc) & ((1ull << n) - 1)); }
};
void huf_uncompress(const uint8_t *src, size_t n_src, uint16_t *out, size_t n_out, const Bytes &ctx) {
    if (n_src == 0) { if (n_out) ctx.fail("PIZ: empty Huffman block"); return; }
    if (n_src < 20) ctx.fail("PIZ: truncated Huffman header");
    uint32_t im, iM, n_bits; std::memcpy(&im, src, 4); std::memcpy(&iM, src + 4, 4); std::memcpy(&n_bits, src + 12, 4);
    const uint32_t ENC = (1u << 16) + 1;
    if (im >= ENC || iM >= ENC || im > iM) ctx.fail("PIZ: invalid Huffman symbol range");
    std::vector<uint8_t> len(ENC, 0);
    BitReader br{ src + 20, src + n_src };
    for (uint32_t i = im; i <= iM; ++i) {
        uint32_t l = br.bits(6);
        if (l == 63) { uint32_t run = br.bits(8) + 6; if (i + run > iM + 1) ctx.fail("PIZ: Huffman table overrun"); i += run - 1; }
        else if (l >= 59) { uint32_t run = l - 59 + 2; if (i + run > iM + 1) ctx.fail("PIZ: Huffman table overrun"); i += run - 1; }
        else len[i] = (uint8_t) l;
    }
    // canonical codes
    uint64_t count[59] = { 0 }, first[59] = { 0 };
    for (uint32_t i = 0; i < ENC; ++i) count[len[i]]++;
    { uint64_t c = 0; for (int l = 58; l > 0; --l) { uint64_t nc = (c + count[l]) >> 1; first[l] = c; c = nc; } }
    std::vector<uint32_t> offs(60, 0), syms;
    { uint32_t acc = 0; for (int l = 1; l < 59; ++l) { offs[l] = acc; acc += (uint32_t) count[l]; } offs[59] = acc; syms.resize(acc); }
    { std::vector<uint32_t> fill(offs.begin(), offs.end()); for (uint32_t i = 0; i < ENC; ++i) if (len[i]) syms[fill[len[i]]++] = i; }
    // 12-bit first-level lookup for the short codes, bit-serial continuation for the long ones
    const int LUT = 12; std::vector<uint32_t> lut(1u << LUT, 0);       // (symbol << 6) | length, 0 = longer than LUT bits
    for (int l = 1; l <= LUT; ++l) for (uint64_t k = 0; k < count[l]; ++k) {
        uint64_t code = first[l] + k; uint32_t sym = syms[offs[l] + (uint32_t) k];
        for (uint32_t fillb = 0; fillb < (1u << (LUT - l)); ++fillb) lut[(size_t) ((code << (LUT - l)) | fillb)] = (sym << 6) | (uint32_t) l;
    }
    const uint8_t *data = br.p - (br.lc / 8);      // the table is byte-padded: the code stream starts at the next whole byte
    BitReader d{ data, src + n_src };
    const uint32_t rlc = iM; size_t o = 0; uint64_t used = 0;
    while (o < n_out) {
        if (used >= n_bits) ctx.fail("PIZ: Huffman stream ends early");
        while (d.lc <= 56) { d.c = (d.c << 8) | (d.p < d.end ? *d.p : 0); ++d.p; d.lc += 8; }      // 57..64 bits buffered
        uint32_t sym; int l;
        uint32_t peek = (uint32_t) ((d.c >> (d.lc - LUT)) & ((1u << LUT) - 1)), e = lut[peek];
        if (e) { sym = e >> 6; l = (int) (e & 63u); }
        else {
            l = LUT; uint64_t code = peek; sym = 0xffffffffu;
            while (l < 58) {
                ++l; if (d.lc < l) ctx.fail("PIZ: Huffman code too long");
                code = (d.c >> (d.lc - l)) & ((1ull << l) - 1);
                if (count[l] && code >= first[l] && code - first[l] < count[l]) { sym = syms[offs[l] + (uint32_t) (code - first[l])]; break; }
            }
            if (sym == 0xffffffffu) ctx.fail("PIZ: invalid Huffman code");
        }
        d.lc -= l; used += (uint64_t) l;
        if (sym == rlc) {
            if (d.lc < 8) { d.c = (d.c << 8) | (d.p < d.end ? *d.p : 0); ++d.p; d.lc += 8; }
            d.lc -= 8; used += 8; uint32_t run = (uint32_t) ((d.c >> d.lc) & 255u);
            if (o == 0 || o + run > n_out) ctx.fail("PIZ: invalid run");
            uint16_t v = out[o - 1]; while (run--) out[o++] = v;
        } else out[o++] = (uint16_t) sym;
    }
}

// ---------------------------------------------------------------------------------------- PIZ: wavelet stage (2-D Haar, integer lifting)
inline void wdec14(uint16_t l, uint16_t h, uint16_t &a, uint16_t &b) {
    int ls = (int16_t) l, hi = (int16_t) h, ai = ls + (hi & 1) + (hi >> 1);
    a = (uint16_t) (int16_t) ai; b = (uint16_t) (int16_t) (ai - hi);
}
inline void wdec16(uint16_t l, uint16_t h, uint16_t &a, uint16_t &b) {
    int m = l, d = h, bb = (m - (d >> 1)) & 0xffff, aa = (d + bb - 0x8000) & 0xffff;
    b = (uint16_t) bb; a = (uint16_t) aa;
}
void wav2_decode(uint16_t *in, int nx, int ox, int ny, int oy, uint16_t mx) {
    const bool w14 = mx < (1 << 14);
    int n = nx > ny ? ny : nx, p = 1, p2;
    while (p <= n) p <<= 1;
    p >>= 1; p2 = p; p >>= 1;
    auto dec = [&](uint16_t l, uint16_t h, uint16_t &a, uint16_t &b) { if (w14) wdec14(l, h, a, b); else wdec16(l, h, a, b); };
    while (p >= 1) {
        uint16_t *py = in, *ey = in + (ptrdiff_t) oy * (ny - p2);
        const int oy1 = oy * p, oy2 = oy * p2, ox1 = ox * p, ox2 = ox * p2;
        uint16_t i00, i01, i10, i11;
        for (; py <= ey; py += oy2) {
            uint16_t *px = py, *ex = py + (ptrdiff_t) ox * (nx - p2);
            for (; px <= ex; px += ox2) {
                uint16_t *p01 = px + ox1, *p10 = px + oy1, *p11 = p10 + ox1;
                dec(*px, *p10, i00, i10); dec(*p01, *p11, i01, i11);
                dec(i00, i01, *px, *p01); dec(i10, i11, *p10, *p11);
            }
            if (nx & p) { uint16_t *p10 = px + oy1; dec(*px, *p10, i00, *p10); *px = i00; }
        }
        if (ny & p) {
            uint16_t *px = py, *ex = py + (ptrdiff_t) ox * (nx - p2);
            for (; px <= ex; px += ox2) { uint16_t *p01 = px + ox1; dec(*px, *p01, i00, *p01); *px = i00; }
        }
        p2 = p; p >>= 1;
    }
}

struct Channel { std::string name; int type; };       // 0 UINT, 1 HALF, 2 FLOAT

void read_exr(const std::string &path, const std::vector<uint8_t> &file, uint32_t *w_out, uint32_t *h_out, std::vector<float> &rgb) {
    Bytes b{ file.data(), file.data() + file.size(), path };
    if (b.get<uint32_t>() != 20000630u) b.fail("not an OpenEXR file");
    const uint32_t version = b.get<uint32_t>();
    if ((version & 0xffu) != 2) b.fail("unsupported OpenEXR version");
    if (version & 0x200u) b.fail("tiled OpenEXR files are not supported (scan-line only)");
    if (version & 0x1800u) b.fail("multi-part / deep OpenEXR files are not supported");
    std::vector<Channel> ch; int compression = -1, line_order = 0; int32_t dw[4] = { 0, 0, -1, -1 };
    for (;;) {
        std::string name = b.cstr(); if (name.empty()) break;
        std::string type = b.cstr(); const uint32_t size = b.get<uint32_t>(); b.need(size);
        Bytes a{ b.p, b.p + size, path }; b.p += size;
        if (name == "channels") {
            for (;;) {
                std::string cn = a.cstr(); if (cn.empty()) break;
                Channel c; c.name = cn; c.type = a.get<int32_t>(); a.get<uint32_t>();      // pLinear + 3 reserved bytes
                const int32_t xs = a.get<int32_t>(), ys = a.get<int32_t>();
                if (xs != 1 || ys != 1) b.fail("sub-sampled OpenEXR channels are not supported");
                if (c.type < 0 || c.type > 2) b.fail("unknown OpenEXR pixel type");
                ch.push_back(c);
            }
        } else if (name == "compression") compression = a.get<uint8_t>();
        else if (name == "dataWindow") { for (int i = 0; i < 4; ++i) dw[i] = a.get<int32_t>(); }
        else if (name == "lineOrder") line_order = a.get<uint8_t>();
        (void) type;
    }
    if (ch.empty() || compression < 0 || dw[2] < dw[0] || dw[3] < dw[1]) b.fail("incomplete OpenEXR header");
    if (compression != 0 && compression != 4) AM_THROW("io: \"%s\": OpenEXR compression %d is not supported (NO_COMPRESSION and PIZ are)", path.c_str(), compression);
    (void) line_order;      // chunks carry their own y coordinate; the offset table gives their positions whatever the order
    const uint32_t W = (uint32_t) (dw[2] - dw[0] + 1), H = (uint32_t) (dw[3] - dw[1] + 1);
    if ((uint64_t) W * H > (1ull << 28)) b.fail("image too large");
    // channel -> colour: R G B (also with a layer prefix "x.R"), or a single luminance channel Y
    int idx[3] = { -1, -1, -1 };
    for (size_t i = 0; i < ch.size(); ++i) {
        std::string n = ch[i].name; size_t dot = n.rfind('.'); if (dot != std::string::npos) n = n.substr(dot + 1);
        if (n == "R" && idx[0] < 0) idx[0] = (int) i; else if (n == "G" && idx[1] < 0) idx[1] = (int) i; else if (n == "B" && idx[2] < 0) idx[2] = (int) i;
    }
    if (idx[0] < 0 || idx[1] < 0 || idx[2] < 0) {
        int y = -1; for (size_t i = 0; i < ch.size(); ++i) if (ch[i].name == "Y") y = (int) i;
        if (y < 0 && ch.size() == 1) y = 0;
        if (y < 0) b.fail("OpenEXR file has neither R, G, B nor a luminance channel");
        idx[0] = idx[1] = idx[2] = y;
    }
    std::vector<size_t> ch_off(ch.size()); size_t line_bytes = 0;
    for (size_t i = 0; i < ch.size(); ++i) { ch_off[i] = line_bytes; line_bytes += (size_t) W * (ch[i].type == 1 ? 2 : 4); }
    const uint32_t lines_per_block = compression == 4 ? 32u : 1u, n_blocks = (H + lines_per_block - 1) / lines_per_block;
    std::vector<uint64_t> offsets(n_blocks); for (auto &o : offsets) o = b.get<uint64_t>();
    rgb.assign((size_t) W * H * 3, 0.f);
    std::vector<uint8_t> raw; std::vector<uint16_t> tmp;
    for (uint32_t blk = 0; blk < n_blocks; ++blk) {
        if (offsets[blk] + 8 > file.size()) b.fail("chunk offset outside the file");
        Bytes c{ file.data() + offsets[blk], file.data() + file.size(), path };
        const int32_t y0 = c.get<int32_t>(); const uint32_t size = c.get<uint32_t>(); c.need(size);
        if (y0 < dw[1] || y0 > dw[3]) b.fail("chunk outside the data window");
        const uint32_t row0 = (uint32_t) (y0 - dw[1]), rows = std::min(lines_per_block, H - row0);
        const size_t raw_size = line_bytes * rows;
        const uint8_t *lines = c.p;
        if (compression == 4 && size < raw_size) {
            // PIZ: [min, max non-zero bitmap byte][bitmap][huffman length][huffman data] -> u16 planes per channel -> wavelet -> LUT -> lines
            Bytes z{ c.p, c.p + size, path };
            const uint16_t mn = z.get<uint16_t>(), mx = z.get<uint16_t>();
            std::vector<uint8_t> bitmap(8192, 0);
            if (mx >= 8192) b.fail("PIZ: invalid bitmap range");
            if (mn <= mx) { z.need((size_t) mx - mn + 1); std::memcpy(bitmap.data() + mn, z.p, (size_t) mx - mn + 1); z.p += (size_t) mx - mn + 1; }
            std::vector<uint16_t> lut(65536, 0); uint32_t k = 0;
            for (uint32_t i = 0; i < 65536; ++i) if (i == 0 || (bitmap[i >> 3] & (1u << (i & 7u)))) lut[k++] = (uint16_t) i;
            const uint16_t max_value = (uint16_t) (k - 1);
            const uint32_t hlen = z.get<uint32_t>(); z.need(hlen);
            const size_t n16 = raw_size / 2; tmp.assign(n16, 0);
            huf_uncompress(z.p, hlen, tmp.data(), n16, b);
            size_t start = 0; std::vector<size_t> ch_start(ch.size());
            for (size_t i = 0; i < ch.size(); ++i) {
                const int sz = ch[i].type == 1 ? 1 : 2; ch_start[i] = start;
                for (int j = 0; j < sz; ++j) wav2_decode(tmp.data() + start + j, (int) W, sz, (int) rows, (int) W * sz, max_value);
                start += (size_t) W * rows * sz;
            }
            for (auto &v : tmp) v = lut[v];
            raw.resize(raw_size); uint8_t *o = raw.data();
            for (uint32_t r = 0; r < rows; ++r) for (size_t i = 0; i < ch.size(); ++i) {
                const size_t n = (size_t) W * (ch[i].type == 1 ? 1 : 2);
                std::memcpy(o, tmp.data() + ch_start[i] + (size_t) r * n, n * 2); o += n * 2;
            }
            lines = raw.data();
        } else if (size != raw_size) b.fail("chunk size does not match the scan-line size");
        for (uint32_t r = 0; r < rows; ++r) for (int k = 0; k < 3; ++k) {
            const Channel &cc = ch[(size_t) idx[k]]; const uint8_t *src = lines + (size_t) r * line_bytes + ch_off[(size_t) idx[k]];
            float *dst = rgb.data() + ((size_t) (row0 + r) * W) * 3 + k;
            for (uint32_t x = 0; x < W; ++x) {
                float v;
                if (cc.type == 1) { uint16_t h; std::memcpy(&h, src + 2 * (size_t) x, 2); v = half_to_float(h); }
                else if (cc.type == 2) std::memcpy(&v, src + 4 * (size_t) x, 4);
                else { uint32_t u; std::memcpy(&u, src + 4 * (size_t) x, 4); v = (float) u; }
                dst[3 * (size_t) x] = v;
            }
        }
    }
    *w_out = W; *h_out = H;
}

void read_pfm(const std::string &path, const std::vector<uint8_t> &file, uint32_t *w_out, uint32_t *h_out, std::vector<float> &rgb) {
    std::string text((const char *) file.data(), std::min<size_t>(file.size(), 128)); std::istringstream hs(text);
    std::string magic; long w = 0, h = 0; double scale = 0; hs >> magic >> w >> h >> scale;
    if (!hs || (magic != "PF" && magic != "Pf") || w <= 0 || h <= 0 || scale == 0) AM_THROW("io: \"%s\": malformed PFM header", path.c_str());
    size_t pos = (size_t) hs.tellg() + 1;        // one whitespace byte after the scale
    const int nc = magic == "PF" ? 3 : 1; const size_t need = (size_t) w * h * nc * 4;
    if (pos + need > file.size()) AM_THROW("io: \"%s\": PFM pixel data is truncated", path.c_str());
    const bool big = scale > 0;
    rgb.assign((size_t) w * h * 3, 0.f);
    for (long y = 0; y < h; ++y) for (long x = 0; x < w; ++x) for (int c = 0; c < 3; ++c) {
        const uint8_t *s = file.data() + pos + (((size_t) (h - 1 - y) * w + x) * nc + (nc == 3 ? c : 0)) * 4;
        uint8_t q[4] = { s[0], s[1], s[2], s[3] }; if (big) { std::swap(q[0], q[3]); std::swap(q[1], q[2]); }
        float v; std::memcpy(&v, q, 4); rgb[((size_t) y * w + x) * 3 + c] = v;
    }
    *w_out = (uint32_t) w; *h_out = (uint32_t) h;
}

} // namespace

void load_image_rgb(const std::string &path, uint32_t *w, uint32_t *h, std::vector<float> &rgb) {
    std::ifstream f(path, std::ios::binary); if (!f) AM_THROW("io: \"%s\": file does not exist!", path.c_str());
    std::vector<uint8_t> file((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    if (file.size() >= 4 && file[0] == 0x76 && file[1] == 0x2f && file[2] == 0x31 && file[3] == 0x01) { read_exr(path, file, w, h, rgb); return; }
    if (file.size() >= 2 && file[0] == 'P' && (file[1] == 'F' || file[1] == 'f')) { read_pfm(path, file, w, h, rgb); return; }
    AM_THROW("io: \"%s\": unsupported image format (this loader reads OpenEXR scan-line files and PFM)", path.c_str());
}

} // namespace amvpt
