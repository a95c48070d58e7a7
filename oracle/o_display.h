// oracle/o_display.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// CPU restatement of the reference's display step: developed quilt -> 8-bit sRGB bitmap ->
// lenticular interlace for a slanted-lens light-field display.  Plain scalar C++, canonical arithmetic (no FMA
// contraction; fmaf only where the reference writes fmadd).
#pragma once
#include <cmath>
#include <cstdint>
#include <algorithm>

namespace odisp {

// dr::linear_to_srgb, single-precision branch (ext/drjit/include/drjit/color.h:19-53): rational fit p(sqrt x) / q(sqrt x)
// evaluated with Estrin's scheme (ext/drjit/include/drjit/math.h:26-41: pairs fused with fmadd, then recursion on x^2).
inline float estrin6(float x, const float c[6]) {
    float a0 = std::fmaf(x, c[1], c[0]), a1 = std::fmaf(x, c[3], c[2]), a2 = std::fmaf(x, c[5], c[4]);
    float x2 = x * x;
    float b0 = std::fmaf(x2, a1, a0), b1 = a2;          // three coefficients left: one fmadd + the odd one
    float x4 = x2 * x2;
    return std::fmaf(x4, b1, b0);
}
inline float linear_to_srgb(float x) {
    float r = 12.92f;
    if (x > 0.0031308f) {
        const float P[6] = { -0.0016829072605308378f, 0.03453868659826638f, 0.7642611304733891f, 2.0041169284241644f, 0.7551545191665577f, -0.016202083165206348f };
        const float Q[6] = { 4.178892964897981e-7f, -0.00004375359692957097f, 0.03467195408529984f, 0.6085338522168684f, 1.8970238036421054f, 1.f };
        float y = std::sqrt(x);
        r = estrin6(y, P) / estrin6(y, Q);
    }
    return r * x;
}

// Bitmap::convert(PixelFormat::RGB, Struct::Type::UInt8, srgb_gamma = true) as called by the viewer (src/mitsuba/program.h:55);
// per component: StructConverter::save (src/core/struct.cpp:1645-1662) with Float = float (include/mitsuba/core/struct.h:369),
// no dithering: gamma, scale by the integer range, clamp in double, round to nearest even.
inline uint8_t to_srgb8(float v) {
    float f = linear_to_srgb(v);
    f *= 255.f;
    double d = (double) f;
    d = std::max(d, 0.0); d = std::min(d, 255.0);     // std::max(NaN, 0) keeps the first argument, as in the reference
    d = std::rint(d);
    return (uint8_t) (int64_t) d;
}

// src/mitsuba/util.h:47-49
inline float clip(float x) { return std::min(std::max(x, 0.f), 1.f); }
inline float mod(float x, float y) { return x - y * std::floor(x / y); }
// lerp2<T = uint8_t, Tp = float>: the float result is converted back to uint8_t by the return type (truncation)
inline uint8_t lerp2(uint8_t a, uint8_t b, float t) { return (uint8_t) ((1.f - t) * (float) a + t * (float) b); }

// interpolate2d<uint8_t> (src/mitsuba/util.h:104-128)
inline uint8_t interpolate2d(const uint8_t *data, float x, float y, uint32_t w, uint32_t h, uint32_t c, uint32_t nch, bool nearest) {
    if (nearest) {
        uint32_t x1 = (uint32_t) (x + 0.5f), y1 = (uint32_t) (y + 0.5f);
        return data[((size_t) x1 + (size_t) y1 * w) * nch + c];
    }
    uint32_t x1 = (uint32_t) x, y1 = (uint32_t) y;
    float tx = x - (float) x1, ty = y - (float) y1;
    uint32_t x2 = x1 + 1 < w ? x1 + 1 : x1, y2 = y1 + 1 < h ? y1 + 1 : y1;
    uint8_t a00 = data[((size_t) x1 + (size_t) y1 * w) * nch + c], a10 = data[((size_t) x2 + (size_t) y1 * w) * nch + c];
    uint8_t a01 = data[((size_t) x1 + (size_t) y2 * w) * nch + c], a11 = data[((size_t) x2 + (size_t) y2 * w) * nch + c];
    return lerp2(lerp2(a00, a10, tx), lerp2(a01, a11, tx), ty);
}

// src/mitsuba/preset.h:12-17 plus the two viewer toggles (program.cpp:130-132)
struct Preset { float center, focus, pitch, tilt, subp, view[2]; int32_t grid[2]; int32_t flip_rgb, quilt, nearest; };

// Program::display_image (src/mitsuba/program.cpp:199-272).  dst rows are dst_w * dst_ch bytes: the reference's "dst_dim[0]" is
// the row length in BYTES (screenshot: width * 3, :213-215), and x0 = dx * 1 / (row_bytes - 1) with dx the byte offset.
inline void interlace(const uint8_t *src, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const Preset &p,
                      uint8_t *dst, uint32_t dst_w, uint32_t dst_h, uint32_t dst_ch) {
    const uint32_t row = dst_w * dst_ch;
    const bool flip_rgb = p.flip_rgb && dst_ch > 2;
    const float grid_n = (float) (p.grid[0] * p.grid[1]), grid_scl = 1.f / grid_n;
    const float igrid[2] = { 1.f / (float) p.grid[0], 1.f / (float) p.grid[1] };
    const float dst_scl[2] = { 1.f / ((float) row - 1.f), 1.f / ((float) dst_h - 1.f) };
    const float src_scl[2] = { (float) src_w - 1.f, (float) src_h - 1.f };
    for (uint32_t dy = 0; dy < dst_h; dy++) {
        float y0 = (float) dy * dst_scl[1];
        for (uint32_t dx = 0; dx < row; dx += dst_ch) {
            float x0 = (float) dx * dst_scl[0];
            for (uint32_t c = 0; c < dst_ch; c++) {
                float xy[2] = { x0, y0 };
                if (c >= 3) { dst[(size_t) dy * row + dx + c] = 255; continue; }
                uint32_t ch = flip_rgb ? c : 2 - c;
                if (!p.quilt) {
                    float z = p.pitch * (xy[0] + (float) c * p.subp + (1.f - xy[1]) * p.tilt) - p.center;
                    z = 1.f - (z - std::floor(z));
                    float z1 = std::floor(z * grid_n);                 // camera id
                    float z2 = std::floor((1.f - z) * grid_n);         // inverse camera id
                    float w = p.focus * (1.f - 2.f * clip(z1 * grid_scl));
                    xy[0] = mod(z1, (float) p.grid[0]) + clip(xy[0] + w);
                    xy[1] = std::floor(z2 * igrid[0]) + xy[1];
                    xy[0] = clip(p.view[0] * igrid[0] * xy[0]); xy[1] = clip(p.view[1] * igrid[1] * xy[1]);
                }
                xy[0] = src_scl[0] * xy[0]; xy[1] = src_scl[1] * xy[1];
                dst[(size_t) dy * row + dx + ch] = interpolate2d(src, xy[0], xy[1], src_w, src_h, c, src_ch, p.nearest != 0);
            }
        }
    }
}

} // namespace odisp
