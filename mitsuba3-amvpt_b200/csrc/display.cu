// Display step of the light-field pipeline: the developed quilt becomes an 8-bit sRGB bitmap, which the
// viewer interlaces for a slanted lenticular display.  The reference does both on the host (Bitmap::convert in
// src/mitsuba/program.h:55, Program::display_image in src/mitsuba/program.cpp:199-272, one std::thread per row block) and
// uploads the result to an SDL surface; here both are kernels writing a headless device buffer (no window: unavailable on this box).
// Parity: bit-exact against oracle/o_display.h, whose leaves are pinned to the reference's own headers.
#include "dmath.cuh"
#include "kernels.cuh"

namespace amvpt {

// dr::linear_to_srgb, single precision (ext/drjit/include/drjit/color.h:19-53); Estrin pairs are the fmadd sites of drjit/math.h:26-41
AM_D float estrin6(float x, float c0, float c1, float c2, float c3, float c4, float c5) {
    const float a0 = __fmaf_rn(x, c1, c0), a1 = __fmaf_rn(x, c3, c2), a2 = __fmaf_rn(x, c5, c4);
    const float x2 = x * x, b0 = __fmaf_rn(x2, a1, a0), x4 = x2 * x2;
    return __fmaf_rn(x4, a2, b0);
}
AM_D float linear_to_srgb(float x) {
    float r = 12.92f;
    if (x > 0.0031308f) {
        const float y = fsqrt(x);
        r = fdiv(estrin6(y, -0.0016829072605308378f, 0.03453868659826638f, 0.7642611304733891f, 2.0041169284241644f, 0.7551545191665577f, -0.016202083165206348f),
                 estrin6(y, 4.178892964897981e-7f, -0.00004375359692957097f, 0.03467195408529984f, 0.6085338522168684f, 1.8970238036421054f, 1.f));
    }
    return r * x;
}
// StructConverter::save for a normalized UInt8 field with the gamma flag (src/core/struct.cpp:1645-1662): scale by 255, clamp,
// round to nearest even.  The reference clamps and rounds in double; v * 255 is a float, so rounding it in float is the same number.
AM_D uint32_t to_srgb8(float v) { return (uint32_t) __float2int_rn(fminf(fmaxf(linear_to_srgb(v) * 255.f, 0.f), 255.f)); }

// HDRFilm::develop + Bitmap::convert(RGB, UInt8, srgb).  A CTA converts tiles of 1024 pixels: four coalesced 16-byte loads per thread
// (all issued before the arithmetic), the 3 bytes of each pixel staged in shared memory, the tile's 3072 bytes stored as three
// coalesced 32-bit words per thread.  Streaming: 16 B read + 3 B written per pixel.
__global__ void __launch_bounds__(256) srgb8_kernel(const float4 *__restrict__ film, uint8_t *__restrict__ out, size_t npix) {
    __shared__ uint32_t sm[768];
    uint8_t *smb = reinterpret_cast<uint8_t *>(sm);
    const size_t n_tiles = (npix + 1023) / 1024; const uint32_t tid = threadIdx.x;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t base = tile * 1024; const uint32_t rem = (uint32_t) (npix - base < 1024 ? npix - base : 1024);
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { const uint32_t i = k * 256 + tid; v[k] = i < rem ? __ldcs(film + base + i) : make_float4(0.f, 0.f, 0.f, 0.f); }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float w = (v[k].w == 0.f) ? 1.f : v[k].w; uint8_t *o = smb + 3 * (k * 256 + tid);
            o[0] = (uint8_t) to_srgb8(fdiv(v[k].x, w)); o[1] = (uint8_t) to_srgb8(fdiv(v[k].y, w)); o[2] = (uint8_t) to_srgb8(fdiv(v[k].z, w));
        }
        __syncthreads();
        if (rem == 1024) {
            uint32_t *o = reinterpret_cast<uint32_t *>(out) + tile * 768;      // 3 * base bytes: a multiple of four
#pragma unroll
            for (int j = 0; j < 3; ++j) o[j * 256 + tid] = sm[j * 256 + tid];
        } else for (uint32_t j = tid; j < 3 * rem; j += 256) out[3 * base + j] = smb[j];    // ragged last tile
        __syncthreads();
    }
}
void launch_srgb8(const float4 *film, uint8_t *out, size_t npix, int sm_count, cudaStream_t st) {
    if (!npix) return;
    const size_t n_tiles = (npix + 1023) / 1024; const unsigned cap = (unsigned) sm_count * 8;
    srgb8_kernel<<<n_tiles < cap ? (unsigned) n_tiles : cap, 256, 0, st>>>(film, out, npix);
}

// src/mitsuba/util.h:47-49, 104-128
AM_D float clip01(float x) { return fminf(fmaxf(x, 0.f), 1.f); }
AM_D uint32_t lerp_u8(uint32_t a, uint32_t b, float t) { return (uint32_t) ((1.f - t) * (float) a + t * (float) b) & 255u; }   // lerp2<uint8_t, float>: truncated by its return type
AM_D uint32_t interpolate2d(const uint8_t *__restrict__ d, float x, float y, uint32_t w, uint32_t h, uint32_t c, uint32_t nch, bool nearest) {
    if (nearest) { const uint32_t x1 = (uint32_t) (x + 0.5f), y1 = (uint32_t) (y + 0.5f); return __ldg(d + ((size_t) x1 + (size_t) y1 * w) * nch + c); }
    const uint32_t x1 = (uint32_t) x, y1 = (uint32_t) y; const float tx = x - (float) x1, ty = y - (float) y1;
    const uint32_t x2 = x1 + 1 < w ? x1 + 1 : x1, y2 = y1 + 1 < h ? y1 + 1 : y1;
    const uint32_t a00 = __ldg(d + ((size_t) x1 + (size_t) y1 * w) * nch + c), a10 = __ldg(d + ((size_t) x2 + (size_t) y1 * w) * nch + c);
    const uint32_t a01 = __ldg(d + ((size_t) x1 + (size_t) y2 * w) * nch + c), a11 = __ldg(d + ((size_t) x2 + (size_t) y2 * w) * nch + c);
    return lerp_u8(lerp_u8(a00, a10, tx), lerp_u8(a01, a11, tx), ty);
}

// Program::display_image (src/mitsuba/program.cpp:199-272): one thread per destination pixel, all its sub-pixels.  Every colour
// channel of a pixel lies under a different lens position (c * subp), so it reads a different view of the quilt: 4 byte gathers per
// sub-pixel from the 8-bit quilt (50 MB for the 4096^2 quilt: L2-resident), dst_ch bytes written per pixel.
template <int DCH> __global__ void __launch_bounds__(256)
interlace_kernel(const uint8_t *__restrict__ src, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const DisplayPreset p,
                 uint8_t *__restrict__ dst, uint32_t dst_w, uint32_t dst_h) {
    const uint32_t px = blockIdx.x * blockDim.x + threadIdx.x, dy = blockIdx.y;
    if (px >= dst_w) return;
    const uint32_t row = dst_w * DCH, dx = px * DCH;        // the reference's dst_dim[0] is the row length in bytes, dx a byte offset
    const bool flip_rgb = p.flip_rgb && DCH > 2;
    const float grid_n = (float) (p.grid[0] * p.grid[1]), grid_scl = fdiv(1.f, grid_n);
    const float ig0 = fdiv(1.f, (float) p.grid[0]), ig1 = fdiv(1.f, (float) p.grid[1]);
    const float x0 = (float) dx * fdiv(1.f, (float) row - 1.f), y0 = (float) dy * fdiv(1.f, (float) dst_h - 1.f);
    const float sx = (float) src_w - 1.f, sy = (float) src_h - 1.f;
    uint32_t out[4] = { 0, 0, 0, 255 };
#pragma unroll
    for (uint32_t c = 0; c < (DCH < 3 ? DCH : 3); ++c) {
        float x = x0, y = y0;
        if (!p.quilt) {
            float z = p.pitch * (x + (float) c * p.subp + (1.f - y) * p.tilt) - p.center;
            z = 1.f - (z - floorf(z));
            const float z1 = floorf(z * grid_n), z2 = floorf((1.f - z) * grid_n);
            const float w = p.focus * (1.f - 2.f * clip01(z1 * grid_scl));
            const float g0 = (float) p.grid[0];
            x = (z1 - g0 * floorf(fdiv(z1, g0))) + clip01(x + w);
            y = floorf(z2 * ig0) + y;
            x = clip01(p.view[0] * ig0 * x); y = clip01(p.view[1] * ig1 * y);
        }
        x = sx * x; y = sy * y;
        out[flip_rgb ? c : 2 - c] = interpolate2d(src, x, y, src_w, src_h, c, src_ch, p.nearest != 0);
    }
    uint8_t *o = dst + (size_t) dy * row + dx;
    if (DCH == 4) *reinterpret_cast<uint32_t *>(o) = out[0] | (out[1] << 8) | (out[2] << 16) | (out[3] << 24);
    else { o[0] = (uint8_t) out[0]; o[1] = (uint8_t) out[1]; o[2] = (uint8_t) out[2]; }
}
void launch_interlace(const uint8_t *src, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const DisplayPreset &p,
                      uint8_t *dst, uint32_t dst_w, uint32_t dst_h, uint32_t dst_ch, cudaStream_t st) {
    if (!dst_w || !dst_h) return;
    dim3 grid((dst_w + 255) / 256, dst_h);
    if (dst_ch == 4) interlace_kernel<4><<<grid, 256, 0, st>>>(src, src_w, src_h, src_ch, p, dst, dst_w, dst_h);
    else interlace_kernel<3><<<grid, 256, 0, st>>>(src, src_w, src_h, src_ch, p, dst, dst_w, dst_h);
}

} // namespace amvpt
