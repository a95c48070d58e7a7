// kernels.cuh -- launch parameter blocks shared by kernels.cu (device) and render.cpp / capi.cpp (host).
#pragma once
#include "dtypes.h"

namespace amvpt {

// One rectangle of the quilt traced this pass with `spp` samples per pixel (<= lane stride spp_per_pass).
// lane_begin = prefix sum of w*h*spp over the preceding tiles of the launch.
struct TileDesc { uint32_t x0, y0, w, h; uint32_t spp, pad; unsigned long long lane_begin; };

struct RenderParams {
    uint32_t seed_value;       // sampler seed + spp_per_pass * pass + render seed (mvpath.cpp:226-231, sampler.cpp:129)
    uint32_t spp_per_pass;     // lane stride of the reference's lane -> pixel map (mvpath.cpp:174-190)
    uint32_t n_tiles, regen_min;   // regen_min: idle lanes per warp that trigger a refill (schedulers 1, 2)
    uint32_t trace_min, bvh_depth; // scheduler 2: the traversal loop is left when fewer lanes than this are traversing; tree depth (stack size)
    const TileDesc *tiles;
    unsigned long long n_lanes;
    float4 *film;              // RGBW, film_w * film_h
    float2 *moments;           // optional (sum lum, sum lum^2) per pixel for the adaptive allocator, or nullptr
    // scheduler 6 ("split"): the launch covers tickets [ticket_base, ticket_base + prim_m); prim holds 3 * prim_m float4 written by
    // primary_vertex_kernel: closest hit of the camera ray, (light weight, pdf), light direction of the first vertex
    const float4 *prim; uint32_t prim_m, pad2; unsigned long long ticket_base;
    uint8_t *adapt_mask;           // reuse branch with `adaptive` > 0: adapt_mask of every ticket of the pass (mvpath_multi.h:306-307), else nullptr
    unsigned long long *counters;   // [0] next-lane ticket, [1] closest rays, [2] shadow rays, [3] nodes, [4] tris, [5] path vertices
};

// Per-lane debug taps of the parity tests (mirrors oracle/amvpt_oracle.cpp orc_trace_lanes)
struct LaneTaps {
    float *sample_pos, *ray_o, *ray_d, *ray_maxt; uint32_t *view; float *radiance; uint32_t *pixel;
    uint32_t max_hits; uint32_t *hit_prim; float *hit_t;
};

enum Counter { CTR_TICKET = 0, CTR_CLOSEST = 1, CTR_SHADOW = 2, CTR_NODES = 3, CTR_TRIS = 4, CTR_VERTICES = 5, CTR_COUNT = 8 };

// Path-slot pool of the wavefront scheduler (wavefront.cu): structure of arrays, n_slots entries each.
// Path-state pool of the wavefront scheduler.  The 16-byte fields of one slot are grouped into three 64-byte records so that the shade stages,
// which visit slots in queue order (scattered: one queue per BSDF plugin type), touch whole 32-byte sectors instead of half-used ones:
//   record A: ro, rd, tp, pp   (the path's next ray and carry state: read and rewritten by every shade)
//   record B: res, rng, hit, - (radiance so far + pixel, sampler, last closest hit)
//   record C: so, sd, sT, sV   (pending next-event shadow ray and its parked contribution)
// The trace and regen stages walk slots in order, so their accesses stay coalesced either way.  -DAM_WF_AOS=0 restores one array per field.
#ifndef AM_WF_AOS
#define AM_WF_AOS 1
#endif
template <class T, int OFF> struct PoolField {
    T *base;
#ifdef __CUDACC__
    __device__ __forceinline__ T &operator[](size_t i) const { return AM_WF_AOS ? base[4 * i + OFF] : base[i]; }
#endif
};
struct WfPool {
    PoolField<float4, 0> ro; PoolField<float4, 1> rd;        // next closest-hit ray: (o, maxt), (d, eta)
    PoolField<float4, 2> tp; PoolField<float4, 0> res; PoolField<float4, 3> pp;  // (throughput, prev_bsdf_pdf), (result, bits pixel), (prev_p, bits depth)
    PoolField<uint4, 1> rng;             // PCG32 state, inc
    PoolField<float4, 2> hit;            // (t, u, v, bits tri) of the last closest-hit ray
    PoolField<float4, 0> so; PoolField<float4, 1> sd; PoolField<float4, 2> sT; PoolField<float4, 3> sV;   // pending next-event shadow ray and the parked operands of result = fmadd(sT, sV, result)
    uint32_t *flags;        // WF_* bits
    float2 *spos;           // sample position on the film (read by the splat when the reconstruction filter is not the box)
    uint32_t *queue;        // [6][n_slots] slot indices per BSDF plugin type
    uint32_t *qcount;       // [6]
    uint32_t *live;         // [64] live-slot counters, one per iteration (ring)
    uint32_t n_slots;
    // ray-task sort (wf_key / wf_scan / wf_scatter kernels): tasks of one iteration bucketed by origin cell (Morton, 4 bits per axis) and
    // direction octant; order[] lists the existing tasks bucket by bucket, sort_total[0] = their number
    uint32_t *sort_hist;    // [WF_SORT_BUCKETS + 1] counts, then exclusive offsets
    uint32_t *sort_key;     // [2 * n_slots] bucket of task t, or WF_SORT_NONE
    uint32_t *sort_rank;    // [2 * n_slots] arrival rank of task t inside its bucket
    uint32_t *order;        // [2 * n_slots]
};
#define WF_SORT_BUCKETS 32768u
#define WF_SORT_NONE 0xffffffffu
uint32_t run_wavefront_pass(const DScene &s, const RenderParams &p, const WfPool &w, uint32_t type_mask, int sm_count, bool stats, cudaStream_t st,
                            double *trace_seconds, uint32_t *trace_launches);

// sample-reuse branch (reuse.cu)
void launch_reuse_pass(const DScene &s, const RenderParams &p, int sm_count, bool stats, cudaStream_t st);
size_t reuse_staged_float4s(uint32_t n_reuse, uint32_t m_cap);    // size of `buf` below
uint32_t run_reuse_staged(const DScene &s, const RenderParams &p, float4 *buf, uint32_t m_cap, int schedule /* 1 staged, 2 pipeline */, int sm_count, cudaStream_t st);
// fill-in pass of the reuse branch (mvpath_multi.h:79-115) over the pass described by p (adapt_mask filled by the kernels above); scratch = at least
// n_lanes / 1024 + 2 uint32; returns the number of launches and the number of fill-in paths traced
uint32_t run_reuse_fill_in(const DScene &s, const RenderParams &p, uint32_t base_seed, uint32_t *scratch, uint32_t **compact, size_t *compact_cap, int sm_count, cudaStream_t st, unsigned long long *n_fill);
void launch_reuse_tap(const DScene &s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, float *out /* n * n_reuse * 9 */, cudaStream_t st);

// host-callable launchers (kernels.cu)
uint32_t run_split_pass(const DScene &s, RenderParams p, float4 *prim /* 3 * m_cap */, uint32_t m_cap, int sm_count, cudaStream_t st);
void launch_render_pass(const DScene &s, const RenderParams &p, int mode, int sm_count, bool stats, cudaStream_t st);
void launch_trace_lanes(const DScene &s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, int trace,
                        const LaneTaps &taps, unsigned long long *counters, cudaStream_t st);
void launch_develop(const float4 *film, float *rgb, size_t npix, cudaStream_t st);
// display step (display.cu): same layout as amvpt_preset of include/amvpt.h (src/mitsuba/preset.h:12-17 + the viewer's two toggles)
struct DisplayPreset { float center, focus, pitch, tilt, subp, view[2]; int32_t grid[2]; int32_t flip_rgb, quilt, nearest; };
void launch_srgb8(const float4 *film, uint8_t *rgb8, size_t npix, int sm_count, cudaStream_t st);
void launch_interlace(const uint8_t *src, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const DisplayPreset &p,
                      uint8_t *dst, uint32_t dst_w, uint32_t dst_h, uint32_t dst_ch, cudaStream_t st);
void launch_view_stats(const float4 *film, const float2 *moments, uint32_t film_w, const TileDesc *tiles, uint32_t n_tiles,
                       float *stats /* n_tiles * 4 */, cudaStream_t st);
void launch_intersect(const DScene &s, const float *o, const float *d, const float *maxt, uint32_t n, int any,
                      uint32_t *prim, float *tuv, unsigned long long *counters, cudaStream_t st);
void launch_bsdf_tap(const DScene &s, int bsdf, const float *wi, const float *wo, const float *s1, const float *s2, uint32_t n,
                     float *out /* n * 16 */, cudaStream_t st);
void launch_math_tap(int op, const float *in, uint32_t n, float *out, cudaStream_t st);
void launch_env_warp_tap(const DScene &s, const float *u, uint32_t n, float *out /* n * 4: uv, pdf, eval(uv) */, cudaStream_t st);
void launch_sampler_tap(uint32_t seed_value, uint32_t lane, uint32_t n, float *out, cudaStream_t st);
void launch_l2_flush(float *buf, size_t n, cudaStream_t st);
// film gather: tiles (lane_begin = pixel offset of the tile in `stage`) -> contiguous staging buffer, or staging buffer added into the film
void launch_film_pack(float4 *film, uint32_t film_w, const TileDesc *tiles, uint32_t n_tiles, float4 *stage, size_t n_px, bool unpack_add, cudaStream_t st);
double run_membench_read(size_t bytes, uint32_t iters, int sm_count, cudaStream_t st);   // GB/s of 16-byte loads over `bytes`, or < 0 on failure

} // namespace amvpt
