// dpath.cuh -- per-lane device code of the AMVPT path loop shared by the schedulers (kernels.cu: persistent megakernels,
// wavefront.cu: staged wavefront): path state, lane -> (pixel, sampler, camera ray), ticket map, the vertex work of
// MVPT::sample_single with deferred next-event visibility, the traversal state machine, the aggregated film splat.
#pragma once
#include "kernels.cuh"
#include "dscene.cuh"

namespace amvpt {

// ------------------------------------------------------------------------------------------------
struct PathState {
    Ray ray; V3 throughput, result, prev_p; float eta, prev_bsdf_pdf; uint32_t depth;
    bool prev_delta, valid_ray, active;
    Pcg32 rng;
};
struct TapCtx { uint32_t max_hits, n_iter; uint32_t *hit_prim; float *hit_t; };
struct StatCtx { uint32_t closest, shadow, nodes, tris, vertices; };

AM_D void path_init(const DScene &s, PathState &ps, const Ray &ray) {
    ps.ray = ray; ps.throughput = v3(1.f); ps.result = v3(0.f); ps.prev_p = v3(0.f); ps.eta = 1.f; ps.prev_bsdf_pdf = 1.f;
    ps.depth = 0; ps.prev_delta = true;
    ps.valid_ray = !s.hide_emitters && s.env_emitter >= 0;      // mvpath_single.h:98
    ps.active = true;
}
// Direct emission of one loop iteration (mvpath_single.h:148-163): the emitter of the intersected shape, or -- GEN only -- the
// environment emitter when the ray left the scene (SurfaceInteraction::emitter, scene.h:647-657; DirectionSample(scene, si, prev_si),
// records.h:172-179: d = -si.wi for an invalid interaction)
template <bool GEN> AM_D void add_emission(const DScene &s, PathState &ps, const SI &si, const DMesh *mesh) {
    if (si.valid && mesh->emitter >= 0) {
        V3 rel = si.p - ps.prev_p; float dist = norm(rel); V3 d = rel / dist;
        float em_pdf = 0.f;
        if (!ps.prev_delta) em_pdf = pdf_emitter_direction(s, mesh, d, si.sh_n, dist);
        float mis_bsdf = mis_weight(ps.prev_bsdf_pdf, em_pdf);
        V3 le = select3(si.wi.z > 0.f && ps.prev_bsdf_pdf > 0.f, ld3(s.emitters[mesh->emitter].radiance));   // area.cpp:82-86
        ps.result = fmadd(ps.throughput, le * mis_bsdf, ps.result);
    } else if (GEN && !si.valid && s.env_emitter >= 0) {
        const V3 d = ps.ray.d;
        float em_pdf = 0.f;
        if (!ps.prev_delta) em_pdf = env_pdf_direction(s, d) * rcp((float) s.n_emitters);      // scene.cpp:351-360
        float mis_bsdf = mis_weight(ps.prev_bsdf_pdf, em_pdf);
        V3 le = select3(ps.prev_bsdf_pdf > 0.f, env_eval(s, d));
        ps.result = fmadd(ps.throughput, le * mis_bsdf, ps.result);
    }
}

// Lane -> (pixel, sampler, camera ray): MVPT::render lane map (src/integrators/mvpath.cpp:174-190) and
// MVPT::render_sample (src/integrators/mvpath_single.h:50-75).
struct LaneSetup { float spx, spy; uint32_t view, pixel; bool in_film; };
AM_D void lane_setup(const DScene &s, uint32_t seed_value, uint32_t lane, uint32_t px, uint32_t py, PathState &ps, LaneSetup &ls, V2 *aperture_out = nullptr) {
    ps.rng = sampler_seed(seed_value, lane);
    float sx = rcp((float) s.film_w), sy = rcp((float) s.film_h);
    float jx = ps.rng.next_f32(), jy = ps.rng.next_f32();
    ls.spx = (float) (int32_t) px + jx; ls.spy = (float) (int32_t) py + jy;
    V2 adjusted = v2(fmadd(ls.spx, sx, -0.f * sx), fmadd(ls.spy, sy, -0.f * sy));
    V2 aperture = v2(.5f, .5f);
    if (s.thinlens) { aperture.x = ps.rng.next_f32(); aperture.y = ps.rng.next_f32(); }
    if (s.shutter_draw) (void) ps.rng.next_f32();      // time = shutter_open + next_1d() * shutter_open_time (:66-68); the scene is static, the draw is what matters
    Ray ray = sensor_sample_ray(s, adjusted, aperture, &ls.view);
    if (aperture_out) *aperture_out = aperture;
    path_init(s, ps, ray);
    // ImageBlock::put, box filter (src/render/imageblock.cpp:210-233): pixel = floor(sample_pos)
    uint32_t ux = (uint32_t) (int32_t) floorf(ls.spx), uy = (uint32_t) (int32_t) floorf(ls.spy);
    ls.in_film = ux < s.film_w && uy < s.film_h;
    ls.pixel = uy * s.film_w + ux;
}

// Map a launch-local lane ticket to (global lane, pixel) through the tile list.
AM_D bool ticket_to_lane(const RenderParams &p, const DScene &s, unsigned long long ticket, uint32_t *lane, uint32_t *px, uint32_t *py) {
    if (ticket >= p.n_lanes) return false;
    uint32_t lo = 0, hi = p.n_tiles - 1;
    while (lo < hi) {                       // last tile whose lane_begin <= ticket
        uint32_t mid = (lo + hi + 1) >> 1;
        if (p.tiles[mid].lane_begin <= ticket) lo = mid; else hi = mid - 1;
    }
    const TileDesc &t = p.tiles[lo];
    unsigned long long r = ticket - t.lane_begin;
    uint32_t k = (uint32_t) (r % t.spp), pl = (uint32_t) (r / t.spp);
    uint32_t ty = pl / t.w, tx = pl - ty * t.w;
    *px = t.x0 + tx; *py = t.y0 + ty;
    *lane = (*py * s.film_w + *px) * p.spp_per_pass + k;
    return true;
}

// Film splat with warp-aggregated atomics: lanes of the (converged) group `mask` that target the same pixel are
// summed in lane order with shuffles and the lowest lane issues one 16-byte vector reduction (RED.128) per pixel.
AM_D void splat_aggregated(unsigned mask, bool do_put, uint32_t pixel, float4 val, float4 *film, float2 *moments) {
    unsigned lane = threadIdx.x & 31u;
    uint32_t key = do_put ? pixel : 0xffffffffu;
    unsigned peers = __match_any_sync(mask, key);
    int iters = __reduce_max_sync(mask, do_put ? __popc(peers) : 0);
    unsigned rem = peers;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f); float m2 = 0.f;
    float lum = 0.2126f * val.x + 0.7152f * val.y + 0.0722f * val.z, lum2 = lum * lum;
    for (int i = 0; i < iters; ++i) {
        int src = rem ? (__ffs(rem) - 1) : (int) lane;
        float vx = __shfl_sync(mask, val.x, src), vy = __shfl_sync(mask, val.y, src),
              vz = __shfl_sync(mask, val.z, src), vw = __shfl_sync(mask, val.w, src);
        float vm = moments ? __shfl_sync(mask, lum2, src) : 0.f;
        if (rem) { acc.x += vx; acc.y += vy; acc.z += vz; acc.w += vw; m2 += vm; }
        rem &= rem - 1;
    }
    if (do_put && lane == (unsigned) (__ffs(peers) - 1)) {
        atomicAdd(film + pixel, acc);
        if (moments) atomicAdd(moments + pixel, make_float2(0.2126f * acc.x + 0.7152f * acc.y + 0.0722f * acc.z, m2));
    }
}

// ImageBlock::put for a reconstruction filter other than the box: coalesced accumulation (src/render/imageblock.cpp:433-505)
// into a block without border, normalize = false.  The footprint is (2n + 1)^2 pixels, n = ceil(radius - 1/2), around
// floor(pos) - n; every covered pixel inside the film receives {R, G, B, 1} * w_y * w_x as one 16-byte reduction.
AM_D float filter_eval(const DScene &s, float x) {
    if (s.filter_type == 1) return max_(estrin10(x * x, s.filter_coeff), 0.f);        // gaussian.cpp:94-97
    if (s.filter_type == 2) return max_(0.f, 1.f - fabs_(x * s.filter_inv_radius));   // tent.cpp:54-56
    x = fabs_(x);
    if (s.filter_type == 5) {                                                          // lanczos.cpp:57-67
        float x1 = AM_PI * x, x2 = fdiv(x1, s.filter_radius), s1, s2, c;
        sincos_(x1, &s1, &c); sincos_(x2, &s2, &c);
        const float result = fdiv(s1 * s2, x1 * x2);
        return x < AM_EPSILON ? 1.f : (x > s.filter_radius ? 0.f : result);
    }
    const float x2 = x * x, x3 = x2 * x; const float *k = s.filter_coeff; float r;
    if (s.filter_type == 3)                                                            // mitchell.cpp:59-78, coefficients set at commit
        r = x < 1.f ? fmadd(k[0], x3, fmadd(k[1], x2, k[2])) : fmadd(k[3], x3, fmadd(k[4], x2, fmadd(k[5], x, k[6])));
    else                                                                               // catmullrom.cpp:39-53 with B = 0, C = 1/2 folded
        r = x < 1.f ? (9.f * x3 + -15.f * x2) + 6.f : ((-3.f * x3 + 15.f * x2) + -24.f * x) + 12.f;
    r = (1.f / 6.f) * r;
    return x < 2.f ? r : 0.f;
}
AM_D void splat_filtered(const DScene &s, float spx, float spy, V3 L, float4 *film) {
    const uint32_t n = (uint32_t) (int32_t) ceilf(s.filter_radius - .5f), count = 2 * n + 1;
    const int32_t pix = (int32_t) floorf(spx) - (int32_t) n, piy = (int32_t) floorf(spy) - (int32_t) n;
    float rx0 = ((float) pix + .5f) - spx, ry = ((float) piy + .5f) - spy;
    uint32_t y = (uint32_t) piy;
    for (uint32_t ys = 0; ys < count; ++ys, ++y, ry += 1.f) {
        if (!(y < s.film_h)) continue;
        const float wy = filter_eval(s, ry);
        float rx = rx0;
        for (uint32_t xs = 0; xs < count; ++xs, rx += 1.f) {
            const uint32_t x = (uint32_t) pix + xs; if (!(x < s.film_w)) continue;
            const float w = wy * filter_eval(s, rx);
            atomicAdd(film + (size_t) y * s.film_w + x, make_float4(L.x * w, L.y * w, L.z * w, 1.f * w));
        }
    }
}

AM_D void flush_stats(const StatCtx &st, unsigned long long *c) {
    // warp-reduce, one atomic per counter per warp
    uint32_t v[5] = { st.closest, st.shadow, st.nodes, st.tris, st.vertices };
    #pragma unroll
    for (int i = 0; i < 5; ++i) {
        uint32_t x = __reduce_add_sync(0xffffffffu, v[i]);
        if ((threadIdx.x & 31u) == 0 && x) atomicAdd(c + CTR_CLOSEST + i, (unsigned long long) x);
    }
}

struct TravState {
    Ray ray;                 // the ray in flight (maxt = original bound, used by the triangle test)
    SlabRay sr; float best;  // pruning bound (closest t so far)
    int32_t node; int sp;
    Hit h; uint32_t best_prim;
    bool any;                // shadow ray: first accepted triangle ends it
};
AM_D void trav_begin(TravState &tv, const Ray &ray, bool any) {
    tv.ray = ray; tv.sr = slab_ray(ray.o, ray.d); tv.best = ray.maxt; tv.node = 0; tv.sp = 0;
    tv.h.t = AM_INF; tv.h.u = 0.f; tv.h.v = 0.f; tv.h.tri = AM_NO_HIT; tv.best_prim = 0; tv.any = any;
}

// Vertex work for a finished closest-hit ray.  Returns true if a shadow ray has to be traced before the next closest ray
// (then *shadow holds it and sT / sV the parked operands of `result = fmadd(sT, sV, result)`).
template <bool STATS, int TYPE = -1, bool GEN = false> AM_D bool shade_vertex(const DScene &s, PathState &ps, const Hit &h, Ray *shadow, V3 *sT, V3 *sV, StatCtx *st) {
    if (STATS) { st->closest++; st->vertices++; }
    const DMesh *mesh;
    SI si = make_si(s, ps.ray, h, &mesh);
    add_emission<GEN>(s, ps, si, mesh);
    bool active_next = (ps.depth + 1 < s.max_depth) && si.valid;
    if (!active_next) { ps.valid_ray = ps.valid_ray || si.valid; ps.active = false; return false; }
    const DBsdf &bsdf = s.bsdfs[mesh->bsdf];
    bool active_em = (bsdf.flags & L_SMOOTH) != 0;
    V2 es; es.x = ps.rng.next_f32(); es.y = ps.rng.next_f32();
    DirectionSample ds; V3 em_weight = v3(0.f), wo = v3(0.f); ds.pdf = 0.f; ds.delta = false;
    if (active_em) {
        ds = sample_emitter_direction<true, GEN>(s, si, es, &em_weight, nullptr, nullptr, shadow);
        if (STATS) st->shadow++;
        active_em = ds.pdf != 0.f;
        wo = si.to_local(ds.d);
    }
    float s1 = ps.rng.next_f32(); V2 s2; s2.x = ps.rng.next_f32(); s2.y = ps.rng.next_f32();
    BsdfResult br = bsdf_eval_pdf_sample_t<TYPE>(bsdf, si.wi, wo, s1, s2);
    if (active_em) {
        float mis_em = (GEN && ds.delta) ? 1.f : mis_weight(ds.pdf, br.pdf);
        *sT = ps.throughput; *sV = br.f * em_weight * mis_em;
    }
    ps.ray = spawn_ray(si.p, si.n, si.to_world(br.bs.wo));
    ps.throughput = ps.throughput * br.weight;
    ps.eta *= br.bs.eta;
    ps.valid_ray = true;
    ps.prev_p = si.p; ps.prev_bsdf_pdf = br.bs.pdf; ps.prev_delta = (br.bs.type & L_DELTA) != 0;
    ps.depth += 1;
    float tmax = hmax(ps.throughput);
    float rr_prob = min_(tmax * sqr(ps.eta), .95f);
    bool rr_active = ps.depth >= s.rr_depth, rr_continue = ps.rng.next_f32() < rr_prob;
    if (rr_active) ps.throughput = ps.throughput * rcp(rr_prob);
    ps.active = (!rr_active || rr_continue) && (tmax != 0.f);
    return active_em;
}

// Optional out-of-line copies of the ray caster, the emitter sampler and the BSDF dispatch for kernels that call them from many
// places (reuse.cu: per camera, per pair of cameras, and in the suffix loop; -DAM_REUSE_OUTLINE=1).  Inlined everywhere, that kernel is 616 KB of SASS and its warps wait for instruction fetch (ncu: stall_no_instruction 8.9 per issue); these
// out-of-line copies keep one body each.  The scene record stays in the kernel parameter bank (__grid_constant__), so
// passing it by reference costs no local copy.
#ifndef AM_REUSE_OUTLINE
#define AM_REUSE_OUTLINE 0     /* measured on B200 (profiles/r01_reuse_variants.txt): inlined + 8 CTAs/SM 490 / 51 Mpaths/s (Cornell 8x8 / 1M triangles), out-of-line + 6 CTAs/SM 339 / 48 */
#endif
#if AM_REUSE_OUTLINE
#define AM_NOINLINE static __device__ __noinline__
#else
#define AM_NOINLINE static __device__ __forceinline__
#endif
AM_NOINLINE Hit trace_closest(const DScene &s, const Ray &ray, StatCtx *st) { return intersect<false>(s, ray, st ? &st->nodes : nullptr, st ? &st->tris : nullptr); }
AM_NOINLINE bool trace_occluded(const DScene &s, const Ray &ray, StatCtx *st) { return intersect<true>(s, ray, st ? &st->nodes : nullptr, st ? &st->tris : nullptr).tri != AM_NO_HIT; }
AM_NOINLINE BsdfResult bsdf_call(const DBsdf &b, V3 wi, V3 wo, float s1, V2 s2) { return bsdf_eval_pdf_sample(b, wi, wo, s1, s2); }
AM_NOINLINE DirectionSample emitter_call(const DScene &s, const SI &si, V2 u, V3 *w, StatCtx *st) {
    return sample_emitter_direction<false>(s, si, u, w, st ? &st->nodes : nullptr, st ? &st->tris : nullptr);
}

// One iteration of the while-loop of MVPT::sample_single (src/integrators/mvpath_single.h:130-275), wavefront
// (JIT) semantics: every lane active at the top of the iteration draws 2 + 1 + 2 + 1 random numbers.
// OUTLINE = true calls the out-of-line copies above instead of inlining ray caster, emitter sampler and BSDF dispatch.
// FIRST = true: the camera ray's hit and the first vertex's light sample (with its visibility) were computed by
// primary_vertex_kernel and are read from prim[i], prim[m + i], prim[2 m + i] instead of being traced here.
template <bool TAP, bool STATS, bool OUTLINE = false, bool FIRST = false, bool GEN = false, int ACCEL = 0> AM_D void path_bounce(const DScene &s, PathState &ps, TapCtx *tap, StatCtx *st,
                                                                                             const float4 *prim = nullptr, uint32_t pm = 0, uint32_t pi = 0) {
    uint32_t *cn = STATS ? &st->nodes : nullptr, *ct = STATS ? &st->tris : nullptr;
    Hit h;
    if (FIRST) { const float4 a = prim[pi]; h.t = a.x; h.u = a.y; h.v = a.z; h.tri = __float_as_uint(a.w); }
    else h = OUTLINE ? trace_closest(s, ps.ray, STATS ? st : nullptr) : intersect<false, false, ACCEL>(s, ps.ray, cn, ct);
    if (STATS) { st->closest++; st->vertices++; }
    const DMesh *mesh;
    SI si = make_si(s, ps.ray, h, &mesh);
    if (TAP) {
        if (tap->n_iter < tap->max_hits) { tap->hit_prim[tap->n_iter] = si.prim; tap->hit_t[tap->n_iter] = h.t; }
        tap->n_iter++;
    }
    // direct emission (:148-163); DirectionSample(scene, si, prev_si) records.h:141-149
    add_emission<GEN>(s, ps, si, mesh);
    bool active_next = (ps.depth + 1 < s.max_depth) && si.valid;
    if (!active_next) {
        // the lane finishes this iteration (its six draws and the BSDF call are unobservable) and leaves the loop
        ps.valid_ray = ps.valid_ray || si.valid;     // :252
        ps.active = false;
        return;
    }
    const DBsdf &bsdf = s.bsdfs[mesh->bsdf];
    // emitter sampling (:179-201)
    bool active_em = (bsdf.flags & L_SMOOTH) != 0;
    V2 es; es.x = ps.rng.next_f32(); es.y = ps.rng.next_f32();
    DirectionSample ds; V3 em_weight = v3(0.f), wo = v3(0.f); ds.pdf = 0.f; ds.delta = false;
    if (active_em) {
        if (FIRST) { const float4 a = prim[(size_t) pm + pi], b = prim[2 * (size_t) pm + pi]; em_weight = v3(a.x, a.y, a.z); ds.pdf = a.w; ds.d = v3(b.x, b.y, b.z); }
        else ds = OUTLINE ? emitter_call(s, si, es, &em_weight, STATS ? st : nullptr) : sample_emitter_direction<false, GEN, ACCEL>(s, si, es, &em_weight, cn, ct);
        if (STATS) st->shadow++;
        active_em = ds.pdf != 0.f;
        wo = si.to_local(ds.d);
    }
    float s1 = ps.rng.next_f32(); V2 s2; s2.x = ps.rng.next_f32(); s2.y = ps.rng.next_f32();
    BsdfResult br = OUTLINE ? bsdf_call(bsdf, si.wi, wo, s1, s2) : bsdf_eval_pdf_sample(bsdf, si.wi, wo, s1, s2);
    if (active_em) {
        float mis_em = (GEN && ds.delta) ? 1.f : mis_weight(ds.pdf, br.pdf);
        ps.result = fmadd(ps.throughput, br.f * em_weight * mis_em, ps.result);
    }
    ps.ray = spawn_ray(si.p, si.n, si.to_world(br.bs.wo));
    ps.throughput = ps.throughput * br.weight;
    ps.eta *= br.bs.eta;
    ps.valid_ray = true;
    ps.prev_p = si.p; ps.prev_bsdf_pdf = br.bs.pdf; ps.prev_delta = (br.bs.type & L_DELTA) != 0;
    ps.depth += 1;
    // Russian roulette (:264-274)
    float tmax = hmax(ps.throughput);
    float rr_prob = min_(tmax * sqr(ps.eta), .95f);
    bool rr_active = ps.depth >= s.rr_depth, rr_continue = ps.rng.next_f32() < rr_prob;
    if (rr_active) ps.throughput = ps.throughput * rcp(rr_prob);
    ps.active = (!rr_active || rr_continue) && (tmax != 0.f);
}

} // namespace amvpt
