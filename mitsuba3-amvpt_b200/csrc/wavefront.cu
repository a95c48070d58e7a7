// wavefront.cu -- scheduler 3: the AMVPT path loop as a staged wavefront over a pool of path slots resident in HBM / L2.
//
// Why: in the persistent megakernels (kernels.cu) the vertex code of six BSDF plugins, the emitter sampling and the IEEE
// division / square-root expansions add up to ~175 KB of SASS, so warps that diverge across materials stall on instruction
// fetch (ncu: stall_no_instruction dominates on the 1M-triangle scene) and run with a handful of live lanes.  Here
//   regen  : one thread per slot -- finished paths are splatted (warp-aggregated RED.128), free slots draw lane tickets
//            (ballot + popc prefix, one atomic per warp) and generate camera rays
//   trace  : one thread per slot -- the pending next-event shadow ray (any hit) and then the closest-hit ray, nothing but
//            BVH traversal in the kernel; hits are appended to one queue per BSDF plugin type (match.any-aggregated atomics)
//   shade<T>: one thread per queue entry, specialised for plugin type T -- the vertex work of mvpath_single.h:143-275
// run in turn until every lane of the pass has been traced.  The arithmetic per lane is the code of dpath.cuh, shared with
// the megakernels, so the results are the same bit for bit; only the order of the film atomics differs.
#include "dpath.cuh"
#include "host.h"

namespace amvpt {

enum : uint32_t { WF_PREV_DELTA = 1u, WF_VALID_RAY = 2u, WF_ALIVE = 4u, WF_HAS_RESULT = 8u, WF_SHADOW = 16u };
#define WF_NO_PIXEL 0xffffffffu
#define WF_N_TYPES 6

AM_D float4 f4(V3 v, float w) { return make_float4(v.x, v.y, v.z, w); }
AM_D V3 xyz(float4 v) { return v3(v.x, v.y, v.z); }

template <bool STATS> __global__ void __launch_bounds__(128)
wf_regen_kernel(const DScene s, const RenderParams p, const WfPool w, uint32_t iter) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;       // pool size is a multiple of the CTA size
    const unsigned lane_id = threadIdx.x & 31u;
    uint32_t f = w.flags[i];
    bool dead = (f & (WF_ALIVE | WF_SHADOW)) == 0;
    unsigned dm = __ballot_sync(0xffffffffu, dead);
    if (dead) {
        bool put = (f & WF_HAS_RESULT) != 0; uint32_t pixel = WF_NO_PIXEL; V3 L = v3(0.f);
        if (put) { float4 r = w.res[i]; pixel = __float_as_uint(r.w); if (f & WF_VALID_RAY) L = xyz(r); }
        splat_aggregated(dm, put && pixel != WF_NO_PIXEL, pixel, make_float4(L.x, L.y, L.z, 1.f), p.film, p.moments);
        f = 0;
        unsigned long long base = 0; unsigned leader = __ffs(dm) - 1;
        if (lane_id == leader) base = atomicAdd(p.counters + CTR_TICKET, (unsigned long long) __popc(dm));
        base = __shfl_sync(dm, base, leader);
        unsigned long long ticket = base + __popc(dm & ((1u << lane_id) - 1u));
        uint32_t lane, px, py;
        if (ticket_to_lane(p, s, ticket, &lane, &px, &py)) {
            PathState ps; LaneSetup ls;
            lane_setup(s, p.seed_value, lane, px, py, ps, ls);
            w.ro[i] = f4(ps.ray.o, ps.ray.maxt); w.rd[i] = f4(ps.ray.d, ps.eta); w.tp[i] = f4(ps.throughput, ps.prev_bsdf_pdf);
            w.res[i] = f4(ps.result, __uint_as_float(ls.in_film ? ls.pixel : WF_NO_PIXEL)); w.pp[i] = f4(ps.prev_p, __uint_as_float(0u));
            w.rng[i] = make_uint4((uint32_t) ps.rng.state, (uint32_t) (ps.rng.state >> 32), (uint32_t) ps.rng.inc, (uint32_t) (ps.rng.inc >> 32));
            f = WF_HAS_RESULT | WF_PREV_DELTA | (s.max_depth != 0 ? WF_ALIVE : 0u);      // mvpath_single.h:86-87
        }
        w.flags[i] = f;
    }
    unsigned live = __ballot_sync(0xffffffffu, (f & (WF_ALIVE | WF_SHADOW | WF_HAS_RESULT)) != 0);
    if (lane_id == 0 && live) atomicAdd(w.live + (iter & 63u), (uint32_t) __popc(live));
    if (i == 0) {
        for (int k = 0; k < WF_N_TYPES; ++k) w.qcount[k] = 0;
        w.live[(iter + 32u) & 63u] = 0;
    }
}

template <bool STATS> __global__ void __launch_bounds__(128)
wf_trace_kernel(const DScene s, const RenderParams p, const WfPool w) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane_id = threadIdx.x & 31u;
    uint32_t f = w.flags[i], f0 = f, key = 0xffu;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    uint32_t *cn = STATS ? &st.nodes : nullptr, *ct = STATS ? &st.tris : nullptr;
    if (f & WF_SHADOW) {
        float4 o = w.so[i], d = w.sd[i];
        Ray r; r.o = xyz(o); r.d = xyz(d); r.maxt = o.w;
        if (intersect<true>(s, r, cn, ct).tri == AM_NO_HIT) {
            // unoccluded: the parked next-event contribution enters `result` (mvpath_single.h:199)
            float4 res = w.res[i]; V3 nr = fmadd(xyz(w.sT[i]), xyz(w.sV[i]), xyz(res));
            w.res[i] = f4(nr, res.w);
        }
        f &= ~WF_SHADOW;
    }
    if (f & WF_ALIVE) {
        float4 o = w.ro[i], d = w.rd[i];
        Ray r; r.o = xyz(o); r.d = xyz(d); r.maxt = o.w;
        Hit h = intersect<false>(s, r, cn, ct);
        if (h.tri == AM_NO_HIT) {
            f &= ~WF_ALIVE;                 // no environment emitter: the path ends, valid_ray stays as it is
            if (STATS) { st.closest++; st.vertices++; }
        } else {
            w.hit[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
            key = __float_as_uint(ldg4(s.tris + 3 * (size_t) h.tri + 2).w) & 7u;
        }
    }
    if (f != f0) w.flags[i] = f;
    // append to the queue of the hit surface's BSDF plugin type: one atomic per (warp, type)
    unsigned peers = __match_any_sync(0xffffffffu, key);
    unsigned leader = __ffs(peers) - 1; uint32_t base = 0;
    if (lane_id == leader && key != 0xffu) base = atomicAdd(w.qcount + key, (uint32_t) __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (key != 0xffu) w.queue[(size_t) key * w.n_slots + base + __popc(peers & ((1u << lane_id) - 1u))] = i;
    if (STATS) flush_stats(st, p.counters);
}

template <bool STATS, int TYPE> __global__ void __launch_bounds__(128)
wf_shade_kernel(const DScene s, const RenderParams p, const WfPool w) {
    const uint32_t n = w.qcount[TYPE], stride = gridDim.x * blockDim.x;
    const uint32_t *queue = w.queue + (size_t) TYPE * w.n_slots;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
        const uint32_t i = queue[j];
        uint32_t f = w.flags[i];
        PathState ps;
        { float4 a = w.ro[i], b = w.rd[i], c = w.tp[i], d = w.res[i], e = w.pp[i]; uint4 g = w.rng[i];
          ps.ray.o = xyz(a); ps.ray.maxt = a.w; ps.ray.d = xyz(b); ps.eta = b.w; ps.throughput = xyz(c); ps.prev_bsdf_pdf = c.w;
          ps.result = xyz(d); ps.prev_p = xyz(e); ps.depth = __float_as_uint(e.w);
          ps.rng.state = (unsigned long long) g.x | ((unsigned long long) g.y << 32); ps.rng.inc = (unsigned long long) g.z | ((unsigned long long) g.w << 32);
          ps.prev_delta = (f & WF_PREV_DELTA) != 0; ps.valid_ray = (f & WF_VALID_RAY) != 0; ps.active = true;
          float4 hh = w.hit[i]; Hit h; h.t = hh.x; h.u = hh.y; h.v = hh.z; h.tri = __float_as_uint(hh.w);
          Ray shadow; V3 sT, sV;
          bool need_shadow = shade_vertex<STATS, TYPE>(s, ps, h, &shadow, &sT, &sV, &st);
          w.res[i] = f4(ps.result, d.w);
          f = WF_HAS_RESULT | (ps.valid_ray ? WF_VALID_RAY : 0u);
          if (need_shadow) { w.so[i] = f4(shadow.o, shadow.maxt); w.sd[i] = f4(shadow.d, 0.f); w.sT[i] = f4(sT, 0.f); w.sV[i] = f4(sV, 0.f); f |= WF_SHADOW; }
          if (ps.active) {
              w.ro[i] = f4(ps.ray.o, ps.ray.maxt); w.rd[i] = f4(ps.ray.d, ps.eta); w.tp[i] = f4(ps.throughput, ps.prev_bsdf_pdf);
              w.pp[i] = f4(ps.prev_p, __uint_as_float(ps.depth));
              w.rng[i] = make_uint4((uint32_t) ps.rng.state, (uint32_t) (ps.rng.state >> 32), (uint32_t) ps.rng.inc, (uint32_t) (ps.rng.inc >> 32));
              f |= WF_ALIVE | (ps.prev_delta ? WF_PREV_DELTA : 0u);
          }
          w.flags[i] = f; }
    }
    if (STATS) flush_stats(st, p.counters);
}

template <bool STATS> static void launch_shade(int type, const DScene &s, const RenderParams &p, const WfPool &w, unsigned grid, cudaStream_t st) {
    switch (type) {
        case BSDF_DIFFUSE: wf_shade_kernel<STATS, BSDF_DIFFUSE><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_CONDUCTOR: wf_shade_kernel<STATS, BSDF_CONDUCTOR><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_DIELECTRIC: wf_shade_kernel<STATS, BSDF_DIELECTRIC><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_ROUGHCONDUCTOR: wf_shade_kernel<STATS, BSDF_ROUGHCONDUCTOR><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_ROUGHDIELECTRIC: wf_shade_kernel<STATS, BSDF_ROUGHDIELECTRIC><<<grid, 128, 0, st>>>(s, p, w); break;
        default: wf_shade_kernel<STATS, BSDF_ROUGHPLASTIC><<<grid, 128, 0, st>>>(s, p, w); break;
    }
}

// One pass (p.n_lanes lane tickets) through the pool.  Returns the number of kernel launches.
uint32_t run_wavefront_pass(const DScene &s, const RenderParams &p, const WfPool &w, uint32_t type_mask, int sm_count, bool stats, cudaStream_t st) {
    AM_CUDA(cudaMemsetAsync(w.flags, 0, sizeof(uint32_t) * w.n_slots, st));
    AM_CUDA(cudaMemsetAsync(w.live, 0, sizeof(uint32_t) * 64, st));
    const unsigned grid = w.n_slots / 128, sgrid = (unsigned) sm_count * 16;
    uint32_t launches = 0, next_check = (uint32_t) (p.n_lanes / w.n_slots) + 1;
    for (uint32_t iter = 0;; ++iter) {
        if (stats) wf_regen_kernel<true><<<grid, 128, 0, st>>>(s, p, w, iter); else wf_regen_kernel<false><<<grid, 128, 0, st>>>(s, p, w, iter);
        if (stats) wf_trace_kernel<true><<<grid, 128, 0, st>>>(s, p, w); else wf_trace_kernel<false><<<grid, 128, 0, st>>>(s, p, w);
        launches += 2;
        for (int t = 0; t < WF_N_TYPES; ++t) if (type_mask & (1u << t)) {
            if (stats) launch_shade<true>(t, s, p, w, sgrid, st); else launch_shade<false>(t, s, p, w, sgrid, st);
            launches++;
        }
        if (iter >= next_check) {
            uint32_t live = 0;
            AM_CUDA(cudaMemcpyAsync(&live, w.live + (iter & 63u), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            AM_CUDA(cudaStreamSynchronize(st));
            if (live == 0) break;
            // the tail shrinks geometrically: look again soon when few slots are live, later when the pool is still full
            next_check = iter + (live * 4u < w.n_slots ? 2u : 6u);
        }
    }
    AM_CUDA(cudaGetLastError());
    return launches;
}

} // namespace amvpt
