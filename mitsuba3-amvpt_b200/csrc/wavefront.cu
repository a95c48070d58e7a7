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
#include <algorithm>
#include <cstdlib>
#include <vector>

namespace amvpt {

enum : uint32_t { WF_PREV_DELTA = 1u, WF_VALID_RAY = 2u, WF_ALIVE = 4u, WF_HAS_RESULT = 8u, WF_SHADOW = 16u };
#define WF_NO_PIXEL 0xffffffffu
#define WF_N_TYPES 6
#define WF_MISS_KEY 7u     /* qcount[6] is the ray-task counter; escaped rays are counted in qcount[7] and stored in queue row 6 */

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
        if (s.filter_type == 0) splat_aggregated(dm, put && pixel != WF_NO_PIXEL, pixel, make_float4(L.x, L.y, L.z, 1.f), p.film, p.moments);
        else if (put) { float2 sp = w.spos[i]; splat_filtered(s, sp.x, sp.y, L, p.film); }
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
            if (s.filter_type != 0) w.spos[i] = make_float2(ls.spx, ls.spy);
            f = WF_HAS_RESULT | WF_PREV_DELTA | (s.max_depth != 0 ? WF_ALIVE | (ps.valid_ray ? WF_VALID_RAY : 0u) : 0u);      // mvpath_single.h:86-87, 98
        }
        w.flags[i] = f;
    }
    unsigned live = __ballot_sync(0xffffffffu, (f & (WF_ALIVE | WF_SHADOW | WF_HAS_RESULT)) != 0);
    if (lane_id == 0 && live) atomicAdd(w.live + (iter & 63u), (uint32_t) __popc(live));
    if (i == 0) {
        for (int k = 0; k <= WF_N_TYPES + 1; ++k) w.qcount[k] = 0;  // type queues + the ray-task counter + the escaped-ray queue
        w.live[(iter + 32u) & 63u] = 0;
    }
}

// Ray tasks of one iteration: task 2 i = the pending shadow ray of slot i, task 2 i + 1 = its closest-hit ray (a task exists
// if the slot's flag says so).  Persistent warps pull tasks from one counter: whenever at least fetch_min lanes of a warp have
// finished their ray they take the next task ids together (one atomic per warp, ballot + popc prefix), so the traversal loop
// keeps running with nearly full warps although rays differ widely in length (Aila & Laine's dynamic fetch).
#ifndef AM_WT_MIN_CTAS
#define AM_WT_MIN_CTAS 10   /* 51 registers: +2 % on the 1M-triangle scene; 12 (40 registers) spills and loses 3 % */
#endif
// ---- ray-task sort.  Rays leave the shade stage in slot order, i.e. uncorrelated in space; traced that way the lanes of a warp walk
// different parts of the tree (ncu: 16 of 32 lanes per instruction, L1 hit 67 %).  Before the trace stage the existing tasks are
// bucketed by (Morton code of the origin's cell in a 16^3 grid over the scene box, direction octant): one pass computes the bucket
// and takes a rank in it (match.any-aggregated atomics on a 32768-entry histogram), one CTA scans the histogram, one pass scatters
// the task ids.  Warps then pull consecutive entries of order[]: neighbours in space, same octant.
AM_D uint32_t morton4(uint32_t x, uint32_t y, uint32_t z) {       // 4 bits per axis -> 12 bits
    auto spread = [](uint32_t v) { v &= 15u; v = (v | (v << 4)) & 0x0c3u; v = (v | (v << 2)) & 0x249u; return v; };
    return spread(x) | (spread(y) << 1) | (spread(z) << 2);
}
AM_D uint32_t ray_bucket(const DScene &s, float4 o, float4 d) {
    const float fx = (o.x - s.box_lo[0]) * s.box_inv[0], fy = (o.y - s.box_lo[1]) * s.box_inv[1], fz = (o.z - s.box_lo[2]) * s.box_inv[2];
    const uint32_t cx = (uint32_t) fminf(fmaxf(fx, 0.f), 15.f), cy = (uint32_t) fminf(fmaxf(fy, 0.f), 15.f), cz = (uint32_t) fminf(fmaxf(fz, 0.f), 15.f);
    const uint32_t oct = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u);
    return (morton4(cx, cy, cz) << 3) | oct;
}
__global__ void __launch_bounds__(256) wf_key_kernel(const DScene s, const WfPool w) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;          // one thread per slot, two tasks
    const unsigned lane_id = threadIdx.x & 31u;
    const uint32_t f = i < w.n_slots ? w.flags[i] : 0u;
    #pragma unroll
    for (int kind = 0; kind < 2; ++kind) {
        const bool exists = kind == 0 ? (f & WF_SHADOW) != 0 : (f & WF_ALIVE) != 0;
        uint32_t key = WF_SORT_NONE;
        if (exists) key = kind == 0 ? ray_bucket(s, w.so[i], w.sd[i]) : ray_bucket(s, w.ro[i], w.rd[i]);
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        uint32_t base = 0; const unsigned leader = __ffs(peers) - 1;
        if (exists && lane_id == leader) base = atomicAdd(w.sort_hist + key, (uint32_t) __popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (i < w.n_slots) { w.sort_key[2 * i + kind] = key; w.sort_rank[2 * i + kind] = base + __popc(peers & ((1u << lane_id) - 1u)); }
    }
}
__global__ void __launch_bounds__(1024) wf_scan_kernel(const WfPool w) {      // exclusive scan of 32768 counts by one CTA; [BUCKETS] = total
    __shared__ uint32_t part[1024];
    const uint32_t t = threadIdx.x, per = WF_SORT_BUCKETS / 1024u; uint32_t v[WF_SORT_BUCKETS / 1024u], sum = 0;
    for (uint32_t k = 0; k < per; ++k) { v[k] = w.sort_hist[t * per + k]; sum += v[k]; }
    part[t] = sum; __syncthreads();
    for (uint32_t off = 1; off < 1024u; off <<= 1) { uint32_t x = t >= off ? part[t - off] : 0u; __syncthreads(); part[t] += x; __syncthreads(); }
    uint32_t run = part[t] - sum;
    for (uint32_t k = 0; k < per; ++k) { w.sort_hist[t * per + k] = run; run += v[k]; }
    if (t == 1023u) w.sort_hist[WF_SORT_BUCKETS] = part[1023];
}
__global__ void __launch_bounds__(256) wf_scatter_kernel(const WfPool w) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; if (t >= 2u * w.n_slots) return;
    const uint32_t key = w.sort_key[t]; if (key == WF_SORT_NONE) return;
    w.order[w.sort_hist[key] + w.sort_rank[t]] = t;
}

// Q4 = true walks the compressed 4-wide tree (DScene::nodes_q4: half the node steps of the binary tree at the same 64 bytes per step);
// SORTED = true pulls tasks from order[] (all of them exist) instead of testing every slot's two task flags
template <bool STATS, bool Q4, bool SORTED> __global__ void __launch_bounds__(128, AM_WT_MIN_CTAS)
wf_trace_kernel(const DScene s, const RenderParams p, const WfPool w, uint32_t fetch_min, uint32_t leaf_min) {
    const unsigned lane_id = threadIdx.x & 31u;
    const uint32_t n_tasks = SORTED ? w.sort_hist[WF_SORT_BUCKETS] : 2u * w.n_slots;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    int32_t stack[AM_STACK_SIZE];
    TravState tv; tv.node = AM_TRAVERSAL_DONE; tv.sp = 0; tv.any = false;
    uint32_t slot = 0; bool busy = false, exhausted = false;
    for (;;) {
        unsigned idle = __ballot_sync(0xffffffffu, !busy);
        if (idle == 0xffffffffu || (!exhausted && __popc(idle) >= (int) fetch_min)) {
            if (exhausted) break;                       // every lane idle and no task left
            uint32_t t = n_tasks;
            if (!busy) {
                uint32_t base = 0; unsigned leader = __ffs(idle) - 1;
                if (lane_id == leader) base = atomicAdd(w.qcount + WF_N_TYPES, (uint32_t) __popc(idle));
                base = __shfl_sync(idle, base, leader);
                t = base + __popc(idle & ((1u << lane_id) - 1u));
                if (t < n_tasks) {
                    const uint32_t task = SORTED ? w.order[t] : t;
                    slot = task >> 1; const uint32_t f = SORTED ? (WF_SHADOW | WF_ALIVE) : w.flags[slot];
                    if ((task & 1u) == 0) {
                        if (f & WF_SHADOW) { float4 o = w.so[slot], d = w.sd[slot]; Ray r; r.o = xyz(o); r.d = xyz(d); r.maxt = o.w; trav_begin(tv, r, true); busy = true; }
                    } else if (f & WF_ALIVE) { float4 o = w.ro[slot], d = w.rd[slot]; Ray r; r.o = xyz(o); r.d = xyz(d); r.maxt = o.w; trav_begin(tv, r, false); busy = true; }
                    if (busy) { stack[0] = AM_TRAVERSAL_DONE; tv.sp = 1; }
                }
            }
            exhausted = __any_sync(0xffffffffu, !busy && t >= n_tasks && (idle >> lane_id & 1u));
            continue;
        }
        uint32_t key = 0xffu;
        // descend: lanes holding an inner node step through the tree together; the loop is left as soon as at least leaf_min
        // lanes hold a leaf (or a finished ray), so that no lane waits long with a leaf while a few others wander through
        // inner nodes, and the few that are interrupted resume after the leaf phase
        {
            const unsigned bm = __ballot_sync(0xffffffffu, busy);
            for (;;) {
                const bool desc = busy && tv.node >= 0;
                const unsigned dm = __ballot_sync(0xffffffffu, desc);
                if (dm == 0 || __popc(bm & ~dm) >= (int) leaf_min) break;
                if (desc) {
                    if (STATS) st.nodes++;
                    tv.node = Q4 ? node_step_q4(s, tv.sr, tv.best, tv.node, stack, tv.sp) : node_step(s, tv.sr, tv.best, tv.node, stack, tv.sp);
                }
            }
        }
        if (busy && tv.node < 0) {
            if (tv.node != AM_TRAVERSAL_DONE) {
                uint32_t code = ~(uint32_t) tv.node, first = code >> 3, count = code & 7u;
                bool occluded = false;
                for (uint32_t k = 0; k < count; ++k) {
                    const float4 *tp = s.tris + 3 * (size_t) (first + k);
                    float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
                    if (STATS) st.tris++;
                    float t, u, v;
                    if (moeller_trumbore(tv.ray, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &u, &v)) {
                        if (tv.any) { occluded = true; break; }
                        uint32_t prim = __float_as_uint(a.w);
                        if (tv.h.tri == AM_NO_HIT || t < tv.h.t || (t == tv.h.t && prim > tv.best_prim)) {
                            tv.h.t = t; tv.h.u = u; tv.h.v = v; tv.h.tri = first + k; tv.best_prim = prim; tv.best = t;
                        }
                    }
                }
                if (occluded) { tv.h.tri = 0; tv.node = AM_TRAVERSAL_DONE; }
                else tv.node = stack[--tv.sp];
            }
            if (tv.node == AM_TRAVERSAL_DONE) {
                busy = false;
                if (tv.any) {
                    if (tv.h.tri == AM_NO_HIT) {
                        // unoccluded: the parked next-event contribution enters `result` (mvpath_single.h:199)
                        float4 res = w.res[slot]; V3 nr = fmadd(xyz(w.sT[slot]), xyz(w.sV[slot]), xyz(res));
                        w.res[slot] = f4(nr, res.w);
                    }
                    atomicAnd(w.flags + slot, ~WF_SHADOW);
                } else if (tv.h.tri == AM_NO_HIT) {
                    atomicAnd(w.flags + slot, ~WF_ALIVE);      // the path ends, valid_ray stays as it is; with an environment
                    if (s.env_emitter >= 0) key = WF_MISS_KEY;  // emitter the escaped ray still collects its radiance (wf_miss_kernel)
                    if (STATS) { st.closest++; st.vertices++; }
                } else {
                    w.hit[slot] = make_float4(tv.h.t, tv.h.u, tv.h.v, __uint_as_float(tv.h.tri));
                    key = __float_as_uint(ldg4(s.tris + 3 * (size_t) tv.h.tri + 2).w) & 7u;
                }
            }
        }
        // closest hits found in this step: append to the queue of the surface's BSDF plugin type, one atomic per (warp, type)
        unsigned fin = __ballot_sync(0xffffffffu, key != 0xffu);
        if (fin) {
            unsigned peers = __match_any_sync(0xffffffffu, key);
            unsigned leader = __ffs(peers) - 1; uint32_t base = 0;
            if (lane_id == leader && key != 0xffu) base = atomicAdd(w.qcount + key, (uint32_t) __popc(peers));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (key != 0xffu) w.queue[(size_t) (key == WF_MISS_KEY ? WF_N_TYPES : key) * w.n_slots + base + __popc(peers & ((1u << lane_id) - 1u))] = slot;
        }
    }
    if (STATS) flush_stats(st, p.counters);
}

// Escaped rays of scenes with an environment emitter: emission of the environment with the MIS weight of the previous vertex
// (mvpath_single.h:148-163 for an invalid interaction).  One thread per entry of the escaped-ray queue.
__global__ void __launch_bounds__(128) wf_miss_kernel(const DScene s, const WfPool w) {
    const uint32_t n = w.qcount[WF_MISS_KEY], stride = gridDim.x * blockDim.x;
    const uint32_t *queue = w.queue + (size_t) WF_N_TYPES * w.n_slots;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
        const uint32_t i = queue[j]; const uint32_t f = w.flags[i];
        PathState ps; const float4 b = w.rd[i], c = w.tp[i], d = w.res[i];
        ps.ray.d = xyz(b); ps.throughput = xyz(c); ps.prev_bsdf_pdf = c.w; ps.result = xyz(d); ps.prev_delta = (f & WF_PREV_DELTA) != 0;
        SI si; si.valid = false;
        add_emission<true>(s, ps, si, nullptr);
        w.res[i] = f4(ps.result, d.w);
    }
}

template <bool STATS, int TYPE, bool GEN> __global__ void __launch_bounds__(128)
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
          bool need_shadow = shade_vertex<STATS, TYPE, GEN>(s, ps, h, &shadow, &sT, &sV, &st);
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

template <bool STATS, bool GEN> static void launch_shade(int type, const DScene &s, const RenderParams &p, const WfPool &w, unsigned grid, cudaStream_t st) {
    switch (type) {
        case BSDF_DIFFUSE: wf_shade_kernel<STATS, BSDF_DIFFUSE, GEN><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_CONDUCTOR: wf_shade_kernel<STATS, BSDF_CONDUCTOR, GEN><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_DIELECTRIC: wf_shade_kernel<STATS, BSDF_DIELECTRIC, GEN><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_ROUGHCONDUCTOR: wf_shade_kernel<STATS, BSDF_ROUGHCONDUCTOR, GEN><<<grid, 128, 0, st>>>(s, p, w); break;
        case BSDF_ROUGHDIELECTRIC: wf_shade_kernel<STATS, BSDF_ROUGHDIELECTRIC, GEN><<<grid, 128, 0, st>>>(s, p, w); break;
        default: wf_shade_kernel<STATS, BSDF_ROUGHPLASTIC, GEN><<<grid, 128, 0, st>>>(s, p, w); break;
    }
}

// One pass (p.n_lanes lane tickets) through the pool.  Returns the number of kernel launches.
uint32_t run_wavefront_pass(const DScene &s, const RenderParams &p, const WfPool &w, uint32_t type_mask, int sm_count, bool stats, cudaStream_t st,
                            double *trace_seconds, uint32_t *trace_launches) {
    AM_CUDA(cudaMemsetAsync(w.flags, 0, sizeof(uint32_t) * w.n_slots, st));
    AM_CUDA(cudaMemsetAsync(w.live, 0, sizeof(uint32_t) * 64, st));
    const unsigned grid = w.n_slots / 128, sgrid = (unsigned) sm_count * 16;
    // measured on B200 (profiles/r02_*): the compressed 4-wide tree and the task sort are opt-in / opt-out through the environment
    const bool q4 = s.nodes_q4 != nullptr && getenv("AMVPT_Q4") && atoi(getenv("AMVPT_Q4")) != 0;
    const bool sorted = w.order != nullptr;      // the sort buffers exist only when AMVPT_SORT=1 was set at pool creation (render.cpp)
    typedef void (*TraceFn)(const DScene, const RenderParams, const WfPool, uint32_t, uint32_t);
    const TraceFn fns[2][2][2] = { { { wf_trace_kernel<false, false, false>, wf_trace_kernel<false, false, true> }, { wf_trace_kernel<false, true, false>, wf_trace_kernel<false, true, true> } },
                                   { { wf_trace_kernel<true, false, false>, wf_trace_kernel<true, false, true> }, { wf_trace_kernel<true, true, false>, wf_trace_kernel<true, true, true> } } };
    const TraceFn trace_fn = fns[stats ? 1 : 0][q4 ? 1 : 0][sorted ? 1 : 0];
    int trace_ctas = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trace_ctas, trace_fn, 128, 0); if (trace_ctas < 1) trace_ctas = 1;
    const unsigned tgrid = std::min<unsigned>((unsigned) (sm_count * trace_ctas), grid);      // persistent: one resident wave
    const uint32_t fetch_min = getenv("AMVPT_FETCH_MIN") ? (uint32_t) atoi(getenv("AMVPT_FETCH_MIN")) : 16u;
    const uint32_t leaf_min = getenv("AMVPT_LEAF_MIN") ? (uint32_t) atoi(getenv("AMVPT_LEAF_MIN")) : 12u;
    uint32_t launches = 0, next_check = (uint32_t) (p.n_lanes / w.n_slots) + 1;
    std::vector<cudaEvent_t> ev;        // optional: device time of the trace stage (bench.py roofline)
    for (uint32_t iter = 0;; ++iter) {
        if (stats) wf_regen_kernel<true><<<grid, 128, 0, st>>>(s, p, w, iter); else wf_regen_kernel<false><<<grid, 128, 0, st>>>(s, p, w, iter);
        if (trace_seconds) { cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b); ev.push_back(a); ev.push_back(b); cudaEventRecord(a, st); }
        if (sorted) {
            cudaMemsetAsync(w.sort_hist, 0, sizeof(uint32_t) * (WF_SORT_BUCKETS + 1), st);
            wf_key_kernel<<<(w.n_slots + 255) / 256, 256, 0, st>>>(s, w); wf_scan_kernel<<<1, 1024, 0, st>>>(w); wf_scatter_kernel<<<(2 * w.n_slots + 255) / 256, 256, 0, st>>>(w);
            launches += 3;
        }
        trace_fn<<<tgrid, 128, 0, st>>>(s, p, w, fetch_min, leaf_min);
        if (trace_seconds) cudaEventRecord(ev.back(), st);
        launches += 2;
        if (s.env_emitter >= 0) { wf_miss_kernel<<<sgrid, 128, 0, st>>>(s, w); launches++; }
        for (int t = 0; t < WF_N_TYPES; ++t) if (type_mask & (1u << t)) {
            if (s.gen_emitters) { if (stats) launch_shade<true, true>(t, s, p, w, sgrid, st); else launch_shade<false, true>(t, s, p, w, sgrid, st); }
            else { if (stats) launch_shade<true, false>(t, s, p, w, sgrid, st); else launch_shade<false, false>(t, s, p, w, sgrid, st); }
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
    if (trace_seconds) {
        AM_CUDA(cudaStreamSynchronize(st));
        for (size_t k = 0; k + 1 < ev.size(); k += 2) { float ms = 0.f; cudaEventElapsedTime(&ms, ev[k], ev[k + 1]); *trace_seconds += ms * 1e-3; cudaEventDestroy(ev[k]); cudaEventDestroy(ev[k + 1]); }
        if (trace_launches) *trace_launches += (uint32_t) (ev.size() / 2);
    }
    return launches;
}

} // namespace amvpt
