// render.cpp -- host side of MVPathIntegrator::render() (src/integrators/mvpath.cpp:7-278, wavefront branch):
// sample-count rounding, pass loop, per-pass sampler seeding, film hand-off; plus the build-defined pieces the
// reference does not have: view sharding across ranks and the adaptive per-view allocator.
#include "host.h"
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>

namespace amvpt {

// Work units of one rank.  world == 1: the whole quilt as one rectangle, lanes in the reference's row-major order.
// Otherwise the quilt is cut into view tiles (and, when there are fewer views than ranks, horizontal bands of a
// view); unit u goes to rank u mod world.  Lane indices stay global, so any partition reproduces the same samples.
std::vector<TileDesc> rank_tiles(const Scene &s, const Shard &sh, bool per_view) {
    std::vector<TileDesc> out;
    if (sh.world <= 1 && !per_view) {
        TileDesc t{}; t.x0 = 0; t.y0 = 0; t.w = s.film_w; t.h = s.film_h; t.pad = 0; out.push_back(t); return out;
    }
    Shard all; std::vector<TileDesc> views = s.view_tiles(all);
    uint32_t n_views = (uint32_t) views.size(), bands = 1;
    if (n_views < sh.world) bands = (sh.world + n_views - 1) / n_views;
    bands = std::min(bands, std::max(s.res_y, 1u));
    for (uint32_t v = 0; v < n_views; ++v) for (uint32_t b = 0; b < bands; ++b) {
        uint32_t unit = v * bands + b;
        if (unit % sh.world != sh.rank) continue;
        TileDesc t = views[v];
        uint32_t y0 = (uint32_t) ((uint64_t) t.h * b / bands), y1 = (uint32_t) ((uint64_t) t.h * (b + 1) / bands);
        t.y0 += y0; t.h = y1 - y0;
        if (t.h) out.push_back(t);
    }
    return out;
}
static unsigned long long finish_tiles(std::vector<TileDesc> &tiles) {
    unsigned long long lanes = 0; std::vector<TileDesc> keep;
    for (auto &t : tiles) { if (!t.spp) continue; t.lane_begin = lanes; lanes += (unsigned long long) t.w * t.h * t.spp; keep.push_back(t); }
    tiles.swap(keep); return lanes;
}

void render(Scene &s, const RenderOptions &opt) {
    if (!s.committed) AM_THROW("render: scene not committed to a device");
    if (s.integrator.is_mvpath && !s.multi_sensor) AM_THROW("This integrator can only be used with MultiSensor !");
    const bool reuse = s.reuse_samples() > 0;      // mvpath.cpp:17
    if (reuse) {
        if (s.d.n_adapt && opt.shard.world > 1) AM_THROW("mvpath: the `adaptive` fill-in numbers its samples over the whole wavefront (dr::compress, mvpath_multi.h:81); it is built for one rank");
        if (s.rfilter_type != 0) AM_THROW("mvpath: sample reuse needs the box reconstruction filter in this build");
        if (opt.adaptive) AM_THROW("render: the adaptive allocator keeps per-pixel moments of the owning view; it cannot be combined with sample reuse");
        if (s.d.n_reuse != s.reuse_samples()) AM_THROW("render: integrator parameters changed after commit");
        if (s.general_emitters()) AM_THROW("mvpath: sample reuse is built for scenes lit by area emitters; point / constant / envmap emitters need sa_reuse = false");
    }
    if (opt.shard.world == 0 || opt.shard.rank >= opt.shard.world) AM_THROW("render: invalid shard %u/%u", opt.shard.rank, opt.shard.world);
    AM_CUDA(cudaSetDevice(s.device));
    cudaStream_t st = opt.stream;
    s.stop = false; s.last_fill_paths = 0; s.alloc_log.clear();
    // mvpath.cpp:33-41
    uint32_t spp = opt.spp ? opt.spp : s.sample_count;
    if (!spp) AM_THROW("render: sample count is zero");
    uint32_t lim = s.integrator.spp_pass_lim;
    uint32_t spp_per_pass = lim ? std::min(lim, spp) : spp;
    uint32_t n_passes = spp / spp_per_pass;
    spp = n_passes * spp_per_pass;
    s.sample_count = spp;
    // mvpath.cpp:133-147
    unsigned long long wavefront = (unsigned long long) s.film_w * s.film_h * spp_per_pass, limit = 0xffffffffull;
    if (wavefront > limit) {
        spp_per_pass /= (uint32_t) ((wavefront + limit - 1) / limit);
        if (!spp_per_pass) AM_THROW("render: film too large for a 32-bit wavefront");
        n_passes = spp / spp_per_pass;
    }
    bool adaptive = opt.adaptive;
    if (adaptive && opt.shard.world > 1 && !opt.allreduce && !opt.comm) AM_THROW("render: the adaptive allocator on %u ranks needs the statistics summed over ranks: pass an amvpt_comm or an allreduce hook", opt.shard.world);
    if (adaptive && s.rfilter_type != 0) AM_THROW("render: the adaptive allocator keeps per-pixel moments and needs the box reconstruction filter");
    std::vector<TileDesc> base = rank_tiles(s, opt.shard, adaptive);
    if (base.size() > s.tiles_cap) {
        s.tiles_cap = (uint32_t) base.size(); s.tiles_dev = (TileDesc *) s.dalloc(sizeof(TileDesc) * s.tiles_cap);
    }
    uint32_t n_views = s.grid_x * s.grid_y;
    RenderStats rs;
    struct Events { cudaEvent_t a = nullptr, b = nullptr; ~Events() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); } } ev;      // released on every exit path
    AM_CUDA(cudaEventCreate(&ev.a)); AM_CUDA(cudaEventCreate(&ev.b)); cudaEvent_t e0 = ev.a, e1 = ev.b;
    AM_CUDA(cudaMemsetAsync(s.counters, 0, sizeof(unsigned long long) * CTR_COUNT, st));

    RenderParams p{};
    p.regen_min = opt.regen_min ? opt.regen_min : (getenv("AMVPT_REGEN_MIN") ? (uint32_t) atoi(getenv("AMVPT_REGEN_MIN")) : 8u); p.spp_per_pass = spp_per_pass; p.tiles = s.tiles_dev; p.film = s.film; p.moments = adaptive ? s.moments : nullptr; p.counters = s.counters;
    p.trace_min = opt.trace_min ? opt.trace_min : (getenv("AMVPT_TRACE_MIN") ? (uint32_t) atoi(getenv("AMVPT_TRACE_MIN")) : 20u);
    p.bvh_depth = s.bvh.depth;
    int mode = opt.scheduler == SCHED_WARP_BATCH ? 0 : (opt.scheduler == SCHED_STREAM ? 2 : (opt.scheduler == SCHED_WAVEFRONT ? 3 : (opt.scheduler == SCHED_SPLIT ? 6 : 1)));
    if (opt.scheduler == SCHED_AUTO) {
        // Measured on B200 (profiles/): the persistent megakernel wins while its working code stays small and rays are short
        // (Cornell box: 1.7 vs 0.8-1.1 Gpaths/s); with several BSDF plugin types or a deep tree its warps stall on
        // instruction fetch and run a handful of lanes, and the wavefront is about twice as fast (1M triangles: 222 vs 116).
        uint32_t types = 0, n_types = 0; for (auto &m : s.meshes) types |= 1u << s.bsdfs[m.bsdf].type;
        for (uint32_t t = types; t; t &= t - 1) n_types++;
        mode = (n_types >= 3 || s.n_tris >= 32768u) ? 3 : 1;
    }
    if (getenv("AMVPT_SCHEDULER")) mode = atoi(getenv("AMVPT_SCHEDULER"));
    // a reconstruction filter other than the box is splatted by the wavefront's regen stage only: carrying the sample position
    // and the footprint loop through the persistent megakernels costs them 8 % on the box-filter path (measured, profiles/)
    if (s.rfilter_type != 0) mode = 3;
    if (s.general_emitters() && mode != 3) mode = 1;                // point / constant / envmap emitters: regenerating megakernel or wavefront
    if (mode == SCHED_SPLIT && (opt.stats || adaptive)) mode = 1;   // the counting pass and the moment film use the unsplit megakernel
    if (reuse) mode = SCHED_REUSE;                 // the reuse branch has its own kernel (reuse.cu)
    rs.scheduler_used = (uint32_t) mode;
    auto launch = [&](std::vector<TileDesc> &tiles, uint32_t pass) {
        unsigned long long lanes = finish_tiles(tiles);
        if (!lanes) return;
        AM_CUDA(cudaMemcpyAsync(s.tiles_dev, tiles.data(), sizeof(TileDesc) * tiles.size(), cudaMemcpyHostToDevice, st));
        AM_CUDA(cudaMemsetAsync(s.counters + CTR_TICKET, 0, sizeof(unsigned long long), st));
        p.seed_value = s.sampler_seed + (spp_per_pass * pass + opt.seed);     // mvpath.cpp:226-231, sampler.cpp:129
        p.n_tiles = (uint32_t) tiles.size(); p.n_lanes = lanes; p.adapt_mask = nullptr;
        if (reuse && s.d.n_adapt) {
            if (s.adapt_cap < lanes) { s.adapt_mask = (uint8_t *) s.dalloc((size_t) lanes); s.adapt_scratch = (uint32_t *) s.dalloc(sizeof(uint32_t) * ((size_t) (lanes / 1024ull) + 4)); s.adapt_cap = lanes; }
            p.adapt_mask = s.adapt_mask;
        }
        if (mode == 3) {
            if (!s.wf.n_slots) {
                // pool size: a multiple of the CTA size; 2^22 slots (~860 MB of state and queues; measured 193 / 209 / 217 Mpaths/s at 2^20 / 2^21 / 2^22 on the 1M-triangle scene)
                uint32_t n = getenv("AMVPT_WF_POOL") ? (uint32_t) atoi(getenv("AMVPT_WF_POOL")) : (1u << 22);
                n = std::max(128u, (n + 127u) / 128u * 128u);
                WfPool &w = s.wf; w.n_slots = n;
                if (AM_WF_AOS) {       // three 64-byte records per slot (kernels.cuh)
                    float4 *A = (float4 *) s.dalloc(sizeof(float4) * 4 * (size_t) n), *B = (float4 *) s.dalloc(sizeof(float4) * 4 * (size_t) n), *Cc = (float4 *) s.dalloc(sizeof(float4) * 4 * (size_t) n);
                    w.ro.base = w.rd.base = w.tp.base = w.pp.base = A; w.res.base = w.hit.base = B; w.rng.base = (uint4 *) B; w.so.base = w.sd.base = w.sT.base = w.sV.base = Cc;
                } else {
                    float4 **f4s[] = { &w.ro.base, &w.rd.base, &w.tp.base, &w.res.base, &w.pp.base, &w.hit.base, &w.so.base, &w.sd.base, &w.sT.base, &w.sV.base };
                    for (float4 **a : f4s) *a = (float4 *) s.dalloc(sizeof(float4) * n);
                    w.rng.base = (uint4 *) s.dalloc(sizeof(uint4) * n);
                }
                w.flags = (uint32_t *) s.dalloc(sizeof(uint32_t) * n); w.spos = (float2 *) s.dalloc(sizeof(float2) * n);
                w.queue = (uint32_t *) s.dalloc(sizeof(uint32_t) * 7 * (size_t) n); w.qcount = (uint32_t *) s.dalloc(sizeof(uint32_t) * 8);      // six BSDF-type queues + the escaped-ray queue
                w.live = (uint32_t *) s.dalloc(sizeof(uint32_t) * 64);
                w.sort_hist = w.sort_key = w.sort_rank = w.order = nullptr;
                if (getenv("AMVPT_SORT") && atoi(getenv("AMVPT_SORT")) != 0) {      // ray-task sort: measured slower on B200 (profiles/r02_hfield16_trace_experiments.txt), opt-in
                w.sort_hist = (uint32_t *) s.dalloc(sizeof(uint32_t) * (WF_SORT_BUCKETS + 1)); w.sort_key = (uint32_t *) s.dalloc(sizeof(uint32_t) * 2 * (size_t) n);
                w.sort_rank = (uint32_t *) s.dalloc(sizeof(uint32_t) * 2 * (size_t) n); w.order = (uint32_t *) s.dalloc(sizeof(uint32_t) * 2 * (size_t) n);
                }
                s.wf_type_mask = 0; for (auto &m : s.meshes) s.wf_type_mask |= 1u << s.bsdfs[m.bsdf].type;
            }
            rs.launches += run_wavefront_pass(s.d, p, s.wf, s.wf_type_mask, s.sm_count, opt.stats, st, opt.time_stages ? &rs.trace_seconds : nullptr, &rs.trace_launches);
        } else if (mode == SCHED_SPLIT) {
            const uint32_t m_cap = getenv("AMVPT_SPLIT_CHUNK") ? std::max(128u, (uint32_t) atoi(getenv("AMVPT_SPLIT_CHUNK"))) : (1u << 23);
            if (!s.split_buf || s.split_cap != m_cap) { s.split_buf = (float4 *) s.dalloc(sizeof(float4) * 3 * (size_t) m_cap); s.split_cap = m_cap; }
            rs.launches += run_split_pass(s.d, p, s.split_buf, m_cap, s.sm_count, st);
            AM_CUDA(cudaGetLastError());
        } else if (mode == SCHED_REUSE) {
            // 0: one persistent kernel; 1: primary / project / shade kernels per chunk; 2 (default): primary / project / select / suffix / finish
            const int schedule = opt.stats ? 0 : (getenv("AMVPT_REUSE_STAGED") ? atoi(getenv("AMVPT_REUSE_STAGED")) : 2);
            if (schedule) {
                // chunk size: 2^22 lanes, less for large groups so that the projection records stay near 1 GB
                const uint32_t nk = s.d.n_reuse - 1u;
                uint32_t m_cap = getenv("AMVPT_REUSE_CHUNK") ? (uint32_t) atoi(getenv("AMVPT_REUSE_CHUNK")) : std::min(1u << 22, (1u << 25) / nk);
                m_cap = std::max(128u, m_cap / 128u * 128u);
                if (!s.reuse_buf || s.reuse_cap != m_cap) { s.reuse_buf = (float4 *) s.dalloc(sizeof(float4) * reuse_staged_float4s(s.d.n_reuse, m_cap)); s.reuse_cap = m_cap; }
                rs.launches += run_reuse_staged(s.d, p, s.reuse_buf, m_cap, schedule, s.sm_count, st);
            } else {
                launch_reuse_pass(s.d, p, s.sm_count, opt.stats, st);
                rs.launches++;
            }
            AM_CUDA(cudaGetLastError());
            if (s.d.n_adapt && !s.d.reuse_debug) {      // fill-in pass for the lanes whose sample no other camera could reuse (mvpath_multi.h:79-115)
                unsigned long long n_fill = 0;
                rs.launches += run_reuse_fill_in(s.d, p, s.sampler_seed, s.adapt_scratch, &s.adapt_compact, &s.adapt_compact_cap, s.sm_count, st, &n_fill);
                AM_CUDA(cudaGetLastError()); s.last_fill_paths += n_fill;
            }
        } else {
            launch_render_pass(s.d, p, mode, s.sm_count, opt.stats, st);
            AM_CUDA(cudaGetLastError());
            rs.launches++;
        }
        rs.paths += lanes;
    };

    // Integrator::should_stop() (include/mitsuba/render/integrator.h:303-306): cancel() or the `timeout` property (seconds of render time).  The
    // reference tests it between work units; here between passes -- with a timeout set the host waits for each pass so that the clock means something
    const auto t_start = std::chrono::steady_clock::now(); const float timeout = s.integrator.timeout;
    auto should_stop = [&]() {
        if (s.stop) return true;
        if (timeout > 0.f) { cudaStreamSynchronize(st); if (std::chrono::duration<float>(std::chrono::steady_clock::now() - t_start).count() > timeout) return true; }
        return false;
    };
    AM_CUDA(cudaEventRecord(e0, st));
    if (!adaptive) {
        for (uint32_t pass = 0; pass < n_passes && !should_stop(); ++pass) {
            std::vector<TileDesc> tiles = base; for (auto &t : tiles) t.spp = spp_per_pass;
            launch(tiles, pass); rs.passes++;
        }
        rs.spp_done = rs.passes * spp_per_pass;
    } else {
        // Adaptive allocation (build-defined).  Budget: spp * (pixels of the whole quilt) samples.  Warm-up passes give
        // every view adaptive_min_spp; then each pass hands view v a sample count n_v <= spp_per_pass proportional to its
        // error estimate (mean relative variance of the pixel means), recomputed from all-reduced statistics so that every
        // rank derives the same allocation.
        unsigned long long view_px = (unsigned long long) s.res_x * s.res_y, budget = (unsigned long long) spp * view_px * n_views, used = 0;
        uint32_t warm = std::max(1u, (std::min(opt.adaptive_min_spp, spp) + spp_per_pass - 1) / spp_per_pass);
        std::vector<float> hstats(4 * (size_t) n_views), err(n_views, 1.f);
        if (n_views > s.tiles_cap) s.view_stats_dev = (float *) s.dalloc(sizeof(float) * 4 * n_views);
        Shard all; std::vector<TileDesc> all_views = s.view_tiles(all);
        std::vector<uint32_t> alloc(n_views, spp_per_pass);
        // the reference's lane map allows spp / spp_per_pass passes of distinct seeds; adaptive passes keep going with
        // fresh seeds (pass index) until the budget is spent
        uint32_t max_passes = 4 * n_passes + 16;
        for (uint32_t pass = 0; pass < max_passes && used < budget && !should_stop(); ++pass) {
            if (pass >= warm) {
                // statistics of the views this rank owns: ONE launch over the owned tile list (bands of a view accumulate into the view's
                // slot on the device), summed over ranks in place (NCCL on the render stream, or the caller's hook), one read-back per pass
                std::vector<TileDesc> mine = base; for (auto &t : mine) t.spp = 1;
                AM_CUDA(cudaMemsetAsync(s.view_stats_dev, 0, sizeof(float) * 4 * n_views, st));
                if (!mine.empty()) {
                    AM_CUDA(cudaMemcpyAsync(s.tiles_dev, mine.data(), sizeof(TileDesc) * mine.size(), cudaMemcpyHostToDevice, st));
                    launch_view_stats(s.film, s.moments, s.film_w, s.tiles_dev, (uint32_t) mine.size(), s.view_stats_dev, st);
                }
                if (opt.comm) comm_allreduce_sum(opt.comm, s.view_stats_dev, 4 * n_views, st);
                else if (opt.allreduce) { AM_CUDA(cudaStreamSynchronize(st)); opt.allreduce(s.view_stats_dev, 4 * n_views, opt.allreduce_user); }      // the hook completes before it returns
                AM_CUDA(cudaMemcpyAsync(hstats.data(), s.view_stats_dev, sizeof(float) * 4 * n_views, cudaMemcpyDeviceToHost, st));
                AM_CUDA(cudaStreamSynchronize(st));
                double total = 0.0; bool any = false;
                for (uint32_t v = 0; v < n_views; ++v) {
                    float e = hstats[4 * v + 1] > 0.f ? hstats[4 * v] / hstats[4 * v + 1] : 0.f;
                    if (opt.adaptive_threshold > 0.f && e <= opt.adaptive_threshold) e = 0.f;
                    err[v] = e; total += e; any |= e > 0.f;
                }
                if (!any) break;
                unsigned long long remaining = budget - used;
                double pass_budget = std::min<double>((double) remaining / (double) view_px, (double) spp_per_pass * n_views);
                unsigned long long give = 0;
                for (uint32_t v = 0; v < n_views; ++v) {
                    uint32_t n = (uint32_t) std::llround(pass_budget * err[v] / total);
                    n = std::min(n, spp_per_pass);
                    if (give + (unsigned long long) n * view_px > remaining) n = (uint32_t) ((remaining - give) / view_px);
                    alloc[v] = n; give += (unsigned long long) n * view_px;
                }
                if (!give) break;
            } else {
                unsigned long long remaining = budget - used, give = 0;
                for (uint32_t v = 0; v < n_views; ++v) {
                    uint32_t n = spp_per_pass;
                    if (give + (unsigned long long) n * view_px > remaining) n = (uint32_t) ((remaining - give) / view_px);
                    alloc[v] = n; give += (unsigned long long) n * view_px;
                }
            }
            for (uint32_t v = 0; v < n_views; ++v) used += (unsigned long long) alloc[v] * view_px;
            s.alloc_log.insert(s.alloc_log.end(), alloc.begin(), alloc.end());
            std::vector<TileDesc> tiles = base; for (auto &t : tiles) t.spp = alloc[t.pad];
            launch(tiles, pass); rs.passes++;
        }
        rs.spp_done = (uint32_t) (used / (view_px * n_views));
    }
    AM_CUDA(cudaEventRecord(e1, st));
    AM_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f; AM_CUDA(cudaEventElapsedTime(&ms, e0, e1)); rs.seconds = ms * 1e-3;
    unsigned long long c[CTR_COUNT]; AM_CUDA(cudaMemcpy(c, s.counters, sizeof(c), cudaMemcpyDeviceToHost));
    rs.closest_rays = c[CTR_CLOSEST]; rs.shadow_rays = c[CTR_SHADOW]; rs.nodes = c[CTR_NODES]; rs.tris = c[CTR_TRIS]; rs.vertices = c[CTR_VERTICES];
    s.last = rs;
}

} // namespace amvpt
