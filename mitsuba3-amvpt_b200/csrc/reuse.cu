// reuse.cu -- cross-view sample reuse with multiple importance sampling, the AMVPT contribution of the integrator
// (SURVEY.md sec.8f next-1): MVPT::render_multisample / sample_multi / camera_selection / mis_weights / sample_suffix
// (src/integrators/mvpath_multi.h:8-689), sensors_visible / tv_pdf / tv_pdf_fast (src/integrators/mvpath.h:243-293),
// PerspectiveCamera / ThinLensCamera ::sample_surface (src/sensors/perspective.cpp:327-385, thinlens.cpp:358-418),
// BSDF::eval_roughness (diffuse.cpp:186, roughconductor.cpp:518, roughdielectric.cpp:607, roughplastic.cpp:508, bsdf.h:596).
//
// One primary sample finds its first hit, projects that point into the other n - 1 cameras of its group (one visibility
// ray each), selects the cameras whose view of the surface is similar enough (Jacobian ratio x BSDF total-variation
// probability), weights every accepted camera with the multi-view balance heuristic, traces ONE suffix path shared by all
// of them and splats n weighted samples into the quilt.  Per lane the arithmetic is the reference's, statement by
// statement, in the JIT variants' semantics (masked virtual calls return zeros; every lane that reaches a draw draws).
//
// Schedule: a persistent kernel, one lane per thread, 32 tickets per warp fetch.  The per-camera records (92 B each) live
// in local memory; the kernel is instantiated for group sizes up to 8 / 64 so that small groups keep a small frame.
// Not built: the `adaptive` fill-in of mvpath_multi.h:79-115 (the host rejects adaptive > 0) and the `debug` output.
#include "dpath.cuh"

namespace amvpt {

struct DSample {             // mvpath.h:150-167
    V3 result, bsdf_val, wi, wo_r; float px, py, weight, pdfM, pdf, pdf_lk, Jp, iJp; uint32_t idx; bool indirect, valid;
};
struct SurfSample { V3 p, d; float uvx, uvy, pdf, Jp; bool face, active; };
struct BsdfData { const DBsdf *bsdf; float alpha, sqr_a, rsqrt_a; bool diffuse, reuse; };

// sample_surface behind the masked virtual call of GridSensor::sample_surface (grid.cpp:331-336)
AM_D SurfSample sensor_sample_surface(const DScene &s, uint32_t idx, V3 it_p, V3 it_n, V2 sample, bool active) {
    SurfSample r; r.p = v3(0.f); r.d = v3(0.f); r.uvx = r.uvy = r.pdf = r.Jp = 0.f; r.face = false; r.active = false;
    if (!active) return r;
    const DSensor &c = s.sensors[idx]; const DSensorAux &a = s.sensor_aux[idx];
    V3 ref_p = xf_point_affine(a.w2c, it_p);
    active = active && (ref_p.z >= c.near_clip) && (ref_p.z <= c.far_clip);
    const float resx = (float) (s.film_w / s.grid_x), resy = (float) (s.film_h / s.grid_y);
    float inv_dist, ictf3; V3 origin = v3(0.f);
    if (!s.thinlens) {
        V3 ss = xf_point_proj(a.c2s, ref_p);
        r.uvx = ss.x - 0.f; r.uvy = ss.y - 0.f;
        active = active && (r.uvx >= 0.f) && (r.uvx <= 1.f) && (r.uvy >= 0.f) && (r.uvy <= 1.f);
        r.uvx *= resx; r.uvy *= resy;
        float dist = norm(ref_p); inv_dist = rcp(dist);
        float ctf = ref_p.z;                  // cos_theta of the un-normalised camera-space vector (perspective.cpp:356-361)
        active = active && ctf > 0.f;
        float ictf = rcp(ctf); ictf3 = ictf * ictf * ictf;
        r.pdf = a.normalization * ictf3;
    } else {
        V2 tmp = square_to_uniform_disk_concentric(sample);
        origin = v3(tmp.x * c.aperture_radius, tmp.y * c.aperture_radius, 0.f);
        V3 local_d = ref_p - origin; float dist = norm(local_d); inv_dist = rcp(dist); local_d = local_d * inv_dist;
        float ctf = local_d.z, ictf = rcp(ctf); ictf3 = ictf * ictf * ictf;
        V3 film_plane = origin * fdiv(1.f, c.focus_distance) + (local_d / local_d.z);
        V3 scr = xf_point_affine(a.c2s, film_plane);
        active = active && (scr.x >= 0.f) && (scr.y >= 0.f) && (scr.x <= 1.f) && (scr.y <= 1.f);
        float pdf_lens = rcp(sqr(c.aperture_radius) * AM_PI), pdf_film = a.normalization * ictf3;
        r.pdf = pdf_lens * pdf_film; r.uvx = scr.x * resx; r.uvy = scr.y * resy;
    }
    r.p = xf_point_affine(c.to_world, origin);
    r.d = (r.p - it_p) * inv_dist;
    float cts = dot(r.d, it_n);
    r.face = cts > 0.f; cts = fabs_(cts);
    r.Jp = (cts * inv_dist * inv_dist) * r.pdf;
    r.active = active;
    return r;
}
// mvpath.h:243-256
template <bool PRIMARY, bool STATS> AM_D SurfSample sensors_visible(const DScene &s, V2 ap, const SI &si, bool prim_face, uint32_t idx, bool active, StatCtx *st) {
    SurfSample r = sensor_sample_surface(s, idx, si.p, si.n, ap, active);
    if (!PRIMARY) {
        r.active = r.active && (r.face == prim_face) && r.Jp > 0.f;
        if (r.active) {                       // the visibility test of a masked-out lane is unobservable
            if (STATS) st->shadow++;
            r.active = !trace_occluded(s, spawn_ray_to(si.p, si.n, r.p), STATS ? st : nullptr);
        }
    }
    return r;
}
AM_D float bsdf_roughness(const DBsdf &b) {
    switch (b.type) {
        case BSDF_DIFFUSE: return 1.f;
        case BSDF_ROUGHCONDUCTOR: case BSDF_ROUGHDIELECTRIC: return fsqrt(0.5f * (sqr(b.alpha_u) + sqr(b.alpha_v)));
        case BSDF_ROUGHPLASTIC: return b.alpha_u;
    }
    return 0.f;
}
// BSDF::eval_pdf behind a masked virtual call.  The fused eval_pdf + sample routines of dbsdf.cuh are called with a dummy
// sample: eval and pdf do not depend on it, and the unused sampling half is dead code after inlining.
AM_D void bsdf_eval_pdf(const DBsdf *b, V3 wi, V3 wo, bool active, V3 &f, float &pdf) {
    f = v3(0.f); pdf = 0.f; if (!active || !b) return;
    if ((b->flags & L_SMOOTH) == 0) return;          // delta BSDFs: eval = pdf = 0
    BsdfResult r = bsdf_call(*b, wi, wo, .5f, v2(.5f, .5f));
    f = r.f; pdf = r.pdf;
}
// BSDF::pdf with BSDFContext(Radiance, Glossy): diffuse lobes are disabled (mvpath_multi.h:371, mvpath.h:277)
AM_NOINLINE float bsdf_pdf_glossy(const DBsdf *b, V3 wi, V3 wo, bool active) {
    if (!active || !b) return 0.f;
    if (b->type == BSDF_ROUGHCONDUCTOR || b->type == BSDF_ROUGHDIELECTRIC) { V3 f; float pdf; bsdf_eval_pdf(b, wi, wo, true, f, pdf); return pdf; }
    if (b->type != BSDF_ROUGHPLASTIC) return 0.f;
    if (b->twosided) { wo.z = mulsign(wo.z, wi.z); wi.z = fabs_(wi.z); }
    V3 H = normalize(wo + wi); Microfacet d(b->distr, b->alpha_u, b->alpha_u);     // roughplastic.cpp:380-422, specular lobe only
    float result = fdiv(d.eval(H) * d.smith_g1(wi, H), 4.f * wi.z);
    result *= 1.f; result += 0.f * square_to_cosine_hemisphere_pdf(wo);
    return result;
}
AM_D float tv_tail(float q, const BsdfData &b, bool active) {
    float p = fmadd(q - 1.f, b.rsqrt_a, 1.f);
    p = sqr(max_(p, 0.f)); p = lerp_(p, q, b.alpha);
    return active ? p : 0.f;
}
AM_D float tv_any(const DScene &s, V3 wo_l, V3 wi_k, float p_k, const BsdfData &b, bool active) {
    if (s.fast_mis) {                                      // mvpath.h:258-271
        float p_l = sqr(normalize(wi_k + wo_l).z);
        float N = fmadd(b.sqr_a, max_(p_k, p_l), 1.f), D = fmadd(b.sqr_a, min_(p_k, p_l), 1.f);
        return tv_tail(sqr(N * rcp(D)), b, active);
    }
    active = active && p_k > 0.f;                          // mvpath.h:277-293
    float p_l = bsdf_pdf_glossy(b.bsdf, wi_k, wo_l, active);
    active = active && p_l > 0.f;
    float p_max = max_(p_l, p_k), p_min = min_(p_l, p_k);
    return tv_tail(p_min * rcp(p_max), b, active);
}
AM_D float pdf_material(const DScene &s, const BsdfData &b, V3 wi, V3 wo_r, bool active) {
    return s.fast_mis ? sqr(normalize(wi + wo_r).z) : bsdf_pdf_glossy(b.bsdf, wi, wo_r, active);
}
AM_D BsdfResult bsdf_zero() {
    BsdfResult r; r.f = v3(0.f); r.pdf = 0.f; r.weight = v3(0.f); r.bs.wo = v3(0.f); r.bs.pdf = 0.f; r.bs.eta = 0.f; r.bs.type = 0; return r;
}

// MVPT::camera_selection (mvpath_multi.h:360-462)
template <bool STATS> AM_D void camera_selection(const DScene &s, Pcg32 &rng, DSample *samples, uint32_t n, const SI &si, const BsdfData &bd, V3 wo, V2 p_app,
                                                 V2 rand_2, float rand_1, bool p_hit, BSDFSample &bsdf_sample, float &direct_pdf, StatCtx *st) {
    DSample &p = samples[0];
    const bool p_face = si.wi.z > 0.f;
    SurfSample pds = sensors_visible<true, STATS>(s, p_app, si, p_face, p.idx, p_hit, st);
    p.pdf = pds.pdf; p.pdf_lk = pds.pdf; p.Jp = pds.Jp; p.iJp = p_hit ? rcp(pds.Jp) : 0.f;
    p.wi = si.wi; p.wo_r = reflect(si.wi);
    p.valid = p_hit; p.indirect = p_hit;
    p.pdfM = pdf_material(s, bd, p.wi, p.wo_r, p_hit);
    if (bd.diffuse) p.pdfM = 1.f;
    float n_direct = 1.f, n_indir = 2.f;
    for (uint32_t k = 1; k < n; ++k) {
        DSample &sk = samples[k];
        SurfSample ds = sensors_visible<false, STATS>(s, p_app, si, p_face, sk.idx, bd.reuse, st);
        bool valid = ds.active;
        sk.wi = si.to_local(ds.d); sk.wo_r = reflect(sk.wi);
        sk.pdfM = pdf_material(s, bd, sk.wi, sk.wo_r, valid);
        float pdf_Mat = tv_any(s, p.wo_r, sk.wi, sk.pdfM, bd, valid);
        if (bd.diffuse) pdf_Mat = 1.f;
        float J = ds.Jp * p.iJp;
        float pdf_J = J > 1.f ? rcp(J) : J;
        float pdf_Sel = pdf_Mat * pdf_J;
        valid = valid && (rng.next_f32() < pdf_Sel);
        sk.Jp = ds.Jp; sk.iJp = valid ? rcp(ds.Jp) : 0.f;
        sk.px = ds.uvx; sk.py = ds.uvy;
        sk.pdf = valid ? ds.pdf : 0.f;
        sk.pdf_lk = valid ? p.pdf * J * pdf_Sel : 0.f;
        sk.valid = valid;
        bool indirect = valid, direct = valid;
        const bool replace = n_indir * rng.next_f32() < 1.f;
        BsdfResult br = bsdf_zero();                       // masked virtual call: zeros when !valid
        if (valid) br = bsdf_call(*bd.bsdf, sk.wi, wo, rand_1, rand_2);
        direct = direct && br.pdf > 0.f;
        sk.bsdf_val = br.f;
        direct_pdf += direct ? br.pdf : 0.f;
        n_direct += direct ? 1.f : 0.f;
        indirect = indirect && (br.bs.type == bsdf_sample.type);
        if (indirect && replace) bsdf_sample.wo = br.bs.wo;
        n_indir += indirect ? 1.f : 0.f;
        sk.indirect = indirect;
    }
    direct_pdf = fdiv(direct_pdf, n_direct);
}
// MVPT::mis_weights (mvpath_multi.h:464-519)
AM_D void mis_weights(const DScene &s, DSample *samples, uint32_t n, const BsdfData &bd) {
    for (uint32_t k = 0; k < n; ++k) {
        const bool not_first = k > 0; DSample &sk = samples[k];
        float pdfSum = sk.pdf_lk;
        if (not_first) pdfSum += sk.pdf;
        const bool cond = not_first ? sk.valid : bd.reuse;
        float inner = 0.f;
        if (cond && !bd.diffuse) {
            for (uint32_t j = 1; j < n; ++j) {
                if (j == k) continue;
                const DSample &sj = samples[j];
                float pdf_J = min_(sqr(sj.Jp * sk.iJp), 1.f);
                float pdf_Mat = tv_any(s, sj.wo_r, sk.wi, sk.pdfM, bd, sj.valid);
                inner = fmadd(sj.pdf, pdf_J * pdf_Mat, inner);
            }
        } else {
            for (uint32_t j = 1; j < n; ++j) {
                if (j == k) continue;
                const DSample &sj = samples[j];
                float pdf_J = min_(sqr(sj.Jp * sk.iJp), 1.f);
                inner = fmadd(sj.pdf, pdf_J, inner);
            }
            inner = cond ? inner : 0.f;
        }
        pdfSum += inner;
        sk.weight = fdiv(sk.pdf_lk, pdfSum);
    }
}

// MVPT::sample_multi (mvpath_multi.h:133-358).  `ps` arrives from lane_setup (sampler state, primary ray) and leaves as the
// state of the finished suffix path.
template <bool STATS> AM_D void sample_multi(const DScene &s, PathState &ps, DSample *samples, uint32_t n, V2 p_app, StatCtx *st, TapCtx *tap) {
    if (s.max_depth == 0) return;
    DSample &p0 = samples[0];
    const Ray p_ray = ps.ray;
    Hit h = trace_closest(s, p_ray, STATS ? st : nullptr);
    if (STATS) { st->closest++; st->vertices++; }
    const DMesh *mesh;
    SI si = make_si(s, p_ray, h, &mesh);
    if (tap) { if (tap->n_iter < tap->max_hits) { tap->hit_prim[tap->n_iter] = si.prim; tap->hit_t[tap->n_iter] = h.t; } tap->n_iter++; }
    const bool p_hit = si.valid, direct_em = p_hit && mesh->emitter >= 0;
    if (direct_em) p0.result = select3(si.wi.z > 0.f, ld3(s.emitters[mesh->emitter].radiance));      // area.cpp:82-86
    const DBsdf *bsdf = p_hit ? s.bsdfs + mesh->bsdf : nullptr;
    const bool bsdf_smooth = bsdf && (bsdf->flags & L_SMOOTH) != 0;
    bool active_em = p_hit && bsdf_smooth;
    V2 es; es.x = ps.rng.next_f32(); es.y = ps.rng.next_f32();
    DirectionSample ds; ds.pdf = 0.f; ds.d = v3(0.f); V3 em_weight = v3(0.f), wo = v3(0.f);
    if (active_em) {
        if (STATS) st->shadow++;
        ds = emitter_call(s, si, es, &em_weight, STATS ? st : nullptr);
        active_em = ds.pdf != 0.f; wo = si.to_local(ds.d);
    }
    const float rand_1 = ps.rng.next_f32(); V2 rand_2; rand_2.x = ps.rng.next_f32(); rand_2.y = ps.rng.next_f32();
    BsdfResult br = bsdf_zero();
    if (bsdf) br = bsdf_call(*bsdf, si.wi, wo, rand_1, rand_2);
    const V3 bsdf_val = br.f, bsdf_weight = br.weight; float direct_pdf = br.pdf; BSDFSample bsdf_sample = br.bs;
    const bool flag_delta = (bsdf_sample.type & L_DELTA) != 0, flag_null = (bsdf_sample.type & L_NULL) != 0, flag_diff = (bsdf_sample.type & L_DIFFUSE_R) != 0;
    const bool delta = flag_delta || flag_null, p_not_delta = !delta && p_hit;
    const bool reuse = !direct_em && p_not_delta && bsdf_smooth;
    const bool should_reuse = n > 1, should_mis = s.sa_mis && should_reuse;
    if (should_mis) {
        BsdfData bd; bd.bsdf = bsdf; bd.alpha = bsdf ? bsdf_roughness(*bsdf) : 0.f;
        bd.sqr_a = fmsub(bd.alpha, bd.alpha, 1.f); bd.rsqrt_a = rsqrt(bd.alpha); bd.diffuse = flag_diff; bd.reuse = reuse;
        p0.bsdf_val = bsdf_val;
        camera_selection<STATS>(s, ps.rng, samples, n, si, bd, wo, p_app, rand_2, rand_1, p_hit, bsdf_sample, direct_pdf, st);
        mis_weights(s, samples, n, bd);
    } else if (should_reuse) {
        p0.valid = p_hit; const bool p_face = si.wi.z > 0.f;
        for (uint32_t k = 1; k < n; ++k) {
            SurfSample sd = sensors_visible<false, STATS>(s, p_app, si, p_face, samples[k].idx, reuse, st);
            samples[k].px = sd.uvx; samples[k].py = sd.uvy; samples[k].valid = sd.active;
        }
    }
    if (active_em) {
        const V3 emis_mis = em_weight * mis_weight(ds.pdf, direct_pdf);
        if (should_mis) {
            for (uint32_t k = 0; k < n; ++k) if (samples[k].valid) samples[k].result = fmadd(samples[k].bsdf_val, emis_mis, samples[k].result);
        } else p0.result = fmadd(bsdf_val, emis_mis, p0.result);
    }
    const Ray ray = spawn_ray(si.p, si.n, si.to_world(bsdf_sample.wo));
    if (should_mis) {
        float n_indir = 0.f, pdf = 0.f;
        for (uint32_t k = 0; k < n; ++k) {
            DSample &sk = samples[k];
            bool valid = sk.indirect;
            V3 fk; float pk; bsdf_eval_pdf(bsdf, sk.wi, bsdf_sample.wo, valid, fk, pk);
            if (k == 0) {
                fk = p_not_delta ? fk : bsdf_weight; pk = p_not_delta ? pk : bsdf_sample.pdf;
                valid = valid && (pk > 0.f || delta);
            }
            const bool pvalid = pk > 0.f;
            valid = valid && (k == 0 ? (pvalid || delta) : pvalid);
            pk = valid ? pk : 0.f; sk.bsdf_val = select3(valid, fk);
            pdf += pk; n_indir += valid ? 1.f : 0.f;
            sk.indirect = sk.indirect && valid;
        }
        bsdf_sample.pdf = p_not_delta ? fdiv(pdf, n_indir) : bsdf_sample.pdf;
    }
    // suffix (mvpath_multi.h:522-689): the loop of sample_single entered after the first bounce
    ps.ray = ray; ps.throughput = should_mis ? v3(1.f) : bsdf_weight; ps.result = v3(0.f); ps.prev_p = si.p; ps.eta = bsdf_sample.eta;
    ps.prev_bsdf_pdf = bsdf_sample.pdf; ps.depth = p_hit ? 1u : 0u; ps.prev_delta = flag_delta; ps.valid_ray = false;
    ps.active = p_hit;
    if (!should_mis) ps.active = ps.active && hmax(ps.throughput) != 0.f;
    if (s.max_depth <= 1) ps.active = false;
    if (tap) { while (ps.active) path_bounce<true, STATS, true>(s, ps, tap, st); }
    else { while (ps.active) path_bounce<false, STATS, true>(s, ps, nullptr, st); }
    const V3 indirect = ps.result;
    if (should_mis) {
        const float pdfW = p_not_delta ? rcp(bsdf_sample.pdf) : 1.f;
        for (uint32_t k = 0; k < n; ++k) if (samples[k].indirect) samples[k].result = fmadd(samples[k].bsdf_val * pdfW, indirect, samples[k].result);
    } else {
        p0.result = p0.result + indirect; p0.weight = 1.f;
        for (uint32_t k = 1; k < n; ++k) { samples[k].weight = 1.f; samples[k].result = p0.result; }
    }
    p0.weight = p_hit ? p0.weight : 1.f;
    p0.valid = true;
}

// MVPT::render_multisample (mvpath_multi.h:8-78) for one lane: camera group, sample_multi, tile offsets of the reused samples
template <bool STATS> AM_D void lane_multi(const DScene &s, uint32_t seed_value, uint32_t lane, uint32_t px, uint32_t py, DSample *samples, uint32_t n,
                                           StatCtx *st, TapCtx *tap) {
    PathState ps; LaneSetup ls; V2 aperture;
    lane_setup(s, seed_value, lane, px, py, ps, ls, &aperture);
    const uint32_t p_idx = ls.view, max_idx = n * (p_idx / n + 1u);
    for (uint32_t k = 0; k < n; ++k) {
        DSample &d = samples[k];
        d.result = v3(0.f); d.bsdf_val = v3(0.f); d.wi = v3(0.f); d.wo_r = v3(0.f); d.px = d.py = d.weight = d.pdfM = d.pdf = d.pdf_lk = d.Jp = d.iJp = 0.f;
        d.indirect = false; d.valid = false;
        const uint32_t i = p_idx + k; d.idx = i < max_idx ? i : i - n;
    }
    samples[0].px = ls.spx; samples[0].py = ls.spy;
    sample_multi<STATS>(s, ps, samples, n, aperture, st, tap);
    const uint32_t sres_x = s.film_w / s.grid_x, sres_y = s.film_h / s.grid_y;
    for (uint32_t k = 1; k < n; ++k) {
        DSample &d = samples[k];
        uint32_t y = d.idx / s.grid_x, x = d.idx - y * s.grid_x;
        if (s.reverse_x) x = (s.grid_x - 1) - x;
        if (s.reverse_y) y = (s.grid_y - 1) - y;
        d.px += (float) (x * sres_x); d.py += (float) (y * sres_y);
    }
}

// ------------------------------------------------------------------------------------------------
#ifndef AM_REUSE_MIN_CTAS
#define AM_REUSE_MIN_CTAS 16    /* 32 registers per thread: the kernel is latency-bound and occupancy beats spills (4 / 8 / 12 / 16 CTAs per SM: 404 / 490 / 568 / 590 Mpaths/s on the Cornell 8x8 quilt, 45 / 51 / 57 / 66 on 1M triangles) */
#endif
template <int NMAX, bool STATS> __global__ void __launch_bounds__(128, AM_REUSE_MIN_CTAS)
render_reuse_kernel(const __grid_constant__ DScene s, const __grid_constant__ RenderParams p) {
    DSample samples[NMAX];
    const unsigned lane_id = threadIdx.x & 31u;
    const uint32_t n = s.n_reuse;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    for (;;) {
        unsigned long long base = 0;
        if (lane_id == 0) base = atomicAdd(p.counters + CTR_TICKET, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= p.n_lanes) break;
        uint32_t lane, px, py;
        if (!ticket_to_lane(p, s, base + lane_id, &lane, &px, &py)) continue;
        lane_multi<STATS>(s, p.seed_value, lane, px, py, samples, n, &st, nullptr);
        // ImageBlock::put, box path (imageblock.cpp:210-233): values = { weight * ray_weight * result, weight }
        for (uint32_t k = 0; k < n; ++k) {
            const DSample &d = samples[k];
            if (!d.valid) continue;
            const uint32_t ux = (uint32_t) (int32_t) floorf(d.px), uy = (uint32_t) (int32_t) floorf(d.py);
            if (!(ux < s.film_w && uy < s.film_h)) continue;
            const float w = d.weight * 1.f;
            atomicAdd(p.film + (size_t) uy * s.film_w + ux, make_float4(w * d.result.x, w * d.result.y, w * d.result.z, d.weight));
        }
    }
    if (STATS) flush_stats(st, p.counters);
}
// Parity tap: the n records of each lane after the tile offset, 9 floats each (mirrors orc_trace_lanes_multi)
template <int NMAX> __global__ void reuse_tap_kernel(const __grid_constant__ DScene s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t count, float *out) {
    DSample samples[NMAX];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= count) return;
    const uint32_t n = s.n_reuse, lane = lane_begin + i, idx = lane / spp_per_pass, py = idx / s.film_w, px = idx - s.film_w * py;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    lane_multi<false>(s, seed_value, lane, px, py, samples, n, &st, nullptr);
    for (uint32_t k = 0; k < n; ++k) {
        float *o = out + 9 * ((size_t) i * n + k); const DSample &d = samples[k];
        o[0] = d.px; o[1] = d.py; o[2] = d.weight; o[3] = d.result.x; o[4] = d.result.y; o[5] = d.result.z;
        o[6] = d.valid ? 1.f : 0.f; o[7] = d.indirect ? 1.f : 0.f; o[8] = (float) d.idx;
    }
}

// Instantiations are kept few (each is ~0.6 MB of SASS and minutes of ptxas): group sizes <= 8 and <= 64; the counting
// variant and the parity tap exist for 64 only.
template <int NMAX, bool STATS> static void launch_reuse_n(const DScene &s, const RenderParams &p, int sm_count, cudaStream_t st) {
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, render_reuse_kernel<NMAX, STATS>, 128, 0);
    render_reuse_kernel<NMAX, STATS><<<sm_count * (per_sm > 0 ? per_sm : 1), 128, 0, st>>>(s, p);
}
void launch_reuse_pass(const DScene &s, const RenderParams &p, int sm_count, bool stats, cudaStream_t st) {
    if (stats) launch_reuse_n<64, true>(s, p, sm_count, st);
    else if (s.n_reuse <= 8) launch_reuse_n<8, false>(s, p, sm_count, st);
    else launch_reuse_n<64, false>(s, p, sm_count, st);
}
void launch_reuse_tap(const DScene &s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, float *out, cudaStream_t st) {
    if (!n) return;
    reuse_tap_kernel<64><<<(n + 63) / 64, 64, 0, st>>>(s, seed_value, spp_per_pass, lane_begin, n, out);
}

} // namespace amvpt
