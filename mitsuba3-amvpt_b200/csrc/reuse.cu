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
// The `adaptive` fill-in of mvpath_multi.h:52-58, 79-115: every schedule records adapt_mask per ticket and scales the primary weight;
// run_reuse_fill_in compacts the mask in lane order (block counts -> scan -> ranks), and a persistent path kernel traces n_adapt plain
// paths per masked lane with the forked sampler's streams.  `debug` puts the mask instead of the image.
#include "dpath.cuh"

namespace amvpt {

struct DSample {             // mvpath.h:150-167
    V3 result, bsdf_val, wi, wo_r; float px, py, weight, pdfM, pdf, pdf_lk, Jp, iJp; uint32_t idx; bool indirect, valid;
};
struct SurfSample { V3 p, d; float uvx, uvy, pdf, Jp; bool face, active; };
struct BsdfData { const DBsdf *bsdf; float alpha, sqr_a, rsqrt_a; bool diffuse, reuse; };
// Staged schedule: results of the primary-ray stage (hit record) and of the projection stage (one record per candidate
// camera) of this lane, written by reuse_primary_kernel / reuse_project_kernel (below)
struct Staged { const float4 *prim, *proj; uint32_t m, i, nk; };
AM_D Hit staged_hit(const Staged &g) { float4 a = g.prim[g.i]; Hit h; h.t = a.x; h.u = a.y; h.v = a.z; h.tri = __float_as_uint(a.w); return h; }
AM_D SurfSample staged_proj(const Staged &g, uint32_t k, bool active) {
    SurfSample r; r.p = v3(0.f); r.d = v3(0.f); r.uvx = r.uvy = r.pdf = r.Jp = 0.f; r.face = false; r.active = false;
    if (!active) return r;                       // masked virtual call (grid.cpp:331-336): zeros
    const size_t o = (size_t) (k - 1) * g.m + g.i;
    float4 e = g.proj[o], f = g.proj[(size_t) g.nk * g.m + o];
    r.uvx = e.x; r.uvy = e.y; r.pdf = e.z; r.Jp = e.w; r.d = v3(f.x, f.y, f.z);
    const uint32_t fl = __float_as_uint(f.w); r.active = (fl & 1u) != 0; r.face = (fl & 2u) != 0;
    return r;
}

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
template <bool STATS, int STAGED> AM_D void camera_selection(const DScene &s, Pcg32 &rng, DSample *samples, uint32_t n, const SI &si, const BsdfData &bd, V3 wo, V2 p_app,
                                                 V2 rand_2, float rand_1, bool p_hit, BSDFSample &bsdf_sample, float &direct_pdf, StatCtx *st, const Staged *stg) {
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
        SurfSample ds = STAGED >= 1 ? staged_proj(*stg, k, bd.reuse) : sensors_visible<false, STATS>(s, p_app, si, p_face, sk.idx, bd.reuse, st);
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
template <bool STATS, int STAGED> AM_D void sample_multi(const DScene &s, PathState &ps, DSample *samples, uint32_t n, V2 p_app, StatCtx *st, TapCtx *tap, const Staged *stg, bool *mis_out = nullptr, bool *adapt_out = nullptr) {
    if (adapt_out) *adapt_out = false;
    if (s.max_depth == 0) return;
    DSample &p0 = samples[0];
    const Ray p_ray = ps.ray;
    Hit h = STAGED >= 1 ? staged_hit(*stg) : trace_closest(s, p_ray, STATS ? st : nullptr);
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
        if (STAGED == 2) { float4 a = stg->prim[4 * (size_t) stg->m + stg->i], b = stg->prim[5 * (size_t) stg->m + stg->i]; em_weight = v3(a.x, a.y, a.z); ds.pdf = a.w; ds.d = v3(b.x, b.y, b.z); }
        else ds = emitter_call(s, si, es, &em_weight, STATS ? st : nullptr);
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
        camera_selection<STATS, STAGED>(s, ps.rng, samples, n, si, bd, wo, p_app, rand_2, rand_1, p_hit, bsdf_sample, direct_pdf, st, stg);
        mis_weights(s, samples, n, bd);
    } else if (should_reuse) {
        p0.valid = p_hit; const bool p_face = si.wi.z > 0.f;
        for (uint32_t k = 1; k < n; ++k) {
            SurfSample sd = STAGED >= 1 ? staged_proj(*stg, k, reuse) : sensors_visible<false, STATS>(s, p_app, si, p_face, samples[k].idx, reuse, st);
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
        if (adapt_out) *adapt_out = p_hit && !flag_null && n_indir <= 1.f;       // mvpath_multi.h:306-307
    }
    // suffix (mvpath_multi.h:522-689): the loop of sample_single entered after the first bounce
    ps.ray = ray; ps.throughput = should_mis ? v3(1.f) : bsdf_weight; ps.result = v3(0.f); ps.prev_p = si.p; ps.eta = bsdf_sample.eta;
    ps.prev_bsdf_pdf = bsdf_sample.pdf; ps.depth = p_hit ? 1u : 0u; ps.prev_delta = flag_delta; ps.valid_ray = false;
    ps.active = p_hit;
    if (!should_mis) ps.active = ps.active && hmax(ps.throughput) != 0.f;
    if (s.max_depth <= 1) ps.active = false;
    if (STAGED == 2) {
        // pipeline schedule: the suffix path runs in reuse_suffix_kernel and enters the samples in reuse_finish_kernel as
        // result = fmadd(coef, indirect, result) (MIS) or result0 + indirect (no MIS); coef = bsdf_val * pdfW is left in bsdf_val
        if (should_mis) {
            const float pdfW = p_not_delta ? rcp(bsdf_sample.pdf) : 1.f;
            for (uint32_t k = 0; k < n; ++k) if (samples[k].indirect) samples[k].bsdf_val = samples[k].bsdf_val * pdfW;
        } else { p0.weight = 1.f; for (uint32_t k = 1; k < n; ++k) samples[k].weight = 1.f; }
        p0.weight = p_hit ? p0.weight : 1.f; p0.valid = true; *mis_out = should_mis;
        return;
    }
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
template <bool STATS, int STAGED = 0> AM_D void lane_multi(const DScene &s, uint32_t seed_value, uint32_t lane, uint32_t px, uint32_t py, DSample *samples, uint32_t n,
                                           StatCtx *st, TapCtx *tap, const Staged *stg = nullptr, PathState *ps_out = nullptr, bool *mis_out = nullptr, bool *adapt_out = nullptr) {
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
    bool adapt = false;
    sample_multi<STATS, STAGED>(s, ps, samples, n, aperture, st, tap, stg, mis_out, &adapt);
    if (adapt_out) *adapt_out = adapt;
    if (ps_out) *ps_out = ps;
    if (s.reuse_debug) {          // mvpath_multi.h:54-56: put(pos, Float(adapt_mask), alpha, 1) and nothing else
        samples[0].result = v3(adapt ? 1.f : 0.f); samples[0].weight = 1.f; samples[0].valid = true; samples[0].indirect = false;
        for (uint32_t k = 1; k < n; ++k) { samples[k].valid = false; samples[k].indirect = false; }
        if (ps_out) ps_out->active = false;
        if (mis_out) *mis_out = true;         // finish stage: result = record 1 as it stands
    } else if (s.n_adapt && adapt) samples[0].weight = samples[0].weight * s.adapt_w;
    const uint32_t sres_x = s.film_w / s.grid_x, sres_y = s.film_h / s.grid_y;
    for (uint32_t k = 1; k < n; ++k) {
        DSample &d = samples[k];
        uint32_t y = d.idx / s.grid_x, x = d.idx - y * s.grid_x;
        if (s.reverse_x) x = (s.grid_x - 1) - x;
        if (s.reverse_y) y = (s.grid_y - 1) - y;
        d.px += (float) (x * sres_x); d.py += (float) (y * sres_y);
    }
}

// ImageBlock::put, box path (imageblock.cpp:210-233): values = { weight * ray_weight * result, weight }
AM_D void splat_samples(const DScene &s, float4 *film, const DSample *samples, uint32_t n) {
    for (uint32_t k = 0; k < n; ++k) {
        const DSample &d = samples[k];
        if (!d.valid) continue;
        const uint32_t ux = (uint32_t) (int32_t) floorf(d.px), uy = (uint32_t) (int32_t) floorf(d.py);
        if (!(ux < s.film_w && uy < s.film_h)) continue;
        const float w = d.weight * 1.f;
        atomicAdd(film + (size_t) uy * s.film_w + ux, make_float4(w * d.result.x, w * d.result.y, w * d.result.z, d.weight));
    }
}

// ------------------------------------------------------------------------------------------------
// Staged schedule (default): the pass is cut into chunks of m lanes and each chunk runs three small kernels instead of the
// one big persistent kernel.  (1) primary: camera ray + closest hit + the hit point's position / geometric normal / side;
// (2) project: ONE THREAD PER (lane, candidate camera) evaluates sample_surface and traces the visibility ray -- these rays
// are 60 % of all rays of the branch, and here they run with full warps through 10 KB of code; (3) shade: everything else
// (light sample, BSDF calls, camera selection with the sampler's draws in the reference's order, MIS weights, mixture pdf,
// the suffix path, splat), reading the hit and the projection records instead of tracing.  Same arithmetic, same bits.
#ifndef AM_REUSE_SHADE_MIN_CTAS
#define AM_REUSE_SHADE_MIN_CTAS 4     /* the select kernel keeps n 92-byte camera records per thread in local memory: at 12 CTAs/SM they spill to DRAM (ncu: 7.4 GB per 4M-lane launch, long-scoreboard 25); 3 / 4 / 6 / 8 / 12 CTAs: 782 / 783 / 755 / 720 / 698 Mpaths/s */
#endif
#ifndef AM_REUSE_SUFFIX_MIN_CTAS
#define AM_REUSE_SUFFIX_MIN_CTAS 5
#endif
__global__ void __launch_bounds__(128)
reuse_primary_kernel(const __grid_constant__ DScene s, const __grid_constant__ RenderParams p, unsigned long long c0, uint32_t m, float4 *prim) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= m) return;
    uint32_t lane, px, py; if (!ticket_to_lane(p, s, c0 + i, &lane, &px, &py)) return;
    PathState ps; LaneSetup ls; V2 ap;
    lane_setup(s, p.seed_value, lane, px, py, ps, ls, &ap);
    Hit h = intersect<false>(s, ps.ray);
    const DMesh *mesh; SI si = make_si(s, ps.ray, h, &mesh);
    const bool cand = si.valid && mesh->emitter < 0 && (s.bsdfs[mesh->bsdf].flags & L_SMOOTH) != 0;
    const uint32_t fl = ls.view | (cand ? 1u << 16 : 0u) | (si.wi.z > 0.f ? 1u << 17 : 0u);
    prim[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.tri));
    prim[(size_t) m + i] = make_float4(si.p.x, si.p.y, si.p.z, __uint_as_float(fl));
    prim[2 * (size_t) m + i] = make_float4(si.n.x, si.n.y, si.n.z, 0.f);
    prim[3 * (size_t) m + i] = make_float4(ap.x, ap.y, 0.f, 0.f);
    // light sample of the primary vertex (mvpath_multi.h:176-189): the sampler's next two numbers, visibility ray included
    V3 em_weight = v3(0.f); DirectionSample ds; ds.pdf = 0.f; ds.d = v3(0.f);
    if (si.valid && (s.bsdfs[mesh->bsdf].flags & L_SMOOTH) != 0 && s.max_depth != 0) {
        V2 es; es.x = ps.rng.next_f32(); es.y = ps.rng.next_f32();
        ds = sample_emitter_direction<false>(s, si, es, &em_weight, nullptr, nullptr);
    }
    prim[4 * (size_t) m + i] = make_float4(em_weight.x, em_weight.y, em_weight.z, ds.pdf);
    prim[5 * (size_t) m + i] = make_float4(ds.d.x, ds.d.y, ds.d.z, 0.f);
}
__global__ void __launch_bounds__(128)
reuse_project_kernel(const __grid_constant__ DScene s, uint32_t m, uint32_t nk, const float4 *prim, float4 *proj) {
    const size_t j = (size_t) blockIdx.x * blockDim.x + threadIdx.x; if (j >= (size_t) m * nk) return;
    const uint32_t i = (uint32_t) (j % m), k = (uint32_t) (j / m) + 1u, n = nk + 1u;
    const float4 b = prim[(size_t) m + i], c = prim[2 * (size_t) m + i], d = prim[3 * (size_t) m + i];
    const uint32_t fl = __float_as_uint(b.w), view = fl & 0xffffu; const bool cand = (fl >> 16) & 1u, p_face = (fl >> 17) & 1u;
    const uint32_t max_idx = n * (view / n + 1u); uint32_t idx = view + k; idx = idx < max_idx ? idx : idx - n;
    SI si; si.p = v3(b.x, b.y, b.z); si.n = v3(c.x, c.y, c.z);
    SurfSample r = sensors_visible<false, false>(s, v2(d.x, d.y), si, p_face, idx, cand, nullptr);
    proj[j] = make_float4(r.uvx, r.uvy, r.pdf, r.Jp);
    proj[(size_t) nk * m + j] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float((r.active ? 1u : 0u) | (r.face ? 2u : 0u)));
}
template <int NMAX> __global__ void __launch_bounds__(128, AM_REUSE_SHADE_MIN_CTAS)
reuse_shade_kernel(const __grid_constant__ DScene s, const __grid_constant__ RenderParams p, unsigned long long c0, uint32_t m, const float4 *prim, const float4 *proj) {
    DSample samples[NMAX];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= m) return;
    uint32_t lane, px, py; if (!ticket_to_lane(p, s, c0 + i, &lane, &px, &py)) return;
    const uint32_t n = s.n_reuse;
    Staged g; g.prim = prim; g.proj = proj; g.m = m; g.i = i; g.nk = n - 1u;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    bool adapt = false;
    lane_multi<false, 1>(s, p.seed_value, lane, px, py, samples, n, &st, nullptr, &g, nullptr, nullptr, &adapt);
    if (p.adapt_mask) p.adapt_mask[c0 + i] = adapt ? 1 : 0;
    splat_samples(s, p.film, samples, n);
}

// Pipeline schedule (AMVPT_REUSE_STAGED=2, default): the shade stage is cut once more.  (3a) select: as (3) but neither the
// light's visibility ray (done by the primary stage) nor the suffix path: it leaves the suffix's start state and, per
// camera, { film position, weight, flags }, the radiance so far and the coefficient of the suffix radiance; (3b) suffix: the
// persistent path-regeneration loop of render_pass_kernel over the stored states; (3c) finish: one thread per
// (lane, camera): result = fmadd(coef, indirect, result) (or result0 + indirect without MIS) and the 16-byte film RED.
struct PipeBufs { float4 *state; uint4 *rng; float4 *ind; float4 *rec; };    // state: 4 m, rng: m, ind: m, rec: 3 n m
template <int NMAX> __global__ void __launch_bounds__(128, AM_REUSE_SHADE_MIN_CTAS)
reuse_select_kernel(const __grid_constant__ DScene s, const __grid_constant__ RenderParams p, unsigned long long c0, uint32_t m, const float4 *prim, const float4 *proj, PipeBufs b) {
    DSample samples[NMAX];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= m) return;
    uint32_t lane, px, py; if (!ticket_to_lane(p, s, c0 + i, &lane, &px, &py)) return;
    const uint32_t n = s.n_reuse;
    Staged g; g.prim = prim; g.proj = proj; g.m = m; g.i = i; g.nk = n - 1u;
    StatCtx st; PathState ps; bool mis = false; ps.active = false; ps.depth = 0; ps.prev_delta = false;
    ps.ray.o = ps.ray.d = ps.throughput = ps.prev_p = v3(0.f); ps.eta = ps.prev_bsdf_pdf = 0.f; ps.rng.state = ps.rng.inc = 0;
    bool adapt = false;
    lane_multi<false, 2>(s, p.seed_value, lane, px, py, samples, n, &st, nullptr, &g, &ps, &mis, &adapt);
    if (p.adapt_mask) p.adapt_mask[c0 + i] = adapt ? 1 : 0;
    const bool act = ps.active && s.max_depth != 0;
    b.state[i] = make_float4(ps.ray.o.x, ps.ray.o.y, ps.ray.o.z, ps.prev_bsdf_pdf);
    b.state[(size_t) m + i] = make_float4(ps.ray.d.x, ps.ray.d.y, ps.ray.d.z, ps.eta);
    b.state[2 * (size_t) m + i] = make_float4(ps.throughput.x, ps.throughput.y, ps.throughput.z, __uint_as_float(ps.depth | (ps.prev_delta ? 1u << 30 : 0u) | (act ? 1u << 31 : 0u)));
    b.state[3 * (size_t) m + i] = make_float4(ps.prev_p.x, ps.prev_p.y, ps.prev_p.z, __uint_as_float(mis ? 1u : 0u));
    b.rng[i] = make_uint4((uint32_t) ps.rng.state, (uint32_t) (ps.rng.state >> 32), (uint32_t) ps.rng.inc, (uint32_t) (ps.rng.inc >> 32));
    b.ind[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t nm = (size_t) n * m;
    for (uint32_t k = 0; k < n; ++k) {
        const DSample &d = samples[k]; const size_t o = (size_t) k * m + i;
        b.rec[o] = make_float4(d.px, d.py, d.weight, __uint_as_float((d.valid ? 1u : 0u) | (d.indirect ? 2u : 0u)));
        b.rec[nm + o] = make_float4(d.result.x, d.result.y, d.result.z, 0.f);
        b.rec[2 * nm + o] = make_float4(d.bsdf_val.x, d.bsdf_val.y, d.bsdf_val.z, 0.f);
    }
}
__global__ void __launch_bounds__(128, AM_REUSE_SUFFIX_MIN_CTAS)
reuse_suffix_kernel(const __grid_constant__ DScene s, uint32_t m, uint32_t regen_min, PipeBufs b, unsigned long long *ticket_ctr) {
    const unsigned lane_id = threadIdx.x & 31u;
    PathState ps; ps.active = false; uint32_t slot = 0xffffffffu; bool exhausted = false;
    StatCtx st;
    for (;;) {
        const bool idle = !ps.active;
        const unsigned idle_mask = __ballot_sync(0xffffffffu, idle);
        if (__popc(idle_mask) >= (int) regen_min || idle_mask == 0xffffffffu) {
            if (idle) {
                if (slot != 0xffffffffu) { b.ind[slot] = make_float4(ps.result.x, ps.result.y, ps.result.z, 0.f); slot = 0xffffffffu; }
                const unsigned fm = __ballot_sync(idle_mask, !exhausted);
                if (!exhausted) {
                    unsigned long long base = 0; const unsigned leader = __ffs(fm) - 1;
                    if (lane_id == leader) base = atomicAdd(ticket_ctr, (unsigned long long) __popc(fm));
                    base = __shfl_sync(fm, base, leader);
                    const unsigned long long ticket = base + __popc(fm & ((1u << lane_id) - 1u));
                    if (ticket < m) {
                        const uint32_t i = (uint32_t) ticket;
                        const float4 c = b.state[2 * (size_t) m + i]; const uint32_t fl = __float_as_uint(c.w);
                        if (fl >> 31) {                    // paths without a suffix (miss, max_depth, zero throughput) keep ind = 0
                            const float4 a = b.state[i], d = b.state[(size_t) m + i], e = b.state[3 * (size_t) m + i]; const uint4 g = b.rng[i];
                            ps.ray.o = v3(a.x, a.y, a.z); ps.ray.d = v3(d.x, d.y, d.z); ps.ray.maxt = AM_LARGEST; ps.prev_bsdf_pdf = a.w; ps.eta = d.w;
                            ps.throughput = v3(c.x, c.y, c.z); ps.depth = fl & 0x3fffffffu; ps.prev_delta = (fl >> 30) & 1u; ps.prev_p = v3(e.x, e.y, e.z);
                            ps.result = v3(0.f); ps.valid_ray = false; ps.active = true;
                            ps.rng.state = (uint64_t) g.x | ((uint64_t) g.y << 32); ps.rng.inc = (uint64_t) g.z | ((uint64_t) g.w << 32);
                            slot = i;
                        }
                    } else exhausted = true;
                }
            }
            if (__all_sync(0xffffffffu, !ps.active && exhausted)) break;
        }
        if (ps.active) path_bounce<false, false>(s, ps, nullptr, &st);
    }
}
__global__ void __launch_bounds__(256)
reuse_finish_kernel(const __grid_constant__ DScene s, uint32_t m, uint32_t n, PipeBufs b, float4 *film) {
    const size_t j = (size_t) blockIdx.x * blockDim.x + threadIdx.x, nm = (size_t) n * m; if (j >= nm) return;
    const uint32_t i = (uint32_t) (j % m);
    const float4 r0 = b.rec[j]; const uint32_t fl = __float_as_uint(r0.w);
    if (!(fl & 1u)) return;
    const float4 ind = b.ind[i]; const bool mis = __float_as_uint(b.state[3 * (size_t) m + i].w) & 1u;
    V3 result;
    if (mis) {
        const float4 r1 = b.rec[nm + j]; result = v3(r1.x, r1.y, r1.z);
        if (fl & 2u) { const float4 r2 = b.rec[2 * nm + j]; result = fmadd(v3(r2.x, r2.y, r2.z), v3(ind.x, ind.y, ind.z), result); }
    } else { const float4 r1 = b.rec[nm + i]; result = v3(r1.x, r1.y, r1.z) + v3(ind.x, ind.y, ind.z); }     // every camera takes the primary's radiance (mvpath_multi.h:343-350)
    const uint32_t ux = (uint32_t) (int32_t) floorf(r0.x), uy = (uint32_t) (int32_t) floorf(r0.y);
    if (!(ux < s.film_w && uy < s.film_h)) return;
    const float w = r0.z * 1.f;
    atomicAdd(film + (size_t) uy * s.film_w + ux, make_float4(w * result.x, w * result.y, w * result.z, r0.z));
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
        bool adapt = false;
        lane_multi<STATS>(s, p.seed_value, lane, px, py, samples, n, &st, nullptr, nullptr, nullptr, nullptr, &adapt);
        if (p.adapt_mask) p.adapt_mask[base + lane_id] = adapt ? 1 : 0;
        splat_samples(s, p.film, samples, n);
    }
    if (STATS) flush_stats(st, p.counters);
}
// Parity tap: the n records of each lane after the tile offset, 9 floats each (mirrors orc_trace_lanes_multi)
template <int NMAX> __global__ void reuse_tap_kernel(const __grid_constant__ DScene s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t count, float *out) {
    DSample samples[NMAX];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= count) return;
    const uint32_t n = s.n_reuse, lane = lane_begin + i, idx = lane / spp_per_pass, py = idx / s.film_w, px = idx - s.film_w * py;
    StatCtx st; st.closest = st.shadow = st.nodes = st.tris = st.vertices = 0;
    bool adapt = false;
    lane_multi<false>(s, seed_value, lane, px, py, samples, n, &st, nullptr, nullptr, nullptr, nullptr, &adapt);
    for (uint32_t k = 0; k < n; ++k) {
        float *o = out + 9 * ((size_t) i * n + k); const DSample &d = samples[k];
        o[0] = d.px; o[1] = d.py; o[2] = d.weight; o[3] = d.result.x; o[4] = d.result.y; o[5] = d.result.z;
        o[6] = (d.valid ? 1.f : 0.f) + (k == 0 && adapt ? 2.f : 0.f); o[7] = d.indirect ? 1.f : 0.f; o[8] = (float) d.idx;
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
// Staged pass: chunks of m lanes, three launches each.  `buf` holds 4 m + 2 (n - 1) m float4.  Returns the number of launches.
size_t reuse_staged_float4s(uint32_t n, uint32_t m_cap) { return (size_t) m_cap * (6 + 2 * (size_t) (n - 1) + 4 + 1 + 1 + 3 * (size_t) n); }
uint32_t run_reuse_staged(const DScene &s, const RenderParams &p, float4 *buf, uint32_t m_cap, int schedule, int sm_count, cudaStream_t st) {
    const uint32_t n = s.n_reuse, nk = n - 1u; uint32_t launches = 0;
    for (unsigned long long c0 = 0; c0 < p.n_lanes; c0 += m_cap) {
        const uint32_t m = (uint32_t) (p.n_lanes - c0 < m_cap ? p.n_lanes - c0 : m_cap);
        float4 *prim = buf, *proj = buf + 6 * (size_t) m_cap;
        if (schedule == 2) {
            PipeBufs b; b.state = proj + 2 * (size_t) nk * m_cap; b.rng = (uint4 *) (b.state + 4 * (size_t) m_cap); b.ind = (float4 *) (b.rng + m_cap); b.rec = b.ind + m_cap;
            reuse_primary_kernel<<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim);
            const size_t pairs = (size_t) m * nk;
            reuse_project_kernel<<<(unsigned) ((pairs + 127) / 128), 128, 0, st>>>(s, m, nk, prim, proj);
            if (n <= 4) reuse_select_kernel<4><<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim, proj, b);
            else if (n <= 8) reuse_select_kernel<8><<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim, proj, b);
            else if (n <= 16) reuse_select_kernel<16><<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim, proj, b);
            else reuse_select_kernel<64><<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim, proj, b);
            cudaMemsetAsync(p.counters + CTR_TICKET, 0, sizeof(unsigned long long), st);
            int per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, reuse_suffix_kernel, 128, 0);
            reuse_suffix_kernel<<<sm_count * (per_sm > 0 ? per_sm : 1), 128, 0, st>>>(s, m, p.regen_min, b, p.counters + CTR_TICKET);
            const size_t nm = (size_t) n * m;
            reuse_finish_kernel<<<(unsigned) ((nm + 255) / 256), 256, 0, st>>>(s, m, n, b, p.film);
            launches += 5;
            continue;
        }
        reuse_primary_kernel<<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim);
        const size_t pairs = (size_t) m * nk;
        reuse_project_kernel<<<(unsigned) ((pairs + 127) / 128), 128, 0, st>>>(s, m, nk, prim, proj);
        if (n <= 8) reuse_shade_kernel<8><<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim, proj);
        else reuse_shade_kernel<64><<<(m + 127) / 128, 128, 0, st>>>(s, p, c0, m, prim, proj);
        launches += 3;
    }
    return launches;
}
// ------------------------------------------------------------------------------------------------ fill-in pass (mvpath_multi.h:79-115)
// idx = dr::compress(adapt_mask) keeps the lane order; dr::repeat(idx, n_adapt) repeats every index n_adapt times in place.
__global__ void __launch_bounds__(1024) adapt_count_kernel(const uint8_t *mask, unsigned long long n, uint32_t *counts) {
    const unsigned long long t = (unsigned long long) blockIdx.x * 1024ull + threadIdx.x;
    const int c = __syncthreads_count(t < n && mask[t] != 0);
    if (threadIdx.x == 0) counts[blockIdx.x] = (uint32_t) c;
}
__global__ void __launch_bounds__(1024) adapt_scan_kernel(uint32_t *counts, uint32_t n_blocks) {      // exclusive scan in place, counts[n_blocks] = total
    __shared__ uint32_t part[1024];
    const uint32_t t = threadIdx.x, per = (n_blocks + 1023u) / 1024u, b0 = t * per; uint32_t sum = 0;
    for (uint32_t k = 0; k < per && b0 + k < n_blocks; ++k) sum += counts[b0 + k];
    part[t] = sum; __syncthreads();
    for (uint32_t off = 1; off < 1024u; off <<= 1) { const uint32_t x = t >= off ? part[t - off] : 0u; __syncthreads(); part[t] += x; __syncthreads(); }
    uint32_t run = part[t] - sum;
    for (uint32_t k = 0; k < per && b0 + k < n_blocks; ++k) { const uint32_t c = counts[b0 + k]; counts[b0 + k] = run; run += c; }
    if (t == 1023u) counts[n_blocks] = part[1023];
}
__global__ void __launch_bounds__(1024) adapt_write_kernel(const uint8_t *mask, unsigned long long n, const uint32_t *offsets, uint32_t *compact) {
    __shared__ uint32_t warp_base[32];
    const unsigned long long t = (unsigned long long) blockIdx.x * 1024ull + threadIdx.x;
    const bool m = t < n && mask[t] != 0; const unsigned lane_id = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned b = __ballot_sync(0xffffffffu, m);
    if (lane_id == 0) warp_base[warp] = (uint32_t) __popc(b);
    __syncthreads();
    if (warp == 0) { uint32_t v = warp_base[lane_id], inc = v; for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o); if ((int) lane_id >= o) inc += x; } warp_base[lane_id] = inc - v; }
    __syncthreads();
    if (m) compact[offsets[blockIdx.x] + warp_base[warp] + (uint32_t) __popc(b & ((1u << lane_id) - 1u))] = (uint32_t) t;      // tickets of one pass fit 32 bits (mvpath.cpp:133-147)
}
// One plain path per entry j of repeat(compress(mask), n_adapt): the primary sample's own camera ray (same jitter / aperture / time:
// regenerated from the primary's stream), the sampler forked and seeded (wavefront2, wavefront2), put with weight adapt_w.
__global__ void __launch_bounds__(128, AM_REUSE_SUFFIX_MIN_CTAS)
reuse_fill_kernel(const __grid_constant__ DScene s, const __grid_constant__ RenderParams p, const uint32_t *compact, unsigned long long wavefront2, uint32_t seed2, unsigned long long *ticket_ctr) {
    const unsigned lane_id = threadIdx.x & 31u;
    PathState ps; LaneSetup ls; ps.active = false; bool has_result = false, exhausted = false; StatCtx st;
    for (;;) {
        const bool idle = !ps.active;
        const unsigned idle_mask = __ballot_sync(0xffffffffu, idle);
        if (__popc(idle_mask) >= (int) p.regen_min || idle_mask == 0xffffffffu) {
            if (idle) {
                if (has_result && ls.in_film) {
                    const V3 L = ps.valid_ray ? ps.result : v3(0.f); const float w = s.adapt_w;
                    atomicAdd(p.film + ls.pixel, make_float4(w * L.x, w * L.y, w * L.z, w));
                }
                has_result = false;
                const unsigned fm = __ballot_sync(idle_mask, !exhausted);
                if (!exhausted) {
                    unsigned long long base = 0; const unsigned leader = __ffs(fm) - 1;
                    if (lane_id == leader) base = atomicAdd(ticket_ctr, (unsigned long long) __popc(fm));
                    base = __shfl_sync(fm, base, leader);
                    const unsigned long long j = base + __popc(fm & ((1u << lane_id) - 1u));
                    if (j < wavefront2) {
                        uint32_t lane, px, py;
                        if (ticket_to_lane(p, s, (unsigned long long) compact[j / s.n_adapt], &lane, &px, &py)) {
                            lane_setup(s, p.seed_value, lane, px, py, ps, ls);
                            ps.rng = sampler_seed(seed2, (uint32_t) j);
                            has_result = true; if (s.max_depth == 0) ps.active = false;
                        }
                    } else exhausted = true;
                }
            }
            if (__all_sync(0xffffffffu, !ps.active && !has_result && exhausted)) break;
        }
        if (ps.active) path_bounce<false, false>(s, ps, nullptr, &st);
    }
}
uint32_t run_reuse_fill_in(const DScene &s, const RenderParams &p, uint32_t base_seed, uint32_t *scratch, uint32_t **compact, size_t *compact_cap, int sm_count, cudaStream_t st, unsigned long long *n_fill) {
    const unsigned long long n = p.n_lanes; if (!n || !s.n_adapt || !p.adapt_mask) return 0;
    const uint32_t n_blocks = (uint32_t) ((n + 1023ull) / 1024ull);
    adapt_count_kernel<<<n_blocks, 1024, 0, st>>>(p.adapt_mask, n, scratch);
    adapt_scan_kernel<<<1, 1024, 0, st>>>(scratch, n_blocks);
    uint32_t total = 0; cudaMemcpyAsync(&total, scratch + n_blocks, sizeof(uint32_t), cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st);
    if (!total) return 2;
    if (*compact_cap < total) { if (*compact) cudaFree(*compact); *compact = nullptr; *compact_cap = 0; const size_t cap = (size_t) total + total / 4 + 1024;
        if (cudaMalloc((void **) compact, cap * sizeof(uint32_t)) != cudaSuccess) return 0; *compact_cap = cap; }
    adapt_write_kernel<<<n_blocks, 1024, 0, st>>>(p.adapt_mask, n, scratch, *compact);
    const unsigned long long wavefront2 = (unsigned long long) total * s.n_adapt;
    cudaMemsetAsync(p.counters + CTR_TICKET, 0, sizeof(unsigned long long), st);
    int per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, reuse_fill_kernel, 128, 0); if (per_sm < 1) per_sm = 1;
    unsigned grid = (unsigned) (sm_count * per_sm); const unsigned long long need = (wavefront2 + 127ull) / 128ull; if (need < grid) grid = (unsigned) need;
    reuse_fill_kernel<<<grid, 128, 0, st>>>(s, p, *compact, wavefront2, base_seed + (uint32_t) wavefront2, p.counters + CTR_TICKET);
    if (n_fill) *n_fill += wavefront2;
    return 4;
}
void launch_reuse_tap(const DScene &s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, float *out, cudaStream_t st) {
    if (!n) return;
    reuse_tap_kernel<64><<<(n + 63) / 64, 64, 0, st>>>(s, seed_value, spp_per_pass, lane_begin, n, out);
}

} // namespace amvpt
