// oracle/o_reuse.h -- TEST INFRASTRUCTURE ONLY (included by amvpt_oracle.cpp inside namespace orc).
// CPU restatement of the cross-view sample-reuse branch of the AMVPT integrator:
//   MVPT::render_multisample / sample_multi / camera_selection / mis_weights / sample_suffix
//   (src/integrators/mvpath_multi.h:8-689), sensors_visible / tv_pdf / tv_pdf_fast (src/integrators/mvpath.h:243-293),
//   PerspectiveCamera / ThinLensCamera ::sample_surface (src/sensors/perspective.cpp:327-385, thinlens.cpp:358-418)
//   through GridSensor::sample_surface (grid.cpp:331-336), BSDF::eval_roughness (diffuse.cpp:186, roughconductor.cpp:518,
//   roughdielectric.cpp:607, roughplastic.cpp:508, twosided.cpp:284, bsdf.h:596).
// Semantics are those of the JIT variants, lane by lane: `dr::any_or<true>` / `dr::none_or<false>` never skip a statement,
// a masked virtual call returns zeros for a masked-out lane, random numbers are drawn by every lane that reaches the draw.
// PARITY UNPINNED: the reference holds no test for this branch; this restatement is the only checker.
// The `adaptive` fill-in (mvpath_multi.h:52-58, 79-115) is restated in render_lane_multi (mask, primary weight, `debug` output) and in
// orc_render_rows (amvpt_oracle.cpp: dr::compress / dr::repeat order, the forked sampler seeded (wavefront, wavefront), the weighted put).

struct SampleData {          // mvpath.h:150-167
    V3 result, bsdf_val, wi, wo_r; V2 pos{ 0.f, 0.f };
    float weight = 0, pdfM = 0, pdf = 0, pdf_lk = 0, Jp = 0, iJp = 0; uint32_t idx = 0; bool indirect = false, valid = false;
};
struct SurfSample { V3 p, d; V2 uv{ 0.f, 0.f }; float pdf = 0, dist = 0, Jp = 0; bool face = false, active = false; };
struct BsdfData { const Bsdf *bsdf = nullptr; float alpha = 0, sqr_a = 0, rsqrt_a = 0; bool diffuse = false, reuse = false; };

// perspective.cpp:327-385 / thinlens.cpp:358-418 behind the masked virtual call of grid.cpp:331-336
static SurfSample sensor_sample_surface(const Scene &s, uint32_t idx, V3 it_p, V3 it_n, V2 sample, bool active) {
    SurfSample r; if (!active) return r;
    const SubSensor &c = s.sensors[idx];
    V3 ref_p = xf_point_affine(c.w2c, it_p);
    active = active && (ref_p.z >= c.near_clip) && (ref_p.z <= c.far_clip);
    float resx = (float) (s.film_w / s.grid_x), resy = (float) (s.film_h / s.grid_y), inv_dist, ictf3;
    V3 origin(0.f);
    if (!s.thinlens) {
        V3 ss = xf_point(c.c2s, ref_p);
        r.uv = { ss.x - 0.f, ss.y - 0.f };
        active = active && (r.uv.x >= 0.f) && (r.uv.x <= 1.f) && (r.uv.y >= 0.f) && (r.uv.y <= 1.f);
        r.uv.x *= resx; r.uv.y *= resy;
        float dist = norm(ref_p); inv_dist = rcp(dist);
        float ctf = ref_p.z;                    // Frame3f::cos_theta of the un-normalised camera-space vector (:356-361)
        active = active && ctf > 0.f;
        float ictf = rcp(ctf); ictf3 = ictf * ictf * ictf;
        r.pdf = c.normalization * ictf3; r.dist = dist;
    } else {
        V2 tmp = square_to_uniform_disk_concentric(sample);
        origin = V3(tmp.x * c.aperture_radius, tmp.y * c.aperture_radius, 0.f);
        V3 local_d = ref_p - origin; float dist = norm(local_d); inv_dist = rcp(dist); local_d = local_d * inv_dist;
        float ctf = local_d.z, ictf = rcp(ctf); ictf3 = ictf * ictf * ictf;
        V3 film_plane = origin * (1.f / c.focus_distance) + (local_d / local_d.z);
        V3 scr = xf_point_affine(c.c2s, film_plane);
        active = active && (scr.x >= 0.f) && (scr.y >= 0.f) && (scr.x <= 1.f) && (scr.y <= 1.f);
        float pdf_lens = rcp(sqr(c.aperture_radius) * Pi), pdf_film = c.normalization * ictf3;
        r.pdf = pdf_lens * pdf_film; r.uv = { scr.x * resx, scr.y * resy }; r.dist = dist;
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
template <bool Primary> static SurfSample sensors_visible(const Scene &s, V2 ap, const SI &si, bool prim_face, uint32_t idx, bool active, bool brute, TraverseStats *st) {
    SurfSample r = sensor_sample_surface(s, idx, si.p, si.n, ap, active);
    if (!Primary) {
        r.active = r.active && (r.face == prim_face) && r.Jp > 0.f;
        if (r.active) r.active = !scene_ray_test(s, spawn_ray_to(si.p, si.n, r.p), brute, st);   // the test of a masked-out lane is unobservable
    }
    return r;
}
// BSDF::eval_roughness
static inline float bsdf_roughness(const Bsdf &b, bool active) {
    switch (b.type) {
        case BSDF_DIFFUSE: return active ? 1.f : 0.f;
        case BSDF_ROUGHCONDUCTOR: case BSDF_ROUGHDIELECTRIC: { float a = std::sqrt(0.5f * (sqr(b.alpha_u) + sqr(b.alpha_v))); return active ? a : 0.f; }
        case BSDF_ROUGHPLASTIC: return active ? b.alpha_u : 0.f;
    }
    return 0.f;
}
// BSDF::eval_pdf behind a masked virtual call (TwoSidedBRDF::eval_pdf twosided.cpp:218-260, single nested BRDF)
static inline void bsdf_eval_pdf(const Bsdf *b, V3 wi, V3 wo, bool active, V3 &f, float &pdf) {
    f = V3(0.f); pdf = 0.f; if (!active || !b) return;
    if (b->twosided) { wo.z = mulsign(wo.z, wi.z); wi.z = fabs_(wi.z); }
    switch (b->type) {
        case BSDF_DIFFUSE: diffuse_eval_pdf(*b, wi, wo, f, pdf); break;
        case BSDF_ROUGHCONDUCTOR: roughconductor_eval_pdf(*b, wi, wo, f, pdf); break;
        case BSDF_ROUGHDIELECTRIC: roughdielectric_eval_pdf(*b, wi, wo, f, pdf); break;
        case BSDF_ROUGHPLASTIC: roughplastic_eval_pdf(*b, wi, wo, f, pdf); break;
        default: break;     // delta BSDFs: eval = pdf = 0
    }
}
// BSDF::pdf with BSDFContext(Radiance, Glossy) (mvpath_multi.h:371, mvpath.h:277): diffuse lobes are disabled
static inline float bsdf_pdf_glossy(const Bsdf *b, V3 wi, V3 wo, bool active) {
    if (!active || !b) return 0.f;
    if (b->twosided) { wo.z = mulsign(wo.z, wi.z); wi.z = fabs_(wi.z); }
    V3 f; float pdf = 0.f;
    switch (b->type) {
        case BSDF_ROUGHCONDUCTOR: roughconductor_eval_pdf(*b, wi, wo, f, pdf); return pdf;      // roughconductor.cpp:390-425
        case BSDF_ROUGHDIELECTRIC: roughdielectric_eval_pdf(*b, wi, wo, f, pdf); return pdf;    // roughdielectric.cpp:433-496
        case BSDF_ROUGHPLASTIC: {                                                               // roughplastic.cpp:380-422, specular only
            V3 H = normalize(wo + wi); Microfacet d((MfType) b->distr, b->alpha_u, b->alpha_u);
            float result = d.eval(H) * d.smith_g1(wi, H) / (4.f * wi.z);
            result *= 1.f; result += 0.f * square_to_cosine_hemisphere_pdf(wo);
            return result;
        }
        default: return 0.f;
    }
}
// mvpath.h:258-271
static inline float tv_pdf_fast(V3 wo_l, V3 wi_k, float p_k, const BsdfData &b, bool active) {
    float p_l = sqr(normalize(wi_k + wo_l).z);
    float N = fmadd(b.sqr_a, fmax_(p_k, p_l), 1.f), D = fmadd(b.sqr_a, fmin_(p_k, p_l), 1.f);
    float q = sqr(N * rcp(D));
    float p = fmadd(q - 1.f, b.rsqrt_a, 1.f);
    p = sqr(fmax_(p, 0.f)); p = lerp(p, q, b.alpha);
    return active ? p : 0.f;
}
// mvpath.h:277-293
static inline float tv_pdf(V3 wo_l, V3 wi_k, float p_k, const BsdfData &b, bool active) {
    active = active && p_k > 0.f;
    float p_l = bsdf_pdf_glossy(b.bsdf, wi_k, wo_l, active);
    active = active && p_l > 0.f;
    float p_max = fmax_(p_l, p_k), p_min = fmin_(p_l, p_k);
    float q = p_min * rcp(p_max);
    float p = fmadd(q - 1.f, b.rsqrt_a, 1.f);
    p = sqr(fmax_(p, 0.f)); p = lerp(p, q, b.alpha);
    return active ? p : 0.f;
}
static inline float tv_any(const Scene &s, V3 wo_l, V3 wi_k, float p_k, const BsdfData &b, bool active) {
    return s.fast_mis ? tv_pdf_fast(wo_l, wi_k, p_k, b, active) : tv_pdf(wo_l, wi_k, p_k, b, active);
}
static inline float pdf_material(const Scene &s, const BsdfData &b, V3 wi, V3 wo_r, bool active) {
    return s.fast_mis ? sqr(normalize(wi + wo_r).z) : bsdf_pdf_glossy(b.bsdf, wi, wo_r, active);
}

// MVPT::sample_suffix (mvpath_multi.h:522-689): the loop of sample_single entered at the first bounce
static V3 sample_suffix(const Scene &s, Pcg32 &rng, Ray ray, const SI &psi, V3 throughput, float bsdf_pdf, float eta, uint32_t depth,
                        bool bsdf_delta, bool active, bool brute, bool *valid_out, Tap *tap, TraverseStats *st) {
    *valid_out = false;
    if (s.max_depth <= 1) return V3(0.f);
    V3 result(0.f), prev_p = psi.p; float prev_bsdf_pdf = bsdf_pdf; bool prev_bsdf_delta = bsdf_delta, valid_ray = false;
    while (active) {
        Hit h = scene_intersect(s, ray, brute, st);
        if (tap && tap->n_iter < tap->max_hits) { tap->hit_prim[tap->n_iter] = h.prim; tap->hit_t[tap->n_iter] = h.t; }
        if (tap) tap->n_iter++;
        SI si = make_si(s, ray, h);
        if (si.valid && si.mesh->emitter >= 0) {
            V3 rel = si.p - prev_p; float dist = norm(rel); V3 d = rel / dist;
            float em_pdf = 0.f;
            if (!prev_bsdf_delta) em_pdf = pdf_emitter_direction(s, *si.mesh, d, si.sh_n, dist, true);
            float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
            V3 le = mask3(s.emitters[si.mesh->emitter].radiance, si.wi.z > 0.f && prev_bsdf_pdf > 0.f);
            result = fmadd(throughput, le * mis_bsdf, result);
        }
        bool active_next = (depth + 1 < s.max_depth) && si.valid;
        if (!active_next) { valid_ray = valid_ray || si.valid; break; }
        const Bsdf &bsdf = s.bsdfs[si.mesh->bsdf];
        bool active_em = (bsdf.flags() & L_SMOOTH) != 0;
        V2 es; es.x = rng.next_f32(); es.y = rng.next_f32();
        DirectionSample ds; V3 em_weight(0.f), wo(0.f);
        if (active_em) { ds = sample_emitter_direction(s, si, es, &em_weight, brute, st); active_em = ds.pdf != 0.f; wo = si.to_local(ds.d); }
        float s1 = rng.next_f32(); V2 s2; s2.x = rng.next_f32(); s2.y = rng.next_f32();
        BsdfResult br = bsdf_eval_pdf_sample(bsdf, si.wi, wo, s1, s2);
        if (active_em) result = fmadd(throughput, br.f * em_weight * mis_weight(ds.pdf, br.pdf), result);
        ray = spawn_ray(si.p, si.n, si.to_world(br.bs.wo));
        throughput = throughput * br.weight; eta *= br.bs.eta;
        valid_ray = true;
        prev_p = si.p; prev_bsdf_pdf = br.bs.pdf; prev_bsdf_delta = (br.bs.type & L_DELTA) != 0;
        depth += 1;
        float tmax = max3(throughput), rr_prob = fmin_(tmax * sqr(eta), .95f);
        bool rr_active = depth >= s.rr_depth, rr_continue = rng.next_f32() < rr_prob;
        if (rr_active) throughput = throughput * rcp(rr_prob);
        active = (!rr_active || rr_continue) && (tmax != 0.f);
    }
    *valid_out = valid_ray;
    return result;
}

// MVPT::camera_selection (mvpath_multi.h:360-462)
static void camera_selection(const Scene &s, Pcg32 &rng, SampleData *samples, uint32_t n, const SI &si, const BsdfData &bd, V3 wo, V2 p_app,
                             V2 rand_2, float rand_1, bool p_hit, BSDFSample &bsdf_sample, float &direct_pdf, bool brute, TraverseStats *st) {
    SampleData &p = samples[0];
    bool p_face = si.wi.z > 0.f;
    SurfSample pds = sensors_visible<true>(s, p_app, si, p_face, p.idx, p_hit, brute, st);
    p.pdf = pds.pdf; p.pdf_lk = pds.pdf; p.Jp = pds.Jp; p.iJp = p_hit ? rcp(pds.Jp) : 0.f;
    p.wi = si.wi; p.wo_r = reflect(si.wi);
    p.valid = p_hit; p.indirect = p_hit;
    p.pdfM = pdf_material(s, bd, p.wi, p.wo_r, p_hit);
    if (bd.diffuse) p.pdfM = 1.f;
    float n_direct = 1.f, n_indir = 2.f;
    for (uint32_t k = 1; k < n; ++k) {
        SampleData &sk = samples[k];
        SurfSample ds = sensors_visible<false>(s, p_app, si, p_face, sk.idx, bd.reuse, brute, st);
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
        sk.pos = ds.uv;
        sk.pdf = valid ? ds.pdf : 0.f;
        sk.pdf_lk = valid ? p.pdf * J * pdf_Sel : 0.f;
        sk.valid = valid;
        bool indirect = valid, direct = valid;
        bool replace = n_indir * rng.next_f32() < 1.f;
        BsdfResult br; br.f = V3(0.f); br.weight = V3(0.f); br.bs.wo = V3(0.f);      // masked virtual call: zeros when !valid
        if (valid) br = bsdf_eval_pdf_sample(*bd.bsdf, sk.wi, wo, rand_1, rand_2);
        direct = direct && br.pdf > 0.f;
        sk.bsdf_val = br.f;
        direct_pdf += direct ? br.pdf : 0.f;
        n_direct += direct ? 1.f : 0.f;
        indirect = indirect && (br.bs.type == bsdf_sample.type);
        if (indirect && replace) bsdf_sample.wo = br.bs.wo;
        n_indir += indirect ? 1.f : 0.f;
        sk.indirect = indirect;
    }
    direct_pdf /= n_direct;
}
// MVPT::mis_weights (mvpath_multi.h:464-519)
static void mis_weights(const Scene &s, SampleData *samples, uint32_t n, const BsdfData &bd) {
    for (uint32_t k = 0; k < n; ++k) {
        bool not_first = k > 0; SampleData &sk = samples[k];
        float pdfSum = sk.pdf_lk;
        if (not_first) pdfSum += sk.pdf;
        bool cond = not_first ? sk.valid : bd.reuse;
        float inner = 0.f;
        if (cond && !bd.diffuse) {
            for (uint32_t j = 1; j < n; ++j) {
                if (j == k) continue;
                const SampleData &sj = samples[j];
                float pdf_J = fmin_(sqr(sj.Jp * sk.iJp), 1.f);
                float pdf_Mat = tv_any(s, sj.wo_r, sk.wi, sk.pdfM, bd, sj.valid);
                inner = fmadd(sj.pdf, pdf_J * pdf_Mat, inner);
            }
        } else {
            for (uint32_t j = 1; j < n; ++j) {
                if (j == k) continue;
                const SampleData &sj = samples[j];
                float pdf_J = fmin_(sqr(sj.Jp * sk.iJp), 1.f);
                inner = fmadd(sj.pdf, pdf_J, inner);
            }
            inner = cond ? inner : 0.f;
        }
        pdfSum += inner;
        sk.weight = sk.pdf_lk / pdfSum;
    }
}

// MVPT::sample_multi (mvpath_multi.h:133-358)
static bool sample_multi(const Scene &s, Pcg32 &rng, SampleData *samples, uint32_t n, const Ray &p_ray, V2 p_app, bool brute, Tap *tap, TraverseStats *st, bool *adapt_mask = nullptr) {
    if (adapt_mask) *adapt_mask = false;
    if (s.max_depth == 0) return false;
    SampleData &ps = samples[0];
    bool valid_ray = false;                                // no environment emitter on this path
    Hit h = scene_intersect(s, p_ray, brute, st);
    if (tap && tap->n_iter < tap->max_hits) { tap->hit_prim[tap->n_iter] = h.prim; tap->hit_t[tap->n_iter] = h.t; }
    if (tap) tap->n_iter++;
    SI si = make_si(s, p_ray, h);
    bool p_hit = si.valid, direct_em = p_hit && si.mesh->emitter >= 0;
    if (direct_em) ps.result = mask3(s.emitters[si.mesh->emitter].radiance, si.wi.z > 0.f);     // area.cpp:82-86
    const Bsdf *bsdf = p_hit ? &s.bsdfs[si.mesh->bsdf] : nullptr;
    bool bsdf_smooth = bsdf && (bsdf->flags() & L_SMOOTH) != 0;
    bool active_em = p_hit && bsdf_smooth;
    V2 es; es.x = rng.next_f32(); es.y = rng.next_f32();
    DirectionSample ds; V3 em_weight(0.f), wo(0.f);
    if (active_em) { ds = sample_emitter_direction(s, si, es, &em_weight, brute, st); active_em = ds.pdf != 0.f; wo = si.to_local(ds.d); }
    float rand_1 = rng.next_f32(); V2 rand_2; rand_2.x = rng.next_f32(); rand_2.y = rng.next_f32();
    BsdfResult br; br.f = V3(0.f); br.weight = V3(0.f); br.bs.wo = V3(0.f);
    if (bsdf) br = bsdf_eval_pdf_sample(*bsdf, si.wi, wo, rand_1, rand_2);
    V3 bsdf_val = br.f, bsdf_weight = br.weight; float direct_pdf = br.pdf; BSDFSample bsdf_sample = br.bs;
    bool flag_delta = (bsdf_sample.type & L_DELTA) != 0, flag_null = (bsdf_sample.type & L_NULL) != 0, flag_diff = (bsdf_sample.type & L_DIFFUSE_R) != 0;
    bool delta = flag_delta || flag_null, not_delta = !delta, p_not_delta = not_delta && p_hit;
    bool reuse = !direct_em && p_not_delta && bsdf_smooth;
    bool should_reuse = n > 1, should_mis = s.sa_mis && should_reuse;
    BsdfData bd;
    if (should_mis) {
        bd.bsdf = bsdf; bd.alpha = bsdf ? bsdf_roughness(*bsdf, true) : 0.f;
        bd.sqr_a = fmsub(bd.alpha, bd.alpha, 1.f); bd.rsqrt_a = rsqrt(bd.alpha); bd.diffuse = flag_diff; bd.reuse = reuse;
        ps.bsdf_val = bsdf_val;
        camera_selection(s, rng, samples, n, si, bd, wo, p_app, rand_2, rand_1, p_hit, bsdf_sample, direct_pdf, brute, st);
        mis_weights(s, samples, n, bd);
    } else if (should_reuse) {
        ps.valid = p_hit; bool p_face = si.wi.z > 0.f;
        for (uint32_t k = 1; k < n; ++k) {
            SurfSample sd = sensors_visible<false>(s, p_app, si, p_face, samples[k].idx, reuse, brute, st);
            samples[k].pos = sd.uv; samples[k].valid = sd.active;
        }
    }
    if (active_em) {
        V3 emis_mis = em_weight * mis_weight(ds.pdf, direct_pdf);
        if (should_mis) {
            for (uint32_t k = 0; k < n; ++k) if (samples[k].valid) samples[k].result = fmadd(samples[k].bsdf_val, emis_mis, samples[k].result);
        } else ps.result = fmadd(bsdf_val, emis_mis, ps.result);
    }
    Ray ray = spawn_ray(si.p, si.n, si.to_world(bsdf_sample.wo));
    if (should_mis) {
        float n_indir = 0.f, pdf = 0.f;
        for (uint32_t k = 0; k < n; ++k) {
            SampleData &sk = samples[k];
            bool valid = sk.indirect;
            V3 fk; float pk; bsdf_eval_pdf(bsdf, sk.wi, bsdf_sample.wo, valid, fk, pk);
            if (k == 0) {
                fk = p_not_delta ? fk : bsdf_weight; pk = p_not_delta ? pk : bsdf_sample.pdf;
                valid = valid && (pk > 0.f || delta);
            }
            bool pvalid = pk > 0.f;
            valid = valid && (k == 0 ? (pvalid || delta) : pvalid);
            pk = valid ? pk : 0.f; sk.bsdf_val = mask3(fk, valid);
            pdf += pk; n_indir += valid ? 1.f : 0.f;
            sk.indirect = sk.indirect && valid;
        }
        bsdf_sample.pdf = p_not_delta ? pdf / n_indir : bsdf_sample.pdf;
        if (adapt_mask) *adapt_mask = p_hit && !flag_null && n_indir <= 1.f;      // mvpath_multi.h:306-307: only the primary camera collected the sample
    }
    V3 throughput = should_mis ? V3(1.f) : bsdf_weight;
    valid_ray = valid_ray || (p_hit && !flag_null);
    bool active = p_hit;
    if (!should_mis) active = active && max3(throughput) != 0.f;
    bool valid = false;
    V3 indirect = sample_suffix(s, rng, ray, si, throughput, bsdf_sample.pdf, bsdf_sample.eta, p_hit ? 1u : 0u, flag_delta, active, brute, &valid, tap, st);
    valid_ray = valid_ray || valid;
    if (should_mis) {
        float pdfW = p_not_delta ? rcp(bsdf_sample.pdf) : 1.f;
        for (uint32_t k = 0; k < n; ++k) if (samples[k].indirect) samples[k].result = fmadd(samples[k].bsdf_val * pdfW, indirect, samples[k].result);
    } else {
        ps.result = ps.result + indirect; ps.weight = 1.f;
        for (uint32_t k = 1; k < n; ++k) { samples[k].weight = 1.f; samples[k].result = ps.result; }
    }
    ps.weight = p_hit ? ps.weight : 1.f;
    ps.valid = true;
    return valid_ray;
}

// MVPT::render_multisample (mvpath_multi.h:8-78) for one lane; the n splats go through `put(x, y, rgb, weight)`
constexpr uint32_t MaxReuse = 64;
template <typename Put> static void render_lane_multi(const Scene &s, uint32_t seed_value, uint32_t lane, uint32_t spp_per_pass, uint32_t n, bool brute, Tap *tap,
                                                     TraverseStats *st, Put put, SampleData *out = nullptr, bool *adapt_out = nullptr) {
    Pcg32 rng = sampler_seed(seed_value, lane);
    uint32_t idx = lane / spp_per_pass, py = idx / s.film_w, px = idx - s.film_w * py;
    float sx = 1.f / (float) s.film_w, sy = 1.f / (float) s.film_h;
    float jx = rng.next_f32(), jy = rng.next_f32();
    float spx = (float) (int32_t) px + jx, spy = (float) (int32_t) py + jy;
    V2 adjusted = { fmadd(spx, sx, -0.f * sx), fmadd(spy, sy, -0.f * sy) };
    V2 aperture = { .5f, .5f };
    if (s.thinlens) { aperture.x = rng.next_f32(); aperture.y = rng.next_f32(); }
    if (s.shutter_draw) (void) rng.next_f32();       // mvpath_multi.h:21-23
    uint32_t p_idx; Ray ray = sensor_sample_ray(s, adjusted, aperture, &p_idx);
    SampleData samples[MaxReuse];
    uint32_t max_idx = n * (p_idx / n + 1u);
    for (uint32_t k = 0; k < n; ++k) { uint32_t i = p_idx + k; samples[k].idx = i < max_idx ? i : i - n; }
    samples[0].pos = { spx, spy };
    bool adapt_mask = false;
    sample_multi(s, rng, samples, n, ray, aperture, brute, tap, st, &adapt_mask);
    if (adapt_out) *adapt_out = adapt_mask;
    // mvpath_multi.h:52-58: the fill-in samples are averaged with the primary one; `debug` shows the mask instead of the image
    const uint32_t n_adapt = std::min(s.adaptive, n - 1u); const float adapt_w = 1.f / (float) (n_adapt + 1u);
    if (s.debug) {
        if (out) for (uint32_t i = 0; i < n; ++i) { out[i] = samples[i]; if (i) out[i].valid = false; else { out[0].result = V3(adapt_mask ? 1.f : 0.f); out[0].weight = 1.f; out[0].valid = true; } }
        int32_t fx = (int32_t) std::floor(spx), fy = (int32_t) std::floor(spy); uint32_t ux = (uint32_t) fx, uy = (uint32_t) fy;
        if (ux < s.film_w && uy < s.film_h) put(ux, uy, V3(adapt_mask ? 1.f : 0.f), 1.f);
        return;
    } else if (n_adapt && adapt_mask) samples[0].weight = samples[0].weight * adapt_w;
    uint32_t sres_x = s.film_w / s.grid_x, sres_y = s.film_h / s.grid_y;
    for (uint32_t i = 0; i < n; ++i) {
        SampleData &sa = samples[i];
        if (i > 0) {
            uint32_t y = sa.idx / s.grid_x, x = sa.idx - y * s.grid_x;
            if (s.reverse_x) x = (s.grid_x - 1) - x;
            if (s.reverse_y) y = (s.grid_y - 1) - y;
            sa.pos.x += (float) (x * sres_x); sa.pos.y += (float) (y * sres_y);
        }
        if (out) out[i] = sa;
        if (!sa.valid) continue;
        // ImageBlock::put box path (imageblock.cpp:210-233): values = {weight * ray_weight * result, weight}
        int32_t fx = (int32_t) std::floor(sa.pos.x), fy = (int32_t) std::floor(sa.pos.y);
        uint32_t ux = (uint32_t) fx, uy = (uint32_t) fy;
        if (ux < s.film_w && uy < s.film_h) put(ux, uy, V3(sa.weight * 1.f) * sa.result, sa.weight);
    }
}
