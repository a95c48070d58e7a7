// dbsdf.cuh -- microfacet distribution and the BSDF models of the AMVPT path (device + host).
// Follows include/mitsuba/render/microfacet.h (sample_visible = true), include/mitsuba/core/frame.h,
// src/bsdfs/{diffuse,conductor,dielectric,roughconductor,roughdielectric,roughplastic,twosided}.cpp and
// BSDF::eval_pdf_sample (src/render/bsdf.cpp:27-37) of the reference, array (JIT) semantics.
#pragma once
#include "dmath.cuh"

namespace amvpt {

enum BsdfType : int32_t { BSDF_DIFFUSE = 0, BSDF_CONDUCTOR = 1, BSDF_DIELECTRIC = 2, BSDF_ROUGHCONDUCTOR = 3,
                          BSDF_ROUGHDIELECTRIC = 4, BSDF_ROUGHPLASTIC = 5 };
enum MfType : int32_t { MF_BECKMANN = 0, MF_GGX = 1 };
// lobe classes of include/mitsuba/render/bsdf.h:31-117 (only class membership is observable on this path)
enum : uint32_t { L_NULL = 1, L_DIFFUSE_R = 2, L_GLOSSY_R = 4, L_GLOSSY_T = 8, L_DELTA_R = 16, L_DELTA_T = 32,
                  L_SMOOTH = L_DIFFUSE_R | L_GLOSSY_R | L_GLOSSY_T, L_DELTA = L_NULL | L_DELTA_R | L_DELTA_T };
#define AM_ROUGH_TRANSMITTANCE_RES 64   /* src/bsdfs/roughplastic.cpp:13 */

// One record per BSDF plugin instance, 384 B, read through the read-only data path.
struct DBsdf {
    int32_t type, twosided, distr, nonlinear;
    float reflectance[3];      // diffuse `reflectance` / roughplastic `diffuse_reflectance`
    float eta[3], k[3];        // conductor complex IOR
    float spec_refl[3], spec_trans[3];
    float ior_eta;             // int_ior / ext_ior
    float alpha_u, alpha_v;
    float inv_eta_2, spec_sampling_weight, internal_reflectance;   // roughplastic.cpp:207-252
    uint32_t flags;
    float pad[3];
    float ext_transmittance[AM_ROUGH_TRANSMITTANCE_RES];
};
AM_HD V3 ld3(const float *p) { return v3(p[0], p[1], p[2]); }

AM_HD uint32_t bsdf_flags(int32_t type) {
    switch (type) {
        case BSDF_DIFFUSE: return L_DIFFUSE_R;
        case BSDF_CONDUCTOR: return L_DELTA_R;
        case BSDF_DIELECTRIC: return L_DELTA_R | L_DELTA_T;
        case BSDF_ROUGHCONDUCTOR: return L_GLOSSY_R;
        case BSDF_ROUGHDIELECTRIC: return L_GLOSSY_R | L_GLOSSY_T;
        case BSDF_ROUGHPLASTIC: return L_GLOSSY_R | L_DIFFUSE_R;
    }
    return 0;
}

// ------------------------------------------------------------------ microfacet distribution
struct Microfacet {
    int32_t type; float au, av;
    // microfacet.h:85-103 + configure() :425-428 (alpha clamped to >= 1e-4)
    AM_HD Microfacet(int32_t t, float a_u, float a_v) : type(t), au(max_(a_u, 1e-4f)), av(max_(a_v, 1e-4f)) {}

    // microfacet.h:185-210
    AM_HD float eval(V3 m) const {
        float alpha_uv = au * av, ct = m.z, ct2 = sqr(ct), result;
        if (type == MF_BECKMANN)
            result = fdiv(exp_(fdiv(-(sqr(fdiv(m.x, au)) + sqr(fdiv(m.y, av))), ct2)), AM_PI * alpha_uv * sqr(ct2));
        else
            result = rcp(AM_PI * alpha_uv * sqr(sqr(fdiv(m.x, au)) + sqr(fdiv(m.y, av)) + sqr(m.z)));
        return (result * ct > 1e-20f) ? result : 0.f;
    }
    // microfacet.h:341-366
    AM_HD float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = fdiv(xy_alpha_2, sqr(v.z)), result;
        if (type == MF_BECKMANN) {
            float a = rsqrt(tan_theta_alpha_2), a_sqr = sqr(a);
            result = (a >= 1.6f) ? 1.f : fdiv(3.535f * a + 2.181f * a_sqr, 1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = fdiv(2.f, 1.f + fsqrt(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    AM_HD float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }     // :329-331
    AM_HD float pdf(V3 wi, V3 m) const { return eval(m) * fdiv(smith_g1(wi, m) * fabs_(dot(wi, m)), wi.z); }   // :219-228

    // microfacet.h:368-420
    AM_HD V2 sample_visible_11(float cos_theta_i, V2 sample) const {
        if (type == MF_BECKMANN) {
            float tan_theta_i = fdiv(safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)), cos_theta_i);
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = erf_(cot_theta_i);
            sample.x = max_(min_(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = max_(min_(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erf_(fsqrt(-log_(sample.x)));
            sample.x *= 1.f + maxval + AM_INV_SQRT_PI * tan_theta_i * exp_(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_(x),
                      value = 1.f + x + AM_INV_SQRT_PI * tan_theta_i * exp_(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= fdiv(value, derivative);
            }
            return v2(erfinv_(x), erfinv_(fmsub(2.f, sample.y, 1.f)));
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            p.y = lerp_(safe_sqrt(1.f - sqr(p.x)), p.y, s);
            float x = p.x, y = p.y, z = safe_sqrt(1.f - squared_norm(p));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = rcp(fmadd(sin_theta_i, y, cos_theta_i * z));
            return v2(fmsub(cos_theta_i, y, sin_theta_i * z) * nrm, x * nrm);
        }
    }
    // microfacet.h:244-327, visible-normal branch :296-326; Frame::sincos_phi frame.h:111-122
    AM_HD V3 sample(V3 wi, V2 u, float *pdf_out) const {
        V3 wi_p = normalize(v3(au * wi.x, av * wi.y, wi.z));
        float st2 = fmadd(wi_p.x, wi_p.x, sqr(wi_p.y)), inv = rsqrt(st2);
        float cos_phi = wi_p.x * inv, sin_phi = wi_p.y * inv;
        if (fabs_(st2) <= 4.f * AM_EPSILON) { cos_phi = 1.f; sin_phi = 0.f; }
        else { cos_phi = clip_(cos_phi, -1.f, 1.f); sin_phi = clip_(sin_phi, -1.f, 1.f); }
        V2 slope = sample_visible_11(wi_p.z, u);
        slope = v2(fmsub(cos_phi, slope.x, sin_phi * slope.y) * au, fmadd(sin_phi, slope.x, cos_phi * slope.y) * av);
        V3 m = normalize(v3(-slope.x, -slope.y, 1.f));
        *pdf_out = fdiv(eval(m) * smith_g1(wi, m) * fabs_(dot(wi, m)), wi.z);
        return m;
    }
};

// ------------------------------------------------------------------ BSDF models
struct BSDFSample { V3 wo; float pdf, eta; uint32_t type; };
struct BsdfResult { V3 f; float pdf; BSDFSample bs; V3 weight; };

AM_HD V3 fresnel_conductor3(float c, const DBsdf &b) {
    return v3(fresnel_conductor(c, b.eta[0], b.k[0]), fresnel_conductor(c, b.eta[1], b.k[1]), fresnel_conductor(c, b.eta[2], b.k[2]));
}
// roughplastic.cpp:365-376
AM_HD float lerp_gather(const float *data, float x, int size) {
    x *= (float) (size - 1);
    uint32_t index = x > 0.f ? (uint32_t) x : 0u;      // negative / NaN saturate to 0 (cvt.rzi.u32.f32 semantics)
    if (index > (uint32_t) (size - 2)) index = size - 2;
    float v0 = data[index], v1 = data[index + 1];
    return lerp_(v0, v1, x - (float) index);
}

// diffuse.cpp:100-184
AM_HD void diffuse_all(const DBsdf &b, V3 wi, V3 wo, V2 s2, BsdfResult &r) {
    V3 refl = ld3(b.reflectance);
    bool act_e = wi.z > 0.f && wo.z > 0.f;
    r.f = select3(act_e, refl * AM_INV_PI * wo.z);
    r.pdf = act_e ? square_to_cosine_hemisphere_pdf(wo) : 0.f;
    r.bs.wo = square_to_cosine_hemisphere(s2);
    r.bs.pdf = square_to_cosine_hemisphere_pdf(r.bs.wo);
    r.bs.eta = 1.f; r.bs.type = L_DIFFUSE_R;
    r.weight = select3(wi.z > 0.f && r.bs.pdf > 0.f, refl);
}
// conductor.cpp:247-317 (eval/pdf of a delta lobe are zero)
AM_HD void conductor_all(const DBsdf &b, V3 wi, BsdfResult &r) {
    r.f = v3(0.f); r.pdf = 0.f;
    r.bs.type = L_DELTA_R; r.bs.wo = reflect(wi); r.bs.eta = 1.f; r.bs.pdf = 1.f;
    r.weight = select3(wi.z > 0.f, ld3(b.spec_refl) * fresnel_conductor3(wi.z, b));
}
// dielectric.cpp:245-370
AM_HD void dielectric_all(const DBsdf &b, V3 wi, float s1, BsdfResult &r) {
    r.f = v3(0.f); r.pdf = 0.f;
    float r_i, cos_t, eta_it, eta_ti; fresnel(wi.z, b.ior_eta, &r_i, &cos_t, &eta_it, &eta_ti);
    float t_i = 1.f - r_i;
    bool sel_r = s1 <= r_i;
    r.bs.pdf = sel_r ? r_i : t_i;
    r.bs.type = sel_r ? L_DELTA_R : L_DELTA_T;
    r.bs.wo = sel_r ? reflect(wi) : refract(wi, cos_t, eta_ti);
    r.bs.eta = sel_r ? 1.f : eta_it;
    V3 w = v3(1.f);
    if (sel_r) w = w * ld3(b.spec_refl); else { w = w * ld3(b.spec_trans); w = w * sqr(eta_ti); }
    r.weight = w;
}
// roughconductor.cpp:225-311 (sample), 427-516 (eval_pdf)
AM_HD void roughconductor_all(const DBsdf &b, V3 wi, V3 wo, V2 s2, BsdfResult &r) {
    Microfacet d(b.distr, b.alpha_u, b.alpha_v);
    V3 spec = ld3(b.spec_refl);
    {
        float ci = wi.z, co = wo.z;
        V3 H = normalize(wo + wi);
        bool active = ci > 0.f && co > 0.f && dot(wi, H) > 0.f && dot(wo, H) > 0.f;
        float D = d.eval(H);
        active = active && D != 0.f;
        float g1 = d.smith_g1(wi, H), G = g1 * d.smith_g1(wo, H);
        V3 value = v3(fdiv(D * G, 4.f * wi.z));
        V3 F = fresnel_conductor3(dot(wi, H), b);
        value = value * spec;
        float p = fdiv(D * g1, 4.f * ci);
        r.f = select3(active, F * value); r.pdf = active ? p : 0.f;
    }
    {
        bool active = wi.z > 0.f;
        V3 m = d.sample(wi, s2, &r.bs.pdf);
        r.bs.wo = reflect(wi, m); r.bs.eta = 1.f; r.bs.type = L_GLOSSY_R;
        active = active && (r.bs.pdf != 0.f) && r.bs.wo.z > 0.f;
        float weight = d.smith_g1(r.bs.wo, m);
        r.bs.pdf = fdiv(r.bs.pdf, 4.f * dot(r.bs.wo, m));
        V3 F = fresnel_conductor3(dot(wi, m), b);
        r.weight = select3(active, F * (v3(weight) * spec));
    }
}
// roughdielectric.cpp:239-355 (sample), 498-605 (eval_pdf)
AM_HD void roughdielectric_all(const DBsdf &b, V3 wi, V3 wo, float s1, V2 s2, BsdfResult &r) {
    Microfacet d(b.distr, b.alpha_u, b.alpha_v);
    float ci = wi.z;
    V3 wi_s = v3(mulsign(wi.x, ci), mulsign(wi.y, ci), mulsign(wi.z, ci));
    {
        float co = wo.z;
        bool active = ci != 0.f;
        bool refl = ci * co > 0.f;
        float m_eta = b.ior_eta, m_inv_eta = rcp(b.ior_eta);
        float eta = ci > 0.f ? m_eta : m_inv_eta, inv_eta = ci > 0.f ? m_inv_eta : m_eta;
        V3 m = normalize(wi + wo * (refl ? 1.f : eta));
        m = v3(mulsign(m.x, m.z), mulsign(m.y, m.z), mulsign(m.z, m.z));
        float dwi = dot(wi, m), dwo = dot(wo, m);
        active = active && dwi * ci > 0.f && dwo * co > 0.f;
        float D = d.eval(m);
        float F, t0, t1, t2; fresnel(dwi, m_eta, &F, &t0, &t1, &t2);
        float G = d.G(wi, wo, m);
        V3 result = v3(0.f);
        if (refl && active) {
            result = v3(fdiv(F * D * G, 4.f * fabs_(ci))) * ld3(b.spec_refl);
        } else if (!refl && active) {
            float scale = sqr(inv_eta);
            float v = fabs_(fdiv(scale * (1.f - F) * D * G * eta * eta * dwi * dwo, ci * sqr(dwi + eta * dwo)));
            result = v3(v) * ld3(b.spec_trans);
        }
        float p = d.pdf(wi_s, m);
        p *= refl ? F : 1.f - F;
        float dwh_dwo = refl ? rcp(4.f * dwo) : fdiv(eta * eta * dwo, sqr(dwi + eta * dwo));
        r.f = result; r.pdf = active ? p * fabs_(dwh_dwo) : 0.f;
    }
    {
        bool active = ci != 0.f;
        V3 m = d.sample(wi_s, s2, &r.bs.pdf);
        active = active && r.bs.pdf != 0.f;
        float F, cos_t, eta_it, eta_ti; fresnel(dot(wi, m), b.ior_eta, &F, &cos_t, &eta_it, &eta_ti);
        bool sel_r = s1 <= F && active;
        V3 weight = v3(1.f);
        r.bs.pdf *= sel_r ? F : 1.f - F;
        bool sel_t = !sel_r && active;
        r.bs.eta = sel_r ? 1.f : eta_it;
        r.bs.type = sel_r ? L_GLOSSY_R : L_GLOSSY_T;
        r.bs.wo = v3(0.f);
        if (sel_r) { r.bs.wo = reflect(wi, m); weight = weight * ld3(b.spec_refl); }
        float dwh_dwo = rcp(4.f * dot(r.bs.wo, m));
        if (sel_t) {
            r.bs.wo = refract(wi, m, cos_t, eta_ti);
            V3 factor = v3(sqr(eta_ti)) * ld3(b.spec_trans);
            weight = weight * factor;
            dwh_dwo = fdiv(sqr(r.bs.eta) * dot(r.bs.wo, m), sqr(dot(wi, m) + r.bs.eta * dot(r.bs.wo, m)));
        }
        weight = weight * d.smith_g1(r.bs.wo, m);
        r.bs.pdf *= fabs_(dwh_dwo);
        r.weight = select3(active, weight);
    }
}
// roughplastic.cpp:256-312 (sample), 314-360 (eval), 380-422 (pdf), 424-500 (eval_pdf)
AM_HD void roughplastic_probs(const DBsdf &b, float ci, float &t_i, float &ps, float &pd) {
    t_i = lerp_gather(b.ext_transmittance, ci, AM_ROUGH_TRANSMITTANCE_RES);
    ps = (1.f - t_i) * b.spec_sampling_weight; pd = t_i * (1.f - b.spec_sampling_weight);
    ps = fdiv(ps, ps + pd); pd = 1.f - ps;
}
AM_HD V3 roughplastic_diffuse_term(const DBsdf &b, float co, float t_i, float t_o) {
    V3 diff = ld3(b.reflectance);
    V3 den = b.nonlinear ? diff * b.internal_reflectance : v3(b.internal_reflectance);
    diff = v3(fdiv(diff.x, 1.f - den.x), fdiv(diff.y, 1.f - den.y), fdiv(diff.z, 1.f - den.z));
    return diff * (AM_INV_PI * b.inv_eta_2 * co * t_i * t_o);
}
AM_HD void roughplastic_all(const DBsdf &b, V3 wi, V3 wo, float s1, V2 s2, BsdfResult &r) {
    Microfacet d(b.distr, b.alpha_u, b.alpha_u);
    float ci = wi.z;
    float t_i, ps, pd; roughplastic_probs(b, ci, t_i, ps, pd);
    V3 spec = ld3(b.spec_refl);
    {   // eval_pdf
        float co = wo.z;
        bool active = ci > 0.f && co > 0.f;
        V3 H = normalize(wo + wi);
        float D = d.eval(H), g1 = d.smith_g1(wi, H);
        float p = fdiv(D * g1, 4.f * ci);
        p *= ps;
        p += pd * square_to_cosine_hemisphere_pdf(wo);
        float F, t0, t1, t2; fresnel(dot(wi, H), b.ior_eta, &F, &t0, &t1, &t2);
        float G = d.smith_g1(wo, H) * g1;
        V3 value = v3(fdiv(F * D * G, 4.f * ci)) * spec;
        float t_o = lerp_gather(b.ext_transmittance, co, AM_ROUGH_TRANSMITTANCE_RES);
        value = value + roughplastic_diffuse_term(b, co, t_i, t_o);
        r.f = select3(active, value); r.pdf = p;      // pdf is returned unmasked (roughplastic.cpp:421)
    }
    {   // sample = lobe choice, then pdf() and eval() of the sampled direction
        bool active = ci > 0.f;
        bool sample_spec = active && (s1 < ps), sample_diff = active && !sample_spec;
        r.bs.eta = 1.f; r.bs.wo = v3(0.f); r.bs.type = 0;
        if (sample_spec) {
            float tmp; V3 m = d.sample(wi, s2, &tmp);
            r.bs.wo = reflect(wi, m); r.bs.type = L_GLOSSY_R;
        }
        if (sample_diff) { r.bs.wo = square_to_cosine_hemisphere(s2); r.bs.type = L_DIFFUSE_R; }
        V3 wo2 = r.bs.wo; float co = wo2.z;
        V3 H = normalize(wo2 + wi);
        float D = d.eval(H), g1 = d.smith_g1(wi, H);
        float p = fdiv(D * g1, 4.f * ci);
        p *= ps;
        p += pd * square_to_cosine_hemisphere_pdf(wo2);
        r.bs.pdf = p;
        active = active && p > 0.f;
        bool act_e = active && ci > 0.f && co > 0.f;
        float F, t0, t1, t2; fresnel(dot(wi, H), b.ior_eta, &F, &t0, &t1, &t2);
        float G = g1 * d.smith_g1(wo2, H);
        V3 value = v3(fdiv(F * D * G, 4.f * ci)) * spec;
        float t_o = lerp_gather(b.ext_transmittance, co, AM_ROUGH_TRANSMITTANCE_RES);
        value = value + roughplastic_diffuse_term(b, co, t_i, t_o);
        value = select3(act_e, value);
        r.weight = select3(active, value / p);
    }
}

// BSDF::eval_pdf_sample (bsdf.cpp:27-37) = eval_pdf + sample; TwoSidedBRDF (twosided.cpp:110-260, one nested
// BRDF used for both sides) mirrors wi / wo about the surface when wi is on the back side.
AM_HD BsdfResult bsdf_eval_pdf_sample(const DBsdf &b, V3 wi_, V3 wo_em, float s1, V2 s2) {
    BsdfResult r;
    V3 wi = wi_, wo = wo_em;
    if (b.twosided) { wo.z = mulsign(wo.z, wi.z); wi.z = fabs_(wi.z); }
    switch (b.type) {
        case BSDF_DIFFUSE: diffuse_all(b, wi, wo, s2, r); break;
        case BSDF_CONDUCTOR: conductor_all(b, wi, r); break;
        case BSDF_DIELECTRIC: dielectric_all(b, wi, s1, r); break;
        case BSDF_ROUGHCONDUCTOR: roughconductor_all(b, wi, wo, s2, r); break;
        case BSDF_ROUGHDIELECTRIC: roughdielectric_all(b, wi, wo, s1, s2, r); break;
        default: roughplastic_all(b, wi, wo, s1, s2, r); break;
    }
    if (b.twosided) r.bs.wo.z = mulsign(r.bs.wo.z, wi_.z);
    return r;
}
// Same with the plugin type known at compile time (the wavefront scheduler sorts vertices by BSDF type and runs one
// specialised kernel per type); TYPE < 0 = dispatch at run time.
template <int TYPE> AM_HD BsdfResult bsdf_eval_pdf_sample_t(const DBsdf &b, V3 wi_, V3 wo_em, float s1, V2 s2) {
    if (TYPE < 0) return bsdf_eval_pdf_sample(b, wi_, wo_em, s1, s2);
    BsdfResult r;
    V3 wi = wi_, wo = wo_em;
    if (b.twosided) { wo.z = mulsign(wo.z, wi.z); wi.z = fabs_(wi.z); }
    if (TYPE == BSDF_DIFFUSE) diffuse_all(b, wi, wo, s2, r);
    else if (TYPE == BSDF_CONDUCTOR) conductor_all(b, wi, r);
    else if (TYPE == BSDF_DIELECTRIC) dielectric_all(b, wi, s1, r);
    else if (TYPE == BSDF_ROUGHCONDUCTOR) roughconductor_all(b, wi, wo, s2, r);
    else if (TYPE == BSDF_ROUGHDIELECTRIC) roughdielectric_all(b, wi, wo, s1, s2, r);
    else roughplastic_all(b, wi, wo, s1, s2, r);
    if (b.twosided) r.bs.wo.z = mulsign(r.bs.wo.z, wi_.z);
    return r;
}

} // namespace amvpt
