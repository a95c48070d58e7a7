// oracle/o_bsdf.h -- TEST INFRASTRUCTURE ONLY.  Restatement (JIT/array semantics) of the BSDF
// plugins on the AMVPT path: src/bsdfs/{diffuse,conductor,dielectric,roughconductor,
// roughdielectric,roughplastic,twosided}.cpp and BSDF::eval_pdf_sample (src/render/bsdf.cpp:27-37).
#pragma once
#include "o_microfacet.h"

namespace orc {

enum BsdfType { BSDF_DIFFUSE = 0, BSDF_CONDUCTOR = 1, BSDF_DIELECTRIC = 2, BSDF_ROUGHCONDUCTOR = 3,
                BSDF_ROUGHDIELECTRIC = 4, BSDF_ROUGHPLASTIC = 5 };
// lobe classes (include/mitsuba/render/bsdf.h:31-117); only class membership matters on the path
enum Lobe { L_NONE = 0, L_NULL = 1, L_DIFFUSE_R = 2, L_GLOSSY_R = 4, L_GLOSSY_T = 8, L_DELTA_R = 16, L_DELTA_T = 32 };
constexpr uint32_t L_SMOOTH = L_DIFFUSE_R | L_GLOSSY_R | L_GLOSSY_T, L_DELTA = L_NULL | L_DELTA_R | L_DELTA_T;

#define MI_ROUGH_TRANSMITTANCE_RES 64   // src/bsdfs/roughplastic.cpp:13

struct Bsdf {
    int type = BSDF_DIFFUSE, twosided = 0;
    V3 reflectance{ .5f };              // diffuse `reflectance` / roughplastic `diffuse_reflectance`
    V3 eta{ 0.f }, k{ 1.f };            // conductor
    V3 spec_refl{ 1.f }, spec_trans{ 1.f };
    float ior_eta = 1.5046f / 1.000277f; // int_ior / ext_ior
    float alpha_u = .1f, alpha_v = .1f; int distr = Beckmann; int nonlinear = 0;
    // roughplastic precomputation (roughplastic.cpp:207-252)
    float inv_eta_2 = 0, spec_sampling_weight = 0, internal_reflectance = 0;
    float ext_transmittance[MI_ROUGH_TRANSMITTANCE_RES];

    uint32_t flags() const {
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
    // roughplastic.cpp:207-252 parameters_changed()
    void prepare() {
        if (type != BSDF_ROUGHPLASTIC) return;
        inv_eta_2 = 1.f / (ior_eta * ior_eta);
        // Texture::mean() of an RGB constant (src/spectra/srgb.cpp mean = hmean of the color)
        float d_mean = (reflectance.x + reflectance.y + reflectance.z) / 3.f;
        float s_mean = (spec_refl.x + spec_refl.y + spec_refl.z) / 3.f;
        spec_sampling_weight = s_mean / (d_mean + s_mean);
        Microfacet d((MfType) distr, alpha_u, alpha_u);
        float acc = 0.f;
        for (int i = 0; i < MI_ROUGH_TRANSMITTANCE_RES; ++i) {
            // dr::linspace(0,1,N): start + i * step, step = 1/(N-1) (drjit/array_router.h linspace: fmadd(i, step, min))
            float mu = fmax_(1e-6f, fmadd((float) i, 1.f / (float) (MI_ROUGH_TRANSMITTANCE_RES - 1), 0.f));
            V3 wi(std::sqrt(1.f - mu * mu), 0.f, mu);
            ext_transmittance[i] = mf_eval_transmittance(d, wi, ior_eta);
            acc += mf_eval_reflectance(d, wi, 1.f / ior_eta) * wi.z;
        }
        internal_reflectance = acc / (float) MI_ROUGH_TRANSMITTANCE_RES * 2.f;
    }
};

struct BSDFSample { V3 wo; float pdf = 0, eta = 0; uint32_t type = 0; };
struct BsdfResult { V3 f; float pdf = 0; BSDFSample bs; V3 weight; };

static inline V3 mask3(V3 v, bool m) { return m ? v : V3(0.f); }

// roughplastic.cpp:365-376
static inline float lerp_gather(const float *data, float x, int size) {
    x *= (float) (size - 1);
    // float -> uint32 of a negative / NaN value is undefined in C++; the reference's CUDA variant saturates it to 0
    // (cvt.rzi.u32.f32).  Only reachable for wi.z < 0, where every observable output is masked to zero.
    uint32_t index = x > 0.f ? (uint32_t) x : 0u; if (index > (uint32_t) (size - 2)) index = size - 2;
    float v0 = data[index], v1 = data[index + 1];
    return lerp(v0, v1, x - (float) index);
}

// ---- diffuse.cpp:100-184
static inline void diffuse_sample(const Bsdf &b, V3 wi, V2 s2, BSDFSample &bs, V3 &w) {
    bool active = wi.z > 0.f;
    bs.wo = square_to_cosine_hemisphere(s2);
    bs.pdf = square_to_cosine_hemisphere_pdf(bs.wo);
    bs.eta = 1.f; bs.type = L_DIFFUSE_R;
    w = mask3(b.reflectance, active && bs.pdf > 0.f);
}
static inline void diffuse_eval_pdf(const Bsdf &b, V3 wi, V3 wo, V3 &f, float &pdf) {
    bool active = wi.z > 0.f && wo.z > 0.f;
    f = mask3(b.reflectance * InvPi * wo.z, active);
    pdf = active ? square_to_cosine_hemisphere_pdf(wo) : 0.f;
}
// ---- conductor.cpp:247-317
static inline void conductor_sample(const Bsdf &b, V3 wi, BSDFSample &bs, V3 &w) {
    bool active = wi.z > 0.f;
    bs.type = L_DELTA_R; bs.wo = reflect(wi); bs.eta = 1.f; bs.pdf = 1.f;
    V3 F(fresnel_conductor(wi.z, b.eta.x, b.k.x), fresnel_conductor(wi.z, b.eta.y, b.k.y),
         fresnel_conductor(wi.z, b.eta.z, b.k.z));
    w = mask3(b.spec_refl * F, active);
}
// ---- dielectric.cpp:245-370
static inline void dielectric_sample(const Bsdf &b, V3 wi, float s1, BSDFSample &bs, V3 &w) {
    float r_i, cos_t, eta_it, eta_ti; fresnel(wi.z, b.ior_eta, &r_i, &cos_t, &eta_it, &eta_ti);
    float t_i = 1.f - r_i;
    bool sel_r = s1 <= r_i;
    bs.pdf = sel_r ? r_i : t_i;
    bs.type = sel_r ? L_DELTA_R : L_DELTA_T;
    bs.wo = sel_r ? reflect(wi) : refract(wi, cos_t, eta_ti);
    bs.eta = sel_r ? 1.f : eta_it;
    w = V3(1.f);
    if (sel_r) w = w * b.spec_refl; else { w = w * b.spec_trans; w = w * sqr(eta_ti); }
}
// ---- roughconductor.cpp:225-311, 427-516
static inline V3 fresnel_conductor3(float c, const Bsdf &b) {
    return V3(fresnel_conductor(c, b.eta.x, b.k.x), fresnel_conductor(c, b.eta.y, b.k.y), fresnel_conductor(c, b.eta.z, b.k.z));
}
static inline void roughconductor_sample(const Bsdf &b, V3 wi, V2 s2, BSDFSample &bs, V3 &w) {
    bool active = wi.z > 0.f;
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_v);
    V3 m = d.sample(wi, s2, &bs.pdf);
    bs.wo = reflect(wi, m); bs.eta = 1.f; bs.type = L_GLOSSY_R;
    active = active && (bs.pdf != 0.f) && bs.wo.z > 0.f;
    float weight = d.smith_g1(bs.wo, m);
    bs.pdf /= 4.f * dot(bs.wo, m);
    V3 F = fresnel_conductor3(dot(wi, m), b);
    V3 wv = V3(weight) * b.spec_refl;
    w = mask3(F * wv, active);
}
static inline void roughconductor_eval_pdf(const Bsdf &b, V3 wi, V3 wo, V3 &f, float &pdf) {
    float ci = wi.z, co = wo.z;
    V3 H = normalize(wo + wi);
    bool active = ci > 0.f && co > 0.f && dot(wi, H) > 0.f && dot(wo, H) > 0.f;
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_v);
    float D = d.eval(H);
    active = active && D != 0.f;
    float g1 = d.smith_g1(wi, H), G = g1 * d.smith_g1(wo, H);
    V3 value = V3(D * G / (4.f * wi.z));
    V3 F = fresnel_conductor3(dot(wi, H), b);
    value = value * b.spec_refl;
    float p = D * g1 / (4.f * ci);
    f = mask3(F * value, active); pdf = active ? p : 0.f;
}
// ---- roughdielectric.cpp:239-355, 498-605 (sample_visible = true)
static inline void roughdielectric_sample(const Bsdf &b, V3 wi, float s1, V2 s2, BSDFSample &bs, V3 &w) {
    float ci = wi.z;
    bool active = ci != 0.f;
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_v);
    V3 wi_s(mulsign(wi.x, ci), mulsign(wi.y, ci), mulsign(wi.z, ci));
    V3 m = d.sample(wi_s, s2, &bs.pdf);
    active = active && bs.pdf != 0.f;
    float F, cos_t, eta_it, eta_ti; fresnel(dot(wi, m), b.ior_eta, &F, &cos_t, &eta_it, &eta_ti);
    bool sel_r = s1 <= F && active;
    V3 weight(1.f);
    bs.pdf *= sel_r ? F : 1.f - F;
    bool sel_t = !sel_r && active;
    bs.eta = sel_r ? 1.f : eta_it;
    bs.type = sel_r ? L_GLOSSY_R : L_GLOSSY_T;
    bs.wo = V3(0.f);
    float dwh_dwo;
    if (sel_r) { bs.wo = reflect(wi, m); weight = weight * b.spec_refl; }
    dwh_dwo = rcp(4.f * dot(bs.wo, m));
    if (sel_t) {
        bs.wo = refract(wi, m, cos_t, eta_ti);
        V3 factor = V3(sqr(eta_ti)) * b.spec_trans;
        weight = weight * factor;
        dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(dot(wi, m) + bs.eta * dot(bs.wo, m));
    }
    weight = weight * d.smith_g1(bs.wo, m);
    bs.pdf *= fabs_(dwh_dwo);
    w = mask3(weight, active);
}
static inline void roughdielectric_eval_pdf(const Bsdf &b, V3 wi, V3 wo, V3 &f, float &pdf) {
    float ci = wi.z, co = wo.z;
    bool active = ci != 0.f;
    bool refl = ci * co > 0.f;
    float m_eta = b.ior_eta, m_inv_eta = rcp(b.ior_eta);
    float eta = ci > 0.f ? m_eta : m_inv_eta, inv_eta = ci > 0.f ? m_inv_eta : m_eta;
    V3 m = normalize(wi + wo * (refl ? 1.f : eta));
    m = V3(mulsign(m.x, m.z), mulsign(m.y, m.z), mulsign(m.z, m.z));
    float dwi = dot(wi, m), dwo = dot(wo, m);
    active = active && dwi * ci > 0.f && dwo * co > 0.f;
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_v);
    float D = d.eval(m);
    float F, t0, t1, t2; fresnel(dwi, m_eta, &F, &t0, &t1, &t2);
    float G = d.G(wi, wo, m);
    V3 result(0.f);
    if (refl && active) {
        result = V3(F * D * G / (4.f * fabs_(ci))) * b.spec_refl;
    } else if (!refl && active) {
        float scale = sqr(inv_eta);
        float v = fabs_((scale * (1.f - F) * D * G * eta * eta * dwi * dwo) / (ci * sqr(dwi + eta * dwo)));
        result = V3(v) * b.spec_trans;
    }
    V3 wi_s(mulsign(wi.x, ci), mulsign(wi.y, ci), mulsign(wi.z, ci));
    float p = d.pdf(wi_s, m);
    p *= refl ? F : 1.f - F;
    float dwh_dwo = refl ? rcp(4.f * dwo) : (eta * eta * dwo) / sqr(dwi + eta * dwo);
    f = result; pdf = active ? p * fabs_(dwh_dwo) : 0.f;
}
// ---- roughplastic.cpp:256-312 (sample), 314-360 (eval), 380-422 (pdf), 424-500 (eval_pdf)
static inline void roughplastic_probs(const Bsdf &b, float ci, float &t_i, float &ps, float &pd) {
    t_i = lerp_gather(b.ext_transmittance, ci, MI_ROUGH_TRANSMITTANCE_RES);
    ps = (1.f - t_i) * b.spec_sampling_weight; pd = t_i * (1.f - b.spec_sampling_weight);
    ps = ps / (ps + pd); pd = 1.f - ps;
}
static inline float roughplastic_pdf(const Bsdf &b, V3 wi, V3 wo, bool active) {
    float ci = wi.z, co = wo.z;
    active = active && ci > 0.f && co > 0.f;
    (void) active;   // the reference's pdf() returns `result` unmasked (roughplastic.cpp:421)
    float t_i, ps, pd; roughplastic_probs(b, ci, t_i, ps, pd);
    V3 H = normalize(wo + wi);
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_u);
    float result = d.eval(H) * d.smith_g1(wi, H) / (4.f * ci);
    result *= ps;
    result += pd * square_to_cosine_hemisphere_pdf(wo);
    return result;
}
static inline V3 roughplastic_diffuse_term(const Bsdf &b, float co, float t_i, float t_o) {
    V3 diff = b.reflectance;
    V3 den = b.nonlinear ? diff * b.internal_reflectance : V3(b.internal_reflectance);
    diff = V3(diff.x / (1.f - den.x), diff.y / (1.f - den.y), diff.z / (1.f - den.z));
    return diff * (InvPi * b.inv_eta_2 * co * t_i * t_o);
}
static inline V3 roughplastic_eval(const Bsdf &b, V3 wi, V3 wo, bool active) {
    float ci = wi.z, co = wo.z;
    active = active && ci > 0.f && co > 0.f;
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_u);
    V3 H = normalize(wo + wi);
    float D = d.eval(H);
    float F, t0, t1, t2; fresnel(dot(wi, H), b.ior_eta, &F, &t0, &t1, &t2);
    float G = d.G(wi, wo, H);
    V3 value = V3(F * D * G / (4.f * ci)) * b.spec_refl;
    float t_i = lerp_gather(b.ext_transmittance, ci, MI_ROUGH_TRANSMITTANCE_RES),
          t_o = lerp_gather(b.ext_transmittance, co, MI_ROUGH_TRANSMITTANCE_RES);
    value = value + roughplastic_diffuse_term(b, co, t_i, t_o);
    return mask3(value, active);
}
static inline void roughplastic_sample(const Bsdf &b, V3 wi, float s1, V2 s2, BSDFSample &bs, V3 &w) {
    float ci = wi.z;
    bool active = ci > 0.f;
    float t_i, ps, pd; roughplastic_probs(b, ci, t_i, ps, pd);
    bool sample_spec = active && (s1 < ps), sample_diff = active && !sample_spec;
    bs.eta = 1.f; bs.wo = V3(0.f); bs.type = 0;
    if (sample_spec) {
        Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_u);
        float tmp; V3 m = d.sample(wi, s2, &tmp);
        bs.wo = reflect(wi, m); bs.type = L_GLOSSY_R;
    }
    if (sample_diff) { bs.wo = square_to_cosine_hemisphere(s2); bs.type = L_DIFFUSE_R; }
    bs.pdf = roughplastic_pdf(b, wi, bs.wo, active);
    active = active && bs.pdf > 0.f;
    V3 result = roughplastic_eval(b, wi, bs.wo, active);
    w = mask3(result / bs.pdf, active);   // Spectrum / Float -> result * rcp(pdf)
}
static inline void roughplastic_eval_pdf(const Bsdf &b, V3 wi, V3 wo, V3 &f, float &pdf) {
    float ci = wi.z, co = wo.z;
    bool active = ci > 0.f && co > 0.f;
    float t_i, ps, pd; roughplastic_probs(b, ci, t_i, ps, pd);
    V3 H = normalize(wo + wi);
    Microfacet d((MfType) b.distr, b.alpha_u, b.alpha_u);
    float D = d.eval(H), g1 = d.smith_g1(wi, H);
    float p = D * g1 / (4.f * ci);
    p *= ps;
    p += pd * square_to_cosine_hemisphere_pdf(wo);
    float F, t0, t1, t2; fresnel(dot(wi, H), b.ior_eta, &F, &t0, &t1, &t2);
    float G = d.smith_g1(wo, H) * g1;
    V3 value = V3(F * D * G / (4.f * ci)) * b.spec_refl;
    float t_o = lerp_gather(b.ext_transmittance, co, MI_ROUGH_TRANSMITTANCE_RES);
    value = value + roughplastic_diffuse_term(b, co, t_i, t_o);
    f = mask3(value, active); pdf = p;
}

// BSDF::eval_pdf_sample (bsdf.cpp:27-37) = eval_pdf + sample; TwoSidedBRDF (twosided.cpp:110-260,
// single nested BRDF branch) flips wi.z / wo.z.
static inline BsdfResult bsdf_eval_pdf_sample(const Bsdf &b, V3 wi_, V3 wo_em, float s1, V2 s2) {
    BsdfResult r;
    V3 wi = wi_, wo = wo_em;
    if (b.twosided) { wo.z = mulsign(wo.z, wi.z); wi.z = fabs_(wi.z); }
    switch (b.type) {
        case BSDF_DIFFUSE: diffuse_eval_pdf(b, wi, wo, r.f, r.pdf); diffuse_sample(b, wi, s2, r.bs, r.weight); break;
        case BSDF_CONDUCTOR: r.f = V3(0.f); r.pdf = 0.f; conductor_sample(b, wi, r.bs, r.weight); break;
        case BSDF_DIELECTRIC: r.f = V3(0.f); r.pdf = 0.f; dielectric_sample(b, wi, s1, r.bs, r.weight); break;
        case BSDF_ROUGHCONDUCTOR: roughconductor_eval_pdf(b, wi, wo, r.f, r.pdf); roughconductor_sample(b, wi, s2, r.bs, r.weight); break;
        case BSDF_ROUGHDIELECTRIC: roughdielectric_eval_pdf(b, wi, wo, r.f, r.pdf); roughdielectric_sample(b, wi, s1, s2, r.bs, r.weight); break;
        case BSDF_ROUGHPLASTIC: roughplastic_eval_pdf(b, wi, wo, r.f, r.pdf); roughplastic_sample(b, wi, s1, s2, r.bs, r.weight); break;
    }
    if (b.twosided) r.bs.wo.z = mulsign(r.bs.wo.z, wi_.z);
    return r;
}

} // namespace orc
