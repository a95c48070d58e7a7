// oracle/o_microfacet.h -- TEST INFRASTRUCTURE ONLY.  Restatement of
// include/mitsuba/render/microfacet.h (MicrofacetDistribution, sample_visible = true path),
// include/mitsuba/core/frame.h:59-122 and include/mitsuba/core/quad.h:27-86.
#pragma once
#include "o_math.h"
#include <vector>

namespace orc {

enum MfType { Beckmann = 0, GGX = 1 };

// frame.h:65, :111-122
static inline float sin_theta_2(V3 v) { return fmadd(v.x, v.x, sqr(v.y)); }
static inline void sincos_phi(V3 v, float *sin_phi, float *cos_phi) {
    float st2 = sin_theta_2(v), inv = rsqrt(sin_theta_2(v));
    float rx = v.x * inv, ry = v.y * inv;
    if (fabs_(st2) <= 4.f * Epsilon) { rx = 1.f; ry = 0.f; }
    else { rx = clip(rx, -1.f, 1.f); ry = clip(ry, -1.f, 1.f); }
    *sin_phi = ry; *cos_phi = rx;
}

struct Microfacet {
    MfType type; float au, av;
    // microfacet.h:85-103 ctor + :425-428 configure()
    Microfacet(MfType t, float a_u, float a_v) : type(t), au(fmax_(a_u, 1e-4f)), av(fmax_(a_v, 1e-4f)) {}

    // microfacet.h:185-210
    float eval(V3 m) const {
        float alpha_uv = au * av, ct = m.z, ct2 = sqr(ct), result;
        if (type == Beckmann)
            result = exp_(-(sqr(m.x / au) + sqr(m.y / av)) / ct2) / (Pi * alpha_uv * sqr(ct2));
        else
            result = rcp(Pi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return (result * ct > 1e-20f) ? result : 0.f;
    }
    // microfacet.h:341-366
    float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == Beckmann) {
            float a = rsqrt(tan_theta_alpha_2), a_sqr = sqr(a);
            result = (a >= 1.6f) ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else {
            result = 2.f / (1.f + std::sqrt(1.f + tan_theta_alpha_2));
        }
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }   // :329-331
    // microfacet.h:219-228 (sample_visible branch)
    float pdf(V3 wi, V3 m) const { return eval(m) * (smith_g1(wi, m) * fabs_(dot(wi, m)) / wi.z); }

    // microfacet.h:368-420
    V2 sample_visible_11(float cos_theta_i, V2 sample) const {
        if (type == Beckmann) {
            float tan_theta_i = safe_sqrt(fnmadd(cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = rcp(tan_theta_i);
            float maxval = erf_(cot_theta_i);
            sample.x = fmax_(fmin_(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = fmax_(fmin_(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erf_(std::sqrt(-log_(sample.x)));
            sample.x *= 1.f + maxval + InvSqrtPi * tan_theta_i * exp_(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_(x),
                      value = 1.f + x + InvSqrtPi * tan_theta_i * exp_(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return { erfinv_(x), erfinv_(fmsub(2.f, sample.y, 1.f)) };
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            p.y = lerp(safe_sqrt(1.f - sqr(p.x)), p.y, s);
            float x = p.x, y = p.y, z = safe_sqrt(1.f - squared_norm(p));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = rcp(fmadd(sin_theta_i, y, cos_theta_i * z));
            return { fmsub(cos_theta_i, y, sin_theta_i * z) * nrm, x * nrm };
        }
    }
    // microfacet.h:244-327 (visible normal sampling branch :296-326)
    V3 sample(V3 wi, V2 u, float *pdf_out) const {
        V3 wi_p = normalize(V3(au * wi.x, av * wi.y, wi.z));
        float sin_phi, cos_phi; sincos_phi(wi_p, &sin_phi, &cos_phi);
        float cos_theta = wi_p.z;
        V2 slope = sample_visible_11(cos_theta, u);
        slope = { fmsub(cos_phi, slope.x, sin_phi * slope.y) * au, fmadd(sin_phi, slope.x, cos_phi * slope.y) * av };
        V3 m = normalize(V3(-slope.x, -slope.y, 1.f));
        *pdf_out = eval(m) * smith_g1(wi, m) * fabs_(dot(wi, m)) / wi.z;
        return m;
    }
};

// include/mitsuba/core/math.h legendre_pd (three-term recurrence, double) + quad.h:27-86
static inline void legendre_pd(int l, double x, double *p, double *d) {
    if (l == 0) { *p = 1; *d = 0; return; }
    if (l == 1) { *p = x; *d = 1; return; }
    double l_ppred = 1, l_pred = x, l_cur = 0, d_ppred = 0, d_pred = 1, d_cur = 0, k0 = 3, k1 = 2, k2 = 1;
    for (int k = 2; k <= l; ++k) {
        l_cur = (k0 * x * l_pred - k2 * l_ppred) / k1;
        d_cur = d_ppred + k0 * l_pred;
        l_ppred = l_pred; l_pred = l_cur;
        d_ppred = d_pred; d_pred = d_cur;
        k2 = k1; k0 += 2; k1 += 1;
    }
    *p = l_cur; *d = d_cur;
}
static inline void gauss_legendre(int n, std::vector<float> &nodes, std::vector<float> &weights) {
    nodes.assign(n, 0.f); weights.assign(n, 0.f);
    n--;
    if (n == 0) { nodes[0] = 0; weights[0] = 2; }
    else if (n == 1) { nodes[0] = (float) -std::sqrt(1.0 / 3.0); nodes[1] = -nodes[0]; weights[0] = weights[1] = 1; }
    int m = (n + 1) / 2;
    for (int i = 0; i < m; ++i) {
        double x = -std::cos((double) (2 * i + 1) / (double) (2 * n + 2) * 3.14159265358979323846);
        int it = 0;
        while (true) {
            if (++it > 20) break;
            double p, d; legendre_pd(n + 1, x, &p, &d);
            double step = p / d;
            x -= step;
            if (std::fabs(step) <= 4 * std::fabs(x) * 0x1p-53) break;
        }
        double p, d; legendre_pd(n + 1, x, &p, &d);
        weights[i] = weights[n - i] = (float) (2 / ((1 - x * x) * (d * d)));
        nodes[i] = (float) x; nodes[n - i] = (float) -x;
    }
    if ((n % 2) == 0) {
        double p, d; legendre_pd(n + 1, 0.0, &p, &d);
        weights[n / 2] = (float) (2.0 / (d * d));
        nodes[n / 2] = 0;
    }
}

// microfacet.h:464-560 eval_reflectance / eval_transmittance (quadrature over the unit square).
// meshgrid(nodes,nodes) enumerates x fastest; the reference accumulates in that order per lane.
static inline float mf_eval_reflectance(const Microfacet &d, V3 wi, float eta) {
    int res = eta > 1 ? 32 : 128;
    std::vector<float> nodes, weights; gauss_legendre(res, nodes, weights);
    float result = 0.f;
    for (int jy = 0; jy < res; ++jy) for (int jx = 0; jx < res; ++jx) {
        V2 node = { fmadd(nodes[jx], 0.5f, 0.5f), fmadd(nodes[jy], 0.5f, 0.5f) };
        float pdf; V3 m = d.sample(wi, node, &pdf);
        V3 wo = reflect(wi, m);
        float f, ct, eit, eti; fresnel(dot(wi, m), eta, &f, &ct, &eit, &eti);
        float smith = d.smith_g1(wo, m) * f;
        if (wo.z <= 0.f || wi.z <= 0.f) smith = 0.f;
        result += smith * (weights[jx] * weights[jy]) * 0.25f;
    }
    return result;
}
static inline float mf_eval_transmittance(const Microfacet &d, V3 wi, float eta) {
    int res = eta > 1 ? 32 : 128;
    std::vector<float> nodes, weights; gauss_legendre(res, nodes, weights);
    float result = 0.f;
    for (int jy = 0; jy < res; ++jy) for (int jx = 0; jx < res; ++jx) {
        V2 node = { fmadd(nodes[jx], 0.5f, 0.5f), fmadd(nodes[jy], 0.5f, 0.5f) };
        float pdf; V3 m = d.sample(wi, node, &pdf);
        float f, ct, eit, eti; fresnel(dot(wi, m), eta, &f, &ct, &eit, &eti);
        V3 wo = refract(wi, m, ct, eti);
        float smith = d.smith_g1(wo, m) * (1.f - f);
        if (wo.z * wi.z >= 0.f) smith = 0.f;
        result += smith * (weights[jx] * weights[jy]) * 0.25f;
    }
    return result;
}

} // namespace orc
