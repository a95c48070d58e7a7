// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Thin C-ABI shim around the REFERENCE'S OWN header-only code, compiled from the
// sources where they lie under /root/reference (see oracle/Makefile, target `_ref`).
// Nothing from the reference is copied into this repository: this file only
// #includes the reference headers and forwards arrays to them.
//
// The reference templates are instantiated with Float = drjit::Packet<float, 1>.
// A 1-wide packet takes the reference's *array* code paths (the ones the JIT
// variants llvm_rgb / cuda_rgb take: fmadd-chain dot products, masked selects,
// array branch of dr::erf, ...) while every arithmetic leaf is IEEE-754 binary32
// (1/x, 1/sqrt(x), fmaf) because the file is compiled with `-mfma
// -ffp-contract=off`.  That is the "canonical arithmetic" of SURVEY.md section 0 (F4):
// JIT semantics, IEEE leaves.  The CPU restatement (oracle/amvpt_oracle.cpp) and
// the CUDA kernels are both checked bit-for-bit against these entry points.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <tuple>

#include <mitsuba/core/random.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/core/frame.h>
#include <mitsuba/core/distr_1d.h>
#include <mitsuba/core/distr_2d.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/microfacet.h>
#include <mitsuba/render/sensor.h>
#include <drjit/random.h>
#include <drjit/packet.h>
#include <drjit/color.h>
#include AMVPT_REF_UTIL_H      // src/mitsuba/util.h of the reference (viewer helpers: clip, mod, lerp2, interpolate2d), path given by the Makefile

using namespace mitsuba;
namespace dr = drjit;

using F    = dr::Packet<float, 1>;
using U32  = dr::uint32_array_t<F>;
using V2   = Vector<F, 2>;
using P2   = Point<F, 2>;
using V3   = Vector<F, 3>;
using P3   = Point<F, 3>;
using N3   = Normal<F, 3>;
using Spec = Color<F, 3>;
using T4   = Transform<Point<F, 4>>;
using ST4  = Transform<Point<float, 4>>;
using Fr   = Frame<F>;
using MF   = MicrofacetDistribution<F, Spec>;

// The two symbols below live in libmitsuba-core (src/core/class.cpp, src/core/logger.cpp), which is not built here
// (SURVEY.md F6); the header-only code reaches them only on error paths.  Minimal stand-ins so the shim links.
#include <stdexcept>
namespace mitsuba {
const Class *m_class = nullptr;
namespace detail {
void (Throw)(LogLevel, const Class *, const char *, int, const std::string &msg) { throw std::runtime_error(msg); }
}
}

static inline float g(const F &v) { return v.entry(0); }
static inline V3 ld3(const float *p) { return V3(F(p[0]), F(p[1]), F(p[2])); }
static inline void st3(float *p, const V3 &v) { p[0] = g(v.x()); p[1] = g(v.y()); p[2] = g(v.z()); }

extern "C" {

// ---- RNG: ext/drjit/include/drjit/random.h:59-150, include/mitsuba/core/random.h:76-91
void ref_pcg32_u32(uint64_t initstate, uint64_t initseq, int n, uint32_t *out) {
    dr::PCG32<uint32_t> rng(1, initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = rng.next_uint32();
}
void ref_pcg32_f32(uint64_t initstate, uint64_t initseq, int n, float *out) {
    dr::PCG32<uint32_t> rng(1, initstate, initseq);
    for (int i = 0; i < n; ++i) out[i] = rng.next_float32();
}
void ref_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) {
    auto [a, b] = sample_tea_32<uint32_t>(v0, v1, rounds);
    out[0] = a; out[1] = b;
}
// PCG32Sampler::seed in JIT mode (src/render/sampler.cpp:125-144): per-lane TEA -> (state, stream)
void ref_sampler_stream(uint32_t seed_value, uint32_t lane, int n, float *out) {
    auto [v0, v1] = sample_tea_32<uint32_t>(seed_value, lane);
    dr::PCG32<uint32_t> rng;
    rng.seed((uint64_t) v0, (uint64_t) v1);
    for (int i = 0; i < n; ++i) out[i] = rng.next_float32();
}

// ---- elementary functions, array code path (ext/drjit/include/drjit/math.h)
// op: 0 sin 1 cos 2 exp 3 log 4 erf 5 erfinv 6 exp2 7 sqrt 8 rcp 9 rsqrt 10 safe_sqrt
void ref_math1(int op, const float *in, int n, float *out) {
    for (int i = 0; i < n; ++i) {
        F x(in[i]), r;
        switch (op) {
            case 0: r = dr::sin(x); break;
            case 1: r = dr::cos(x); break;
            case 2: r = dr::exp(x); break;
            case 3: r = dr::log(x); break;
            case 4: r = dr::erf(x); break;
            case 5: r = dr::erfinv(x); break;
            case 6: r = dr::exp2(x); break;
            case 7: r = dr::sqrt(x); break;
            case 8: r = dr::rcp(x); break;
            case 9: r = dr::rsqrt(x); break;
            case 10: r = dr::safe_sqrt(x); break;
            case 12: r = dr::acos(x); break;
            case 13: r = dr::atan2(x, F(0.3f)); break;
            case 14: r = dr::atan2(F(0.3f), x); break;
            case 15: r = dr::atan2(x, F(-0.3f)); break;
            case 16: r = dr::safe_acos(x); break;
            default: r = x;
        }
        out[i] = g(r);
    }
}
void ref_sincos(const float *in, int n, float *s, float *c) {
    for (int i = 0; i < n; ++i) {
        auto [ss, cc] = dr::sincos(F(in[i]));
        s[i] = g(ss); c[i] = g(cc);
    }
}

// ---- Hierarchical2D<Float, 0> (include/mitsuba/core/distr_2d.h:344-775): the importance sampler of the environment map
void ref_hier2d(const float *data, uint32_t sx, uint32_t sy, const float *u, int n, float *uv, float *pdf, float *eval_back) {
    Hierarchical2D<F, 0> h(data, Vector<uint32_t, 2>(sx, sy));
    for (int i = 0; i < n; ++i) {
        auto [p, d] = h.sample(P2(F(u[2 * i]), F(u[2 * i + 1])));
        uv[2 * i] = g(p.x()); uv[2 * i + 1] = g(p.y()); pdf[i] = g(d); eval_back[i] = g(h.eval(p));
    }
}
// ---- DiscreteDistribution<Float> (include/mitsuba/core/distr_1d.h:25-290): the area CDF of Mesh::sample_position (sample_reuse_pmf)
void ref_discrete_sample_reuse(const float *pmf, uint32_t n_pmf, const float *u, int n, uint32_t *index, float *reused, float *pmf_out, float *sum, float *normalization) {
    DiscreteDistribution<F> d(pmf, n_pmf);
    *sum = g(F(d.sum())); *normalization = g(F(d.normalization()));
    for (int i = 0; i < n; ++i) {
        auto [idx, r, p] = d.sample_reuse_pmf(F(u[i]));
        index[i] = idx.entry(0); reused[i] = g(r); pmf_out[i] = g(p);
    }
}
void ref_square_to_uniform_sphere(const float *u, int n, float *out) {
    for (int i = 0; i < n; ++i) st3(out + 3 * i, warp::square_to_uniform_sphere(P2(F(u[2 * i]), F(u[2 * i + 1]))));
}
void ref_square_to_bilinear(const float *v4, const float *u, int n, float *out2, float *pdf, float *pdf_eval) {
    for (int i = 0; i < n; ++i) {
        F a(v4[4 * i]), b(v4[4 * i + 1]), c(v4[4 * i + 2]), d(v4[4 * i + 3]);
        auto [p, q] = warp::square_to_bilinear(a, b, c, d, P2(F(u[2 * i]), F(u[2 * i + 1])));
        out2[2 * i] = g(p.x()); out2[2 * i + 1] = g(p.y()); pdf[i] = g(q); pdf_eval[i] = g(warp::square_to_bilinear_pdf(a, b, c, d, p));
    }
}

// ---- vector helpers (ext/drjit/include/drjit/array_router.h:645-676, array_base.h:597-625)
void ref_dot(const float *a, const float *b, int n, float *out) {
    for (int i = 0; i < n; ++i) out[i] = g(dr::dot(ld3(a + 3 * i), ld3(b + 3 * i)));
}
void ref_cross(const float *a, const float *b, int n, float *out) {
    for (int i = 0; i < n; ++i) st3(out + 3 * i, dr::cross(ld3(a + 3 * i), ld3(b + 3 * i)));
}
void ref_normalize(const float *a, int n, float *out) {
    for (int i = 0; i < n; ++i) st3(out + 3 * i, dr::normalize(ld3(a + 3 * i)));
}
void ref_norm(const float *a, int n, float *out) {
    for (int i = 0; i < n; ++i) out[i] = g(dr::norm(ld3(a + 3 * i)));
}
// include/mitsuba/core/vector.h:116-136
void ref_coordinate_system(const float *nrm, int n, float *s, float *t) {
    for (int i = 0; i < n; ++i) {
        auto [a, b] = coordinate_system(ld3(nrm + 3 * i));
        st3(s + 3 * i, a); st3(t + 3 * i, b);
    }
}

// ---- warps (include/mitsuba/core/warp.h:54-90, 153-156, 412-436)
void ref_square_to_uniform_disk_concentric(const float *u, int n, float *out) {
    for (int i = 0; i < n; ++i) {
        auto p = warp::square_to_uniform_disk_concentric(P2(F(u[2 * i]), F(u[2 * i + 1])));
        out[2 * i] = g(p.x()); out[2 * i + 1] = g(p.y());
    }
}
void ref_square_to_cosine_hemisphere(const float *u, int n, float *out, float *pdf) {
    for (int i = 0; i < n; ++i) {
        V3 v = warp::square_to_cosine_hemisphere(P2(F(u[2 * i]), F(u[2 * i + 1])));
        st3(out + 3 * i, v);
        pdf[i] = g(warp::square_to_cosine_hemisphere_pdf(v));
    }
}
void ref_square_to_uniform_triangle(const float *u, int n, float *out) {
    for (int i = 0; i < n; ++i) {
        auto p = warp::square_to_uniform_triangle(P2(F(u[2 * i]), F(u[2 * i + 1])));
        out[2 * i] = g(p.x()); out[2 * i + 1] = g(p.y());
    }
}

// ---- Fresnel / reflect / refract (include/mitsuba/render/fresnel.h:30-135, 276-300)
void ref_fresnel(const float *cos_i, float eta, int n, float *out4) {
    for (int i = 0; i < n; ++i) {
        auto [r, ct, eit, eti] = fresnel(F(cos_i[i]), F(eta));
        out4[4 * i] = g(r); out4[4 * i + 1] = g(ct); out4[4 * i + 2] = g(eit); out4[4 * i + 3] = g(eti);
    }
}
void ref_fresnel_conductor(const float *cos_i, float eta, float k, int n, float *out) {
    for (int i = 0; i < n; ++i)
        out[i] = g(fresnel_conductor(F(cos_i[i]), dr::Complex<F>(F(eta), F(k))));
}
void ref_reflect_m(const float *wi, const float *m, int n, float *out) {
    for (int i = 0; i < n; ++i) st3(out + 3 * i, reflect(ld3(wi + 3 * i), N3(ld3(m + 3 * i))));
}
void ref_refract_m(const float *wi, const float *m, const float *cos_t, const float *eta_ti, int n, float *out) {
    for (int i = 0; i < n; ++i)
        st3(out + 3 * i, refract(ld3(wi + 3 * i), N3(ld3(m + 3 * i)), F(cos_t[i]), F(eta_ti[i])));
}

// ---- microfacet distribution (include/mitsuba/render/microfacet.h:185-420), sample_visible = true
void ref_mf_eval(int type, float au, float av, const float *m, int n, float *out) {
    MF d((MicrofacetType) type, F(au), F(av), true);
    for (int i = 0; i < n; ++i) out[i] = g(d.eval(ld3(m + 3 * i)));
}
void ref_mf_smith_g1(int type, float au, float av, const float *v, const float *m, int n, float *out) {
    MF d((MicrofacetType) type, F(au), F(av), true);
    for (int i = 0; i < n; ++i) out[i] = g(d.smith_g1(ld3(v + 3 * i), ld3(m + 3 * i)));
}
void ref_mf_pdf(int type, float au, float av, const float *wi, const float *m, int n, float *out) {
    MF d((MicrofacetType) type, F(au), F(av), true);
    for (int i = 0; i < n; ++i) out[i] = g(d.pdf(ld3(wi + 3 * i), ld3(m + 3 * i)));
}
void ref_mf_sample(int type, float au, float av, const float *wi, const float *u, int n, float *m_out, float *pdf) {
    MF d((MicrofacetType) type, F(au), F(av), true);
    for (int i = 0; i < n; ++i) {
        auto [m, p] = d.sample(ld3(wi + 3 * i), P2(F(u[2 * i]), F(u[2 * i + 1])));
        st3(m_out + 3 * i, V3(m)); pdf[i] = g(p);
    }
}
// roughplastic table (src/bsdfs/roughplastic.cpp:221-252): reference quadrature, reference packet width
void ref_rough_tables(int type, float alpha, float eta, int res, float *ext_transmittance, float *int_reflectance) {
    using FloatX = dr::DynamicArray<float>;
    using Vector3fX = Vector<FloatX, 3>;
    using FloatP = dr::Packet<float>;
    MicrofacetDistribution<FloatP, Color<FloatP, 3>> distr((MicrofacetType) type, alpha);
    FloatX mu = dr::maximum(1e-6f, dr::linspace<FloatX>(0, 1, res));
    FloatX zero = dr::zeros<FloatX>(res);
    Vector3fX wi = Vector3fX(dr::sqrt(1 - mu * mu), zero, mu);
    auto t = eval_transmittance(distr, wi, eta);
    for (int i = 0; i < res; ++i) ext_transmittance[i] = t[i];
    *int_reflectance = dr::mean(eval_reflectance(distr, wi, 1.f / eta) * wi.z()) * 2.f;
}
void ref_gauss_legendre(int n, float *nodes, float *weights) {
    auto [a, b] = quad::gauss_legendre<dr::DynamicArray<float>>(n);
    for (int i = 0; i < n; ++i) { nodes[i] = a[i]; weights[i] = b[i]; }
}

// ---- transforms (include/mitsuba/core/transform.h, ext/drjit/include/drjit/matrix.h:105-120, 440-503)
static void st44(float *o, const dr::Matrix<float, 4> &m) {
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) o[4 * r + c] = m(r, c);
}
static dr::Matrix<float, 4> ld44(const float *p) {
    dr::Matrix<float, 4> m;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m(r, c) = p[4 * r + c];
    return m;
}
void ref_look_at(const float *origin, const float *target, const float *up, float *mat, float *inv_t) {
    ST4 t = ST4::look_at(Point<float, 3>(origin[0], origin[1], origin[2]), Point<float, 3>(target[0], target[1], target[2]),
                         Vector<float, 3>(up[0], up[1], up[2]));
    st44(mat, t.matrix); st44(inv_t, t.inverse_transpose);
}
void ref_mat_inverse(const float *m, float *out) { st44(out, dr::inverse(ld44(m))); }
void ref_mat_inverse_transpose(const float *m, float *out) { st44(out, dr::inverse_transpose(ld44(m))); }
void ref_mat_mul(const float *a, const float *b, float *out) { st44(out, ld44(a) * ld44(b)); }

// PerspectiveCamera::update_camera_transforms (src/sensors/perspective.cpp:173-181) on reference
// perspective_projection (include/mitsuba/render/sensor.h:320-356): returns sample_to_camera.matrix
void ref_sample_to_camera(int film_w, int film_h, float x_fov, float near_clip, float far_clip, float lens_shift,
                          float *camera_to_sample, float *sample_to_camera) {
    Vector<int, 2> size(film_w, film_h), off(0, 0);
    ST4 c2s = perspective_projection<float>(size, size, off, x_fov, near_clip, far_clip);
    c2s.matrix.entry(0, 2) += lens_shift;
    c2s.inverse_transpose = dr::inverse_transpose(c2s.matrix);
    ST4 s2c = c2s.inverse();
    st44(camera_to_sample, c2s.matrix); st44(sample_to_camera, s2c.matrix);
}
double ref_parse_fov_axis(double fov, int axis /*0 x, 1 y*/, double aspect) {
    // src/render/sensor.cpp:177-231 ("x" and "y" branches)
    if (axis == 0) return fov;
    return dr::rad_to_deg(2.0 * dr::atan(dr::tan(0.5 * dr::deg_to_rad(fov)) * aspect));
}
// GridSensor ctor, line placement branch (src/sensors/grid.cpp:208-226): sub-camera to_world matrices
void ref_grid_to_world(const float *to_world, const float *to_world_inv_t, const float *offset, float *out_mat) {
    auto itrafo = ld44(to_world_inv_t);
    itrafo.entry(3, 0) += offset[0];
    itrafo.entry(3, 1) += offset[1];
    itrafo.entry(3, 2) += offset[2];
    auto ntrafo = dr::inverse(dr::transpose(itrafo));
    (void) to_world;
    st44(out_mat, ntrafo);
}
// PerspectiveCamera::sample_ray (src/sensors/perspective.cpp:205-241), reference Transform operators
void ref_perspective_sample_ray(const float *sample_to_camera, const float *to_world, float near_clip, float far_clip,
                                const float *pos, int n, float *o, float *d, float *maxt) {
    T4 s2c; T4 tw;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) {
        s2c.matrix(r, c) = F(sample_to_camera[4 * r + c]);
        tw.matrix(r, c)  = F(to_world[4 * r + c]);
    }
    for (int i = 0; i < n; ++i) {
        P3 near_p = s2c * P3(F(pos[2 * i]), F(pos[2 * i + 1]), F(0.f));
        V3 dd = dr::normalize(V3(near_p));
        P3 oo = tw.translation();
        V3 dw = tw * dd;
        F inv_z = dr::rcp(dd.z());
        F near_t = F(near_clip) * inv_z, far_t = F(far_clip) * inv_z;
        oo += dw * near_t;
        st3(o + 3 * i, V3(oo)); st3(d + 3 * i, dw); maxt[i] = g(far_t - near_t);
    }
}

// Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:467-488) restated on reference vector ops
void ref_moeller_trumbore(const float *o, const float *d, const float *maxt, const float *p0, const float *p1,
                          const float *p2, int n, float *tuv, int *hit) {
    for (int i = 0; i < n; ++i) {
        V3 ro = ld3(o + 3 * i), rd = ld3(d + 3 * i), a = ld3(p0 + 3 * i), b = ld3(p1 + 3 * i), c = ld3(p2 + 3 * i);
        V3 e1 = b - a, e2 = c - a;
        V3 pvec = dr::cross(rd, e2);
        F inv_det = dr::rcp(dr::dot(e1, pvec));
        V3 tvec = ro - a;
        F u = dr::dot(tvec, pvec) * inv_det;
        auto active = u >= 0.f && u <= 1.f;
        V3 qvec = dr::cross(tvec, e1);
        F v = dr::dot(rd, qvec) * inv_det;
        active &= v >= 0.f && u + v <= 1.f;
        F t = dr::dot(e2, qvec) * inv_det;
        active &= t >= 0.f && t <= F(maxt[i]);
        tuv[3 * i] = g(t); tuv[3 * i + 1] = g(u); tuv[3 * i + 2] = g(v);
        hit[i] = active.entry(0) ? 1 : 0;
    }
}
// Frame::to_local / to_world (include/mitsuba/core/frame.h:34-41)
void ref_frame(const float *s, const float *t, const float *nn, const float *v, int n, float *loc, float *wld) {
    for (int i = 0; i < n; ++i) {
        Fr f; f.s = ld3(s + 3 * i); f.t = ld3(t + 3 * i); f.n = ld3(nn + 3 * i);
        st3(loc + 3 * i, f.to_local(ld3(v + 3 * i)));
        st3(wld + 3 * i, f.to_world(ld3(v + 3 * i)));
    }
}
float ref_const(int which) {
    switch (which) {
        case 0: return math::RayEpsilon<float>;
        case 1: return math::ShadowEpsilon<float>;
        case 2: return dr::Largest<float>;
        case 3: return dr::Epsilon<float>;
        case 4: return dr::InvPi<float>;
        case 5: return dr::Pi<float>;
    }
    return 0.f;
}

// ---- display step leaves: dr::linear_to_srgb (ext/drjit/include/drjit/color.h:19-53), viewer helpers (src/mitsuba/util.h:47-49, 104-128)
void ref_linear_to_srgb(const float *in, int n, float *out) {
    for (int i = 0; i < n; ++i) { F x(in[i]); F r = dr::linear_to_srgb(x); out[i] = r[0]; }
}
void ref_clip_mod(const float *x, const float *y, int n, float *out_clip, float *out_mod) {
    for (int i = 0; i < n; ++i) { out_clip[i] = mitsuba::clip(x[i]); out_mod[i] = mitsuba::mod(x[i], y[i]); }
}
void ref_interpolate2d_u8(const uint8_t *data, uint32_t w, uint32_t h, uint32_t nch, const float *xy, int n, uint32_t c, int nearest, uint8_t *out) {
    for (int i = 0; i < n; ++i)
        out[i] = mitsuba::interpolate2d<uint8_t>(data, mitsuba::vec2f(xy[2 * i], xy[2 * i + 1]), mitsuba::vec2u(w, h), c, nch, nearest != 0);
}

} // extern "C"
