// oracle/o_math.h -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's leaf
// arithmetic in the canonical form of SURVEY.md sec.0 F4: the reference's *array* (JIT)
// code paths with IEEE-754 binary32 leaves.  Compile with -mfma -ffp-contract=off.
// Every function cites the reference file:line it follows (relative to /root/reference).
// Pinned bit-for-bit against oracle/_ref (the reference's own headers) by tests/test_oracle_vs_ref.py.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

namespace orc {

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// ext/drjit/include/drjit/array_utils.h:200-219 (fmadd_/rcp_/rsqrt_ scalar leaves)
static inline float fmadd(float a, float b, float c) { return std::fmaf(a, b, c); }
static inline float fmsub(float a, float b, float c) { return std::fmaf(a, b, -c); }
static inline float fnmadd(float a, float b, float c) { return std::fmaf(-a, b, c); }
static inline float rcp(float a) { return 1.f / a; }
static inline float rsqrt(float a) { return 1.f / std::sqrt(a); }
static inline float sqr(float a) { return a * a; }
static inline float fabs_(float a) { return u2f(f2u(a) & 0x7fffffffu); }
// array_utils.h:192-198: maximum_(a,b) = a < b ? b : a ; minimum_(a,b) = b < a ? b : a
static inline float fmax_(float a, float b) { return a < b ? b : a; }
static inline float fmin_(float a, float b) { return b < a ? b : a; }
// ext/drjit/include/drjit/array_router.h:386-423
static inline float mulsign(float a, float b) { return u2f(f2u(a) ^ (f2u(b) & 0x80000000u)); }
static inline float mulsign_neg(float a, float b) { return u2f(f2u(a) ^ (~f2u(b) & 0x80000000u)); }
static inline float copysign_(float a, float b) { return u2f((f2u(a) & 0x7fffffffu) | (f2u(b) & 0x80000000u)); }
static inline float sign_(float v) { return copysign_(1.f, v); }
static inline float safe_sqrt(float a) { return std::sqrt(fmax_(a, 0.f)); }      // array_router.h:1729-1731
static inline float lerp(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); } // :350-352
static inline float clip(float v, float lo, float hi) { return fmax_(fmin_(v, hi), lo); }    // :355-358

constexpr float Pi = 3.14159265358979323846f, InvPi = 0.31830988618379067154f;
constexpr float InvSqrtPi = 0.56418958354775628695f, InvSqrtTwo = 0.70710678118654752440f;
constexpr float InvLogTwo = 1.44269504088896340736f;
constexpr float Epsilon = 0x1p-24f;                       // drjit/array_constants.h:52
constexpr float RayEpsilon = Epsilon * 1500;              // include/mitsuba/core/math.h:17-21
constexpr float ShadowEpsilon = RayEpsilon * 10;          // :22
constexpr float Largest = std::numeric_limits<float>::max();
constexpr float Inf = std::numeric_limits<float>::infinity();

struct V2 { float x, y; };
struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit V3(float a) : x(a), y(a), z(a) {}
};
static inline V3 operator+(V3 a, V3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
static inline V3 operator-(V3 a, V3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
static inline V3 operator*(V3 a, V3 b) { return { a.x * b.x, a.y * b.y, a.z * b.z }; }
static inline V3 operator*(V3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
// drjit/array_router.h:218-232: (array of depth 2) / (Float) is lowered to a * rcp(b)
static inline V3 operator/(V3 a, float s) { float r = rcp(s); return { a.x * r, a.y * r, a.z * r }; }
static inline V3 operator-(V3 a) { return { -a.x, -a.y, -a.z }; }
static inline V3 fmadd(V3 a, float b, V3 c) { return { fmadd(a.x, b, c.x), fmadd(a.y, b, c.y), fmadd(a.z, b, c.z) }; }
static inline V3 fmadd(V3 a, V3 b, V3 c) { return { fmadd(a.x, b.x, c.x), fmadd(a.y, b.y, c.y), fmadd(a.z, b.z, c.z) }; }

// drjit/array_base.h:597-625 (array branch): result = a0*b0; result = fmadd(ai, bi, result)
static inline float dot(V3 a, V3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
// drjit/array_router.h:645-654 (depth > 1 branch)
static inline float squared_norm(V3 v) { return fmadd(v.z, v.z, fmadd(v.y, v.y, v.x * v.x)); }
static inline float squared_norm(V2 v) { return fmadd(v.y, v.y, v.x * v.x); }
static inline float norm(V3 v) { return std::sqrt(squared_norm(v)); }
static inline V3 normalize(V3 v) { return v * rsqrt(squared_norm(v)); }             // :660-662
// drjit/array_router.h:664-676 (x86 branch): fmsub(v1.yzx, v2.zxy, v1.zxy * v2.yzx)
static inline V3 cross(V3 a, V3 b) {
    return { fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x) };
}
static inline float max3(V3 v) { return fmax_(fmax_(v.x, v.y), v.z); }   // dr::max = horizontal maximum chain

// drjit/math.h:26-41 estrin_impl (recursion unrolled into a loop; identical operation order)
static inline float estrin(float x, const float *c, int n) {
    float buf[16];
    for (int i = 0; i < n; ++i) buf[i] = c[i];
    while (n > 1) {
        int n_fma = n / 2, n_rec = (n - 1) / 2;
        for (int i = 0; i < n_fma; ++i) buf[i] = fmadd(x, buf[2 * i + 1], buf[2 * i]);
        if (n_rec == n_fma) buf[n_rec] = buf[n - 1];
        n = n_rec + 1;
        x = x * x;
    }
    return buf[0];
}

// drjit/math.h:76-175 detail::sincos<true,true> (Single branch)
static inline void sincos(float x, float *s_out, float *c_out) {
    float xa = fabs_(x);
    int32_t j = (int32_t) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    uint32_t sign_sin = ((uint32_t) j << 29) ^ f2u(x);
    uint32_t sign_cos = (uint32_t) (~(j - 2)) << 29;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == Inf) z = u2f(f2u(z) | 0xffffffffu);
    static const float cs[3] = { -1.6666654611e-1f, 8.3321608736e-3f, -1.9515295891e-4f };
    static const float cc[3] = { 4.166664568298827e-2f, -1.388731625493765e-3f, 2.443315711809948e-5f };
    float s = estrin(z, cs, 3) * z, c = estrin(z, cc, 3) * z;
    s = fmadd(s, y, y);
    c = fmadd(c, z, fmadd(z, -0.5f, 1.f));
    bool polymask = (j & 2) == 0;
    *s_out = mulsign(polymask ? s : c, u2f(sign_sin));
    *c_out = mulsign(polymask ? c : s, u2f(sign_cos));
}

// drjit/math.h:615-632 ldexp (generic branch)
static inline float ldexp_(float x, float n) { return x * u2f((uint32_t) (((int32_t) n + 0x7f) << 23)); }

// drjit/math.h:690-760 exp (Single branch)
static inline float exp_(float x) {
    bool over = x > 88.3762588501f, under = x < -88.3762588501f;
    float n = std::floor(fmadd(InvLogTwo, x, 0.5f));
    float y = x;
    y = fmadd(n, -0.693359375f, y);
    y = fmadd(n, 2.12194440e-4f, y);
    static const float c[6] = { 5.0000001201e-1f, 1.6666665459e-1f, 4.1665795894e-2f,
                                8.3334519073e-3f, 1.3981999507e-3f, 1.9875691500e-4f };
    float z = estrin(y, c, 6);
    z = fmadd(z, y * y, y + 1.f);
    return over ? Inf : (under ? 0.f : ldexp_(z, n));
}

// drjit/math.h:763-830 exp2 (Single branch)
static inline float exp2_(float x) {
    bool over = x > 127.f, under = x < -127.f;
    float n = std::floor(x), y = x - n;
    if (y > .5f) { n += 1.f; y -= 1.f; }
    static const float c[6] = { 6.931472028550421e-1f, 2.402264791363012e-1f, 5.550332471162809e-2f,
                                9.618437357674640e-3f, 1.339887440266574e-3f, 1.535336188319500e-4f };
    float z = estrin(y, c, 6);
    z = fmadd(y, z, 1.f);
    return over ? Inf : (under ? 0.f : ldexp_(z, n));
}

// drjit/math.h:639-683 frexp + :833-905 log (Single branch)
static inline float log_(float x) {
    bool valid = x >= 0.f;
    uint32_t xi = f2u(x), ebits = xi & 0x7f800000u;
    bool is_normal = (x != 0.f) && (ebits != 0x7f800000u);
    int32_t ei = is_normal ? ((int32_t) (ebits >> 23) - 0x7f) : 0;
    uint32_t mant = (xi & ~0x7f800000u) | f2u(.5f);
    float xm = is_normal ? u2f(mant) : x, e = (float) ei;
    bool ge = xm >= InvSqrtTwo;
    if (ge) e += 1.f;
    xm += (ge ? 0.f : xm) - 1.f;
    static const float c[9] = { 3.3333331174e-1f, -2.4999993993e-1f, 2.0000714765e-1f, -1.6668057665e-1f,
                                1.4249322787e-1f, -1.2420140846e-1f, 1.1676998740e-1f, -1.1514610310e-1f,
                                7.0376836292e-2f };
    float y = estrin(xm, c, 9);
    float z = xm * xm;
    y *= xm * z;
    y = fmadd(e, -2.121944400546905827679e-4f, y);
    float r = xm + fmadd(-.5f, z, y);
    r = fmadd(e, 0.693359375f, r);
    if (x == Inf) r = Inf;
    if (x == 0.f) r = -Inf;
    return valid ? r : u2f(0xffffffffu);
}

// drjit/math.h:1497-1551 erf -- ARRAY branch: `std::is_same_v<Value, float>` is false for JIT/packet
// arrays, so the double-precision coefficient sets are used, rounded to float.
static inline float erf_(float x) {
    float xa = fabs_(x), x2 = x * x;
    static const float a0[12] = {
        (float) 0x1.20dd750429b6dp+0,  (float) -0x1.812746b0379bcp-2, (float) 0x1.ce2f21a040d12p-4,
        (float) -0x1.b82ce311fa924p-6, (float) 0x1.565bccf92b298p-8,  (float) -0x1.c02db03dd71b8p-11,
        (float) 0x1.f9a2baa8fee07p-14, (float) -0x1.f4ca4d6f3e31bp-17, (float) 0x1.b97fd3d992af4p-20,
        (float) -0x1.5c0726f04e805p-23, (float) 0x1.d71b0f1b15b0ap-27, (float) -0x1.abae491c540bp-31 };
    static const float a1[8] = {
        (float) -0x1.a0be83b09c3d7p+0, (float) -0x1.8bb29648c7afep+1, (float) -0x1.639eb89a5975p+1,
        (float) -0x1.7b48b8cd14d9fp+0, (float) -0x1.fb25a03ddc781p-2, (float) -0x1.9cdb7dcacdfb3p-4,
        (float) -0x1.64f7fbe544f07p-7, (float) -0x1.9a3c3874b3919p-12 };
    static const float a2[8] = {
        (float) 0x1p+0, (float) 0x1.55b5d06f1c2dep+0, (float) 0x1.b998ffae5528ep-1, (float) 0x1.46884d56cb49bp-2,
        (float) 0x1.18e8848a2cc38p-4, (float) 0x1.ee7f90e8d480cp-8, (float) 0x1.1c6a194029df4p-12,
        (float) -0x1.03d1306b29028p-31 };
    float c0 = estrin(x2, a0, 12);
    float c1 = estrin(xa, a1, 8) / estrin(xa, a2, 8);
    float xb = 1.f - exp2_(c1 * xa);
    return xa < 1.f ? x * c0 : copysign_(std::isfinite(xb) ? xb : 1.f, x);
}

// drjit/math.h:1557-1578 erfinv
static inline float erfinv_(float x) {
    float w = -log_((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = std::sqrt(w) - 3.f;
    static const float p1[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                                 -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    static const float p2[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                                 -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    return (w < 5.f ? estrin(w1, p1, 9) : estrin(w2, p2, 9)) * x;
}

// include/mitsuba/core/vector.h:116-136
static inline void coordinate_system(V3 n, V3 *s, V3 *t) {
    float sign = sign_(n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    *s = V3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    *t = V3(b, fmadd(n.y, n.y * a, sign), -n.y);
}

// include/mitsuba/core/warp.h:54-90
static inline V2 square_to_uniform_disk_concentric(V2 u) {
    float x = fmsub(2.f, u.x, 1.f), y = fmsub(2.f, u.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabs_(x) < fabs_(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * Pi * rp / r;
    if (q13) phi = 0.5f * Pi - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincos(phi, &s, &c);
    return { r * c, r * s };
}
// warp.h:412-436
static inline V3 square_to_cosine_hemisphere(V2 u) {
    V2 p = square_to_uniform_disk_concentric(u);
    return { p.x, p.y, safe_sqrt(1.f - squared_norm(p)) };
}
static inline float square_to_cosine_hemisphere_pdf(V3 v) { return InvPi * v.z; }
// warp.h:153-156
static inline V2 square_to_uniform_triangle(V2 u) {
    float t = safe_sqrt(1.f - u.x);
    return { 1.f - t, t * u.y };
}

// include/mitsuba/render/fresnel.h:30-73
static inline void fresnel(float cos_i, float eta, float *r_out, float *cos_t, float *eta_it_o, float *eta_ti_o) {
    bool outside = cos_i >= 0.f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_t_sqr = fnmadd(fnmadd(cos_i, cos_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabs_(cos_i), ct = safe_sqrt(cos_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || (ci == 0.f);
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fnmadd(eta_it, ct, ci) / fmadd(eta_it, ct, ci);
    float a_p = fnmadd(eta_it, ci, ct) / fmadd(eta_it, ci, ct);
    float r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = r_sc;
    *r_out = r; *cos_t = mulsign_neg(ct, cos_i); *eta_it_o = eta_it; *eta_ti_o = eta_ti;
}
// fresnel.h:93-118
static inline float fresnel_conductor(float cos_i, float eta_r, float eta_i) {
    float c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - s2;
    float a2pb2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    float a = safe_sqrt(.5f * (a2pb2 + temp_1));
    float term_1 = a2pb2 + c2, term_2 = 2.f * cos_i * a;
    float r_s = (term_1 - term_2) / (term_1 + term_2);
    float term_3 = a2pb2 * c2 + s4, term_4 = term_2 * s2;
    float r_p = r_s * (term_3 - term_4) / (term_3 + term_4);
    return 0.5f * (r_s + r_p);
}
// fresnel.h:276-300
static inline V3 reflect(V3 wi) { return { -wi.x, -wi.y, wi.z }; }
static inline V3 reflect(V3 wi, V3 m) {
    float k = 2.f * dot(wi, m);
    return { fmsub(m.x, k, wi.x), fmsub(m.y, k, wi.y), fmsub(m.z, k, wi.z) };
}
static inline V3 refract(V3 wi, float cos_t, float eta_ti) { return { -eta_ti * wi.x, -eta_ti * wi.y, cos_t }; }
static inline V3 refract(V3 wi, V3 m, float cos_t, float eta_ti) {
    float k = fmadd(dot(wi, m), eta_ti, cos_t);
    return { fmsub(m.x, k, wi.x * eta_ti), fmsub(m.y, k, wi.y * eta_ti), fmsub(m.z, k, wi.z * eta_ti) };
}

} // namespace orc
