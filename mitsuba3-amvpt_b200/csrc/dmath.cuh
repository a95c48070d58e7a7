// dmath.cuh -- leaf arithmetic of the AMVPT render path for sm_100a device code (and, through the
// AM_HD macro, for the host-side scene preparation that must produce the very same bits).
//
// Canonical arithmetic (DESIGN.md "Arithmetic contract"): the reference's array/JIT code paths with
// IEEE-754 binary32 leaves.  This translation unit is compiled with `-fmad=false`; a fused
// multiply-add happens only where the reference source says dr::fmadd / fmsub / fnmadd, and
// reciprocal / division / square root are the correctly rounded __frcp_rn / __fdiv_rn / __fsqrt_rn
// (the reference's own CUDA variant uses the .approx.ftz forms, its LLVM variant uses fdiv; the
// IEEE forms are the ones its CPU variants and our parity oracle share).
// Reference files are cited relative to the upstream tree (xacond00/mitsuba3-amvpt).
#pragma once
#include <cstdint>
#include <cmath>
#include <cstring>

#if defined(__CUDACC__)
#define AM_HD __host__ __device__ __forceinline__
#define AM_D __device__ __forceinline__
#else
#define AM_HD inline
#define AM_D inline
#endif

namespace amvpt {

#if defined(__CUDA_ARCH__)
AM_HD float fmadd(float a, float b, float c) { return __fmaf_rn(a, b, c); }
AM_HD float rcp(float a) { return __frcp_rn(a); }
AM_HD float fsqrt(float a) { return __fsqrt_rn(a); }
AM_HD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
AM_HD uint32_t f2u(float f) { return __float_as_uint(f); }
AM_HD float u2f(uint32_t u) { return __uint_as_float(u); }
#else
AM_HD float fmadd(float a, float b, float c) { return fmaf(a, b, c); }
AM_HD float rcp(float a) { return 1.f / a; }
AM_HD float fsqrt(float a) { return sqrtf(a); }
AM_HD float fdiv(float a, float b) { return a / b; }
AM_HD uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
AM_HD float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
#endif

// ext/drjit/include/drjit/array_utils.h:192-219
AM_HD float fmsub(float a, float b, float c) { return fmadd(a, b, -c); }
AM_HD float fnmadd(float a, float b, float c) { return fmadd(-a, b, c); }
AM_HD float rsqrt(float a) { return rcp(fsqrt(a)); }
AM_HD float sqr(float a) { return a * a; }
AM_HD float fabs_(float a) { return u2f(f2u(a) & 0x7fffffffu); }
AM_HD float max_(float a, float b) { return a < b ? b : a; }
AM_HD float min_(float a, float b) { return b < a ? b : a; }
// ext/drjit/include/drjit/array_router.h:386-423, 350-358, 1729-1731
AM_HD float mulsign(float a, float b) { return u2f(f2u(a) ^ (f2u(b) & 0x80000000u)); }
AM_HD float mulsign_neg(float a, float b) { return u2f(f2u(a) ^ (~f2u(b) & 0x80000000u)); }
AM_HD float copysign_(float a, float b) { return u2f((f2u(a) & 0x7fffffffu) | (f2u(b) & 0x80000000u)); }
AM_HD float safe_sqrt(float a) { return fsqrt(max_(a, 0.f)); }
AM_HD float lerp_(float a, float b, float t) { return fmadd(b, t, fnmadd(a, t, a)); }
AM_HD float clip_(float v, float lo, float hi) { return max_(min_(v, hi), lo); }
AM_HD bool finite_(float v) { return (f2u(v) & 0x7f800000u) != 0x7f800000u; }

#define AM_PI 3.14159265358979323846f
#define AM_INV_PI 0.31830988618379067154f
#define AM_INV_SQRT_PI 0.56418958354775628695f
#define AM_INV_SQRT_TWO 0.70710678118654752440f
#define AM_INV_LOG_TWO 1.44269504088896340736f
#define AM_EPSILON 5.9604644775390625e-08f            /* 2^-24, drjit/array_constants.h:52 */
#define AM_RAY_EPSILON (AM_EPSILON * 1500.f)          /* include/mitsuba/core/math.h:17-21 */
#define AM_SHADOW_EPSILON (AM_RAY_EPSILON * 10.f)     /* :22 */
#define AM_LARGEST 3.402823466e+38f
#define AM_INF u2f(0x7f800000u)

struct V2 { float x, y; };
struct V3 { float x, y, z; };
AM_HD V3 v3(float a, float b, float c) { V3 r; r.x = a; r.y = b; r.z = c; return r; }
AM_HD V3 v3(float a) { return v3(a, a, a); }
AM_HD V2 v2(float a, float b) { V2 r; r.x = a; r.y = b; return r; }
AM_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
AM_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
AM_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
AM_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
AM_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
// drjit/array_router.h:218-232: (nested array) / (scalar lane) is lowered to a * rcp(b)
AM_HD V3 operator/(V3 a, float s) { float r = rcp(s); return v3(a.x * r, a.y * r, a.z * r); }
AM_HD V3 fmadd(V3 a, float b, V3 c) { return v3(fmadd(a.x, b, c.x), fmadd(a.y, b, c.y), fmadd(a.z, b, c.z)); }
AM_HD V3 fmadd(V3 a, V3 b, V3 c) { return v3(fmadd(a.x, b.x, c.x), fmadd(a.y, b.y, c.y), fmadd(a.z, b.z, c.z)); }
AM_HD V3 select3(bool m, V3 v) { return m ? v : v3(0.f); }
// drjit/array_base.h:597-625: a0*b0, then an fmadd chain
AM_HD float dot(V3 a, V3 b) { return fmadd(a.z, b.z, fmadd(a.y, b.y, a.x * b.x)); }
AM_HD float squared_norm(V3 v) { return fmadd(v.z, v.z, fmadd(v.y, v.y, v.x * v.x)); }   // array_router.h:645-654
AM_HD float squared_norm(V2 v) { return fmadd(v.y, v.y, v.x * v.x); }
AM_HD float norm(V3 v) { return fsqrt(squared_norm(v)); }
AM_HD V3 normalize(V3 v) { return v * rsqrt(squared_norm(v)); }                         // :660-662
// array_router.h:664-676: fmsub(a.yzx, b.zxy, a.zxy * b.yzx)
AM_HD V3 cross(V3 a, V3 b) {
    return v3(fmsub(a.y, b.z, a.z * b.y), fmsub(a.z, b.x, a.x * b.z), fmsub(a.x, b.y, a.y * b.x));
}
AM_HD float hmax(V3 v) { return max_(max_(v.x, v.y), v.z); }

// ---------------------------------------------------------------------------------------------
// Elementary functions: Dr.Jit's own polynomial kernels (ext/drjit/include/drjit/math.h), which is
// what every variant of the reference evaluates unless its JIT substitutes an .approx instruction.
// Estrin evaluation order follows math.h:26-41 (pairwise fmadd, squaring x per level).
AM_HD float estrin3(float x, float c0, float c1, float c2) {
    float x2 = x * x;
    return fmadd(x2, c2, fmadd(x, c1, c0));
}
AM_HD float estrin6(float x, const float *c) {
    float x2 = x * x;
    float a = fmadd(x, c[1], c[0]), b = fmadd(x, c[3], c[2]), d = fmadd(x, c[5], c[4]);
    float x4 = x2 * x2;
    return fmadd(x4, d, fmadd(x2, b, a));
}
AM_HD float estrin8(float x, const float *c) {
    float x2 = x * x;
    float a = fmadd(x, c[1], c[0]), b = fmadd(x, c[3], c[2]), d = fmadd(x, c[5], c[4]), e = fmadd(x, c[7], c[6]);
    float x4 = x2 * x2;
    float f = fmadd(x2, b, a), g = fmadd(x2, e, d);
    return fmadd(x4, g, f);
}
AM_HD float estrin9(float x, const float *c) {
    float x2 = x * x;
    float a = fmadd(x, c[1], c[0]), b = fmadd(x, c[3], c[2]), d = fmadd(x, c[5], c[4]), e = fmadd(x, c[7], c[6]), t = c[8];
    float x4 = x2 * x2;
    float f = fmadd(x2, b, a), g = fmadd(x2, e, d);     // level 2: 5 values -> (f, g, t)
    float x8 = x4 * x4;
    float h = fmadd(x4, g, f);                          // level 3: 3 values -> (h, t)
    return fmadd(x8, t, h);
}
AM_HD float estrin10(float x, const float *c) {
    float x2 = x * x;
    float r0 = fmadd(x, c[1], c[0]), r1 = fmadd(x, c[3], c[2]), r2 = fmadd(x, c[5], c[4]), r3 = fmadd(x, c[7], c[6]), r4 = fmadd(x, c[9], c[8]);
    float x4 = x2 * x2;
    float q0 = fmadd(x2, r1, r0), q1 = fmadd(x2, r3, r2);      // level 2: 5 values -> (q0, q1, r4)
    float x8 = x4 * x4;
    return fmadd(x8, r4, fmadd(x4, q1, q0));                   // level 3: 3 values -> (s0, r4); level 4
}
AM_HD float estrin12(float x, const float *c) {
    float x2 = x * x;
    float p0 = fmadd(x, c[1], c[0]), p1 = fmadd(x, c[3], c[2]), p2 = fmadd(x, c[5], c[4]),
          p3 = fmadd(x, c[7], c[6]), p4 = fmadd(x, c[9], c[8]), p5 = fmadd(x, c[11], c[10]);
    float x4 = x2 * x2;
    float q0 = fmadd(x2, p1, p0), q1 = fmadd(x2, p3, p2), q2 = fmadd(x2, p5, p4);
    float x8 = x4 * x4;
    float r0 = fmadd(x4, q1, q0);
    return fmadd(x8, q2, r0);
}

// generic form of math.h:26-41 estrin_impl (pairwise fmadd per level, an odd tail coefficient is carried up unchanged)
template <int N> AM_HD float estrin_n(float x, const float *c) {
    float v[N]; for (int i = 0; i < N; ++i) v[i] = c[i];
    int n = N;
    while (n > 1) {
        const int n_fma = n / 2, n_rec = (n - 1) / 2;
        for (int i = 0; i < n_fma; ++i) v[i] = fmadd(x, v[2 * i + 1], v[2 * i]);
        if (n_rec == n_fma) v[n_rec] = v[n - 1];
        n = n_rec + 1; x = x * x;
    }
    return v[0];
}
// math.h:458-510 acos (single precision branch) and array_router.h:1775 safe_acos
AM_HD float acos_(float x) {
    const float xa = fabs_(x), x2 = x * x;
    const bool big = xa > 0.5f;
    const float x1 = 0.5f * (1.f - xa), x3 = big ? x1 : x2, x4 = big ? fsqrt(x1) : xa;
    const float c[5] = { 1.666675242e-1f, 7.4953002686e-2f, 4.5470025998e-2f, 2.4181311049e-2f, 4.2163199048e-2f };
    float z1 = estrin_n<5>(x3, c);
    z1 = fmadd(z1, x3 * x4, x4);
    float z2 = z1 + z1;
    z2 = x < 0.f ? AM_PI - z2 : z2;
    const float z3 = (AM_PI * .5f) - copysign_(z1, x);
    return big ? z2 : z3;
}
AM_HD float safe_acos_(float x) { return acos_(clip_(x, -1.f, 1.f)); }
// math.h:512-597 atan2 (single precision branch: MiniMax fit of atan(sqrt(x)) / sqrt(x))
AM_HD float atan2_(float y, float x) {
    const float ax = fabs_(x), ay = fabs_(y), mn = min_(ay, ax), mx = max_(ax, ay), sm = fdiv(mn, mx), z = sm * sm;
    const float c[7] = { 0.99999934166683966009f, -0.33326497518773606976f, 0.19881342388439013552f, -0.13486708938456973185f,
                         0.083863120428809689910f, -0.037006525670417265220f, 0.0078613793713198150252f };
    float t = estrin_n<7>(z, c) * sm;
    t = ay > ax ? AM_PI * .5f - t : t;
    t = x < 0.f ? AM_PI - t : t;
    const float r = y < 0.f ? -t : t;
    return mx != 0.f ? r : 0.f;
}

// math.h:76-175 detail::sincos (single precision branch)
AM_HD void sincos_(float x, float *s_out, float *c_out) {
    float xa = fabs_(x);
    int32_t j = (int32_t) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    uint32_t sign_sin = ((uint32_t) j << 29) ^ f2u(x);
    uint32_t sign_cos = (uint32_t) (~(j - 2)) << 29;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == AM_INF) z = u2f(0xffffffffu);
    float s = estrin3(z, -1.6666654611e-1f, 8.3321608736e-3f, -1.9515295891e-4f) * z;
    float c = estrin3(z, 4.166664568298827e-2f, -1.388731625493765e-3f, 2.443315711809948e-5f) * z;
    s = fmadd(s, y, y);
    c = fmadd(c, z, fmadd(z, -0.5f, 1.f));
    bool polymask = (j & 2) == 0;
    *s_out = mulsign(polymask ? s : c, u2f(sign_sin));
    *c_out = mulsign(polymask ? c : s, u2f(sign_cos));
}
// math.h:186-270 detail::tancot (single precision branch)
AM_HD float tan_(float x) {
    float xa = fabs_(x);
    int32_t j = (int32_t) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == AM_INF) z = u2f(0xffffffffu);
    const float c[6] = { 3.33331568548e-1f, 1.33387994085e-1f, 5.34112807005e-2f,
                         2.44301354525e-2f, 3.11992232697e-3f, 9.38540185543e-3f };
    float r = estrin6(z, c);
    r = fmadd(r, z * y, y);
    if (xa < 1e-4f) r = y;
    if ((j & 2) != 0) r = rcp(r);
    uint32_t sign = ((uint32_t) j << 30) ^ f2u(x);
    return mulsign(r, u2f(sign));
}
// math.h:615-632 ldexp
AM_HD float ldexp_(float x, float n) { return x * u2f((uint32_t) (((int32_t) n + 0x7f) << 23)); }
// math.h:690-760 exp
AM_HD float exp_(float x) {
    bool over = x > 88.3762588501f, under = x < -88.3762588501f;
    float n = floorf(fmadd(AM_INV_LOG_TWO, x, 0.5f));
    float y = x;
    y = fmadd(n, -0.693359375f, y);
    y = fmadd(n, 2.12194440e-4f, y);
    const float c[6] = { 5.0000001201e-1f, 1.6666665459e-1f, 4.1665795894e-2f,
                         8.3334519073e-3f, 1.3981999507e-3f, 1.9875691500e-4f };
    float z = estrin6(y, c);
    z = fmadd(z, y * y, y + 1.f);
    return over ? AM_INF : (under ? 0.f : ldexp_(z, n));
}
// math.h:763-830 exp2
AM_HD float exp2_(float x) {
    bool over = x > 127.f, under = x < -127.f;
    float n = floorf(x), y = x - n;
    if (y > .5f) { n += 1.f; y -= 1.f; }
    const float c[6] = { 6.931472028550421e-1f, 2.402264791363012e-1f, 5.550332471162809e-2f,
                         9.618437357674640e-3f, 1.339887440266574e-3f, 1.535336188319500e-4f };
    float z = estrin6(y, c);
    z = fmadd(y, z, 1.f);
    return over ? AM_INF : (under ? 0.f : ldexp_(z, n));
}
// math.h:639-683 frexp + :833-905 log
AM_HD float log_(float x) {
    bool valid = x >= 0.f;
    uint32_t xi = f2u(x), ebits = xi & 0x7f800000u;
    bool is_normal = (x != 0.f) && (ebits != 0x7f800000u);
    int32_t ei = is_normal ? ((int32_t) (ebits >> 23) - 0x7f) : 0;
    uint32_t mant = (xi & ~0x7f800000u) | 0x3f000000u;
    float xm = is_normal ? u2f(mant) : x, e = (float) ei;
    bool ge = xm >= AM_INV_SQRT_TWO;
    if (ge) e += 1.f;
    xm += (ge ? 0.f : xm) - 1.f;
    const float c[9] = { 3.3333331174e-1f, -2.4999993993e-1f, 2.0000714765e-1f, -1.6668057665e-1f,
                         1.4249322787e-1f, -1.2420140846e-1f, 1.1676998740e-1f, -1.1514610310e-1f,
                         7.0376836292e-2f };
    float y = estrin9(xm, c);
    float z = xm * xm;
    y *= xm * z;
    y = fmadd(e, -2.121944400546905827679e-4f, y);
    float r = xm + fmadd(-.5f, z, y);
    r = fmadd(e, 0.693359375f, r);
    if (x == AM_INF) r = AM_INF;
    if (x == 0.f) r = -AM_INF;
    return valid ? r : u2f(0xffffffffu);
}
// math.h:1497-1551 erf, array branch (double-precision coefficient set rounded to binary32)
AM_HD float erf_(float x) {
    float xa = fabs_(x), x2 = x * x;
    const float a0[12] = {
        (float) 0x1.20dd750429b6dp+0,  (float) -0x1.812746b0379bcp-2, (float) 0x1.ce2f21a040d12p-4,
        (float) -0x1.b82ce311fa924p-6, (float) 0x1.565bccf92b298p-8,  (float) -0x1.c02db03dd71b8p-11,
        (float) 0x1.f9a2baa8fee07p-14, (float) -0x1.f4ca4d6f3e31bp-17, (float) 0x1.b97fd3d992af4p-20,
        (float) -0x1.5c0726f04e805p-23, (float) 0x1.d71b0f1b15b0ap-27, (float) -0x1.abae491c540bp-31 };
    const float a1[8] = {
        (float) -0x1.a0be83b09c3d7p+0, (float) -0x1.8bb29648c7afep+1, (float) -0x1.639eb89a5975p+1,
        (float) -0x1.7b48b8cd14d9fp+0, (float) -0x1.fb25a03ddc781p-2, (float) -0x1.9cdb7dcacdfb3p-4,
        (float) -0x1.64f7fbe544f07p-7, (float) -0x1.9a3c3874b3919p-12 };
    const float a2[8] = {
        (float) 0x1p+0, (float) 0x1.55b5d06f1c2dep+0, (float) 0x1.b998ffae5528ep-1, (float) 0x1.46884d56cb49bp-2,
        (float) 0x1.18e8848a2cc38p-4, (float) 0x1.ee7f90e8d480cp-8, (float) 0x1.1c6a194029df4p-12,
        (float) -0x1.03d1306b29028p-31 };
    float c0 = estrin12(x2, a0);
    float c1 = fdiv(estrin8(xa, a1), estrin8(xa, a2));
    float xb = 1.f - exp2_(c1 * xa);
    return xa < 1.f ? x * c0 : copysign_(finite_(xb) ? xb : 1.f, x);
}
// math.h:1557-1578 erfinv
AM_HD float erfinv_(float x) {
    float w = -log_((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = fsqrt(w) - 3.f;
    const float p1[9] = { 1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                          -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f };
    const float p2[9] = { 2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                          -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f };
    return (w < 5.f ? estrin9(w1, p1) : estrin9(w2, p2)) * x;
}

// include/mitsuba/core/vector.h:116-136
AM_HD void coordinate_system(V3 n, V3 *s, V3 *t) {
    float sign = copysign_(1.f, n.z), a = -rcp(sign + n.z), b = n.x * n.y * a;
    *s = v3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    *t = v3(b, fmadd(n.y, n.y * a, sign), -n.y);
}
// include/mitsuba/core/warp.h:54-90
AM_HD V2 square_to_uniform_disk_concentric(V2 u) {
    float x = fmsub(2.f, u.x, 1.f), y = fmsub(2.f, u.y, 1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabs_(x) < fabs_(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = fdiv(0.25f * AM_PI * rp, r);
    if (q13) phi = 0.5f * AM_PI - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincos_(phi, &s, &c);
    return v2(r * c, r * s);
}
// warp.h:412-436
AM_HD V3 square_to_cosine_hemisphere(V2 u) {
    V2 p = square_to_uniform_disk_concentric(u);
    return v3(p.x, p.y, safe_sqrt(1.f - squared_norm(p)));
}
AM_HD float square_to_cosine_hemisphere_pdf(V3 v) { return AM_INV_PI * v.z; }
// warp.h:153-156
AM_HD V2 square_to_uniform_triangle(V2 u) {
    float t = safe_sqrt(1.f - u.x);
    return v2(1.f - t, t * u.y);
}

// warp.h:23, 250-255, 267-276
AM_HD float circ_(float x) { return safe_sqrt(fnmadd(x, x, 1.f)); }
AM_HD V3 square_to_uniform_sphere(V2 u) {
    const float z = fnmadd(2.f, u.y, 1.f), r = circ_(z);
    float s, c; sincos_(2.f * AM_PI * u.x, &s, &c);
    return v3(r * c, r * s, z);
}
#define AM_INV_FOUR_PI 0.07957747154594766788f
#define AM_INV_TWO_PI 0.15915494309189533577f
// warp.h:446-453, 478-494, 515-521: importance sampling of a linear / bilinear interpolant
AM_HD float interval_to_linear(float v0, float v1, float sample) {
    const float r = fdiv(v0 - safe_sqrt(lerp_(sqr(v0), sqr(v1), sample)), v0 - v1);
    return fabs_(v0 - v1) > 1e-4f * (v0 + v1) ? r : sample;
}
AM_HD V2 square_to_bilinear(float v00, float v10, float v01, float v11, V2 sample, float *pdf) {
    const float r0 = v00 + v10, r1 = v01 + v11;
    sample.y = interval_to_linear(r0, r1, sample.y);
    const float c0 = lerp_(v00, v01, sample.y), c1 = lerp_(v10, v11, sample.y);
    sample.x = interval_to_linear(c0, c1, sample.x);
    *pdf = lerp_(c0, c1, sample.x);
    return sample;
}
AM_HD float square_to_bilinear_pdf(float v00, float v10, float v01, float v11, V2 sample) {
    return lerp_(lerp_(v00, v10, sample.x), lerp_(v01, v11, sample.x), sample.y);
}

// include/mitsuba/render/fresnel.h:30-73
AM_HD void fresnel(float cos_i, float eta, float *r_out, float *cos_t, float *eta_it_o, float *eta_ti_o) {
    bool outside = cos_i >= 0.f;
    float rcp_eta = rcp(eta), eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float cos_t_sqr = fnmadd(fnmadd(cos_i, cos_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabs_(cos_i), ct = safe_sqrt(cos_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || (ci == 0.f);
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fdiv(fnmadd(eta_it, ct, ci), fmadd(eta_it, ct, ci));
    float a_p = fdiv(fnmadd(eta_it, ci, ct), fmadd(eta_it, ci, ct));
    float r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = r_sc;
    *r_out = r; *cos_t = mulsign_neg(ct, cos_i); *eta_it_o = eta_it; *eta_ti_o = eta_ti;
}
// fresnel.h:93-118
AM_HD float fresnel_conductor(float cos_i, float eta_r, float eta_i) {
    float c2 = cos_i * cos_i, s2 = 1.f - c2, s4 = s2 * s2;
    float temp_1 = eta_r * eta_r - eta_i * eta_i - s2;
    float a2pb2 = safe_sqrt(temp_1 * temp_1 + 4.f * eta_i * eta_i * eta_r * eta_r);
    float a = safe_sqrt(.5f * (a2pb2 + temp_1));
    float term_1 = a2pb2 + c2, term_2 = 2.f * cos_i * a;
    float r_s = fdiv(term_1 - term_2, term_1 + term_2);
    float term_3 = a2pb2 * c2 + s4, term_4 = term_2 * s2;
    float r_p = fdiv(r_s * (term_3 - term_4), term_3 + term_4);
    return 0.5f * (r_s + r_p);
}
// fresnel.h:276-300
AM_HD V3 reflect(V3 wi) { return v3(-wi.x, -wi.y, wi.z); }
AM_HD V3 reflect(V3 wi, V3 m) {
    float k = 2.f * dot(wi, m);
    return v3(fmsub(m.x, k, wi.x), fmsub(m.y, k, wi.y), fmsub(m.z, k, wi.z));
}
AM_HD V3 refract(V3 wi, float cos_t, float eta_ti) { return v3(-eta_ti * wi.x, -eta_ti * wi.y, cos_t); }
AM_HD V3 refract(V3 wi, V3 m, float cos_t, float eta_ti) {
    float k = fmadd(dot(wi, m), eta_ti, cos_t);
    return v3(fmsub(m.x, k, wi.x * eta_ti), fmsub(m.y, k, wi.y * eta_ti), fmsub(m.z, k, wi.z * eta_ti));
}

// ---------------------------------------------------------------------------------------------
// Sampler: include/mitsuba/core/random.h:76-91 (TEA), ext/drjit/include/drjit/random.h:59-150 (PCG32),
// src/render/sampler.cpp:125-144 (per-lane seeding of the independent sampler in wavefront mode).
AM_HD void tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *o0, uint32_t *o1) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    *o0 = v0; *o1 = v1;
}
struct Pcg32 {
    uint64_t state, inc;
    AM_HD void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u;
        next_u32(); state += initstate; next_u32();
    }
    AM_HD uint32_t next_u32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dull + inc;
        uint32_t xs = (uint32_t) (((old >> 18) ^ old) >> 27), rot = (uint32_t) (old >> 59);
        return (xs >> rot) | (xs << ((0u - rot) & 31u));
    }
    AM_HD float next_f32() { return u2f((next_u32() >> 9) | 0x3f800000u) - 1.f; }
};
AM_HD Pcg32 sampler_seed(uint32_t seed_value, uint32_t lane) {
    uint32_t v0, v1; tea32(seed_value, lane, 4, &v0, &v1);
    Pcg32 r; r.seed(v0, v1); return r;
}

} // namespace amvpt
