// htransform.h -- host-side camera set-up arithmetic, bit-compatible with the reference's scalar code:
// dr::Matrix<float,4> product / inverse (ext/drjit/include/drjit/matrix.h:105-120, 440-503), Transform4f builders
// (include/mitsuba/core/transform.h:195-300), perspective_projection (include/mitsuba/render/sensor.h:320-356),
// parse_fov (src/render/sensor.cpp:177-231), and the double-precision scene-description transforms of the XML
// loader (src/core/xml.cpp:838-926).  Host translation units are compiled with -mfma -ffp-contract=off.
#pragma once
#include "dmath.cuh"

namespace amvpt {

struct M4 { float m[4][4]; };     // row-major
struct M4d { double m[4][4]; };
struct Xf { M4 mat, inv_t; };     // Transform = (matrix, inverse transpose), transform.h:28-77

inline M4 m4_identity() { M4 r{}; for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f; return r; }
inline M4 m4_transpose(const M4 &a) { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i]; return r; }
// matrix.h:105-120: row_i = a(i,0) * b.row(0), then fmadd(a(i,j), b.row(j), row_i)
inline M4 m4_mul(const M4 &a, const M4 &b) {
    M4 r;
    for (int i = 0; i < 4; ++i) for (int c = 0; c < 4; ++c) {
        float acc = a.m[i][0] * b.m[0][c];
        for (int j = 1; j < 4; ++j) acc = fmadd(a.m[i][j], b.m[j][c], acc);
        r.m[i][c] = acc;
    }
    return r;
}
// matrix.h:440-499: 4x4 inverse-transpose by cofactor expansion on shuffled rows.  Written here as an explicit
// lane program: every step is one of {mul, fmsub, fmadd, fnmadd} on 4-vectors followed by a fixed shuffle.
struct Q4 { float v[4]; };
inline Q4 q_sh(const Q4 &a, int i0, int i1, int i2, int i3) { return { { a.v[i0], a.v[i1], a.v[i2], a.v[i3] } }; }
inline Q4 q_swap_pairs(const Q4 &a) { return q_sh(a, 1, 0, 3, 2); }
inline Q4 q_swap_halves(const Q4 &a) { return q_sh(a, 2, 3, 0, 1); }
inline Q4 q_mul(const Q4 &a, const Q4 &b) { Q4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * b.v[i]; return r; }
inline Q4 q_fma(const Q4 &a, const Q4 &b, const Q4 &c) { Q4 r; for (int i = 0; i < 4; ++i) r.v[i] = fmadd(a.v[i], b.v[i], c.v[i]); return r; }
inline Q4 q_fms(const Q4 &a, const Q4 &b, const Q4 &c) { Q4 r; for (int i = 0; i < 4; ++i) r.v[i] = fmsub(a.v[i], b.v[i], c.v[i]); return r; }
inline Q4 q_fnma(const Q4 &a, const Q4 &b, const Q4 &c) { Q4 r; for (int i = 0; i < 4; ++i) r.v[i] = fnmadd(a.v[i], b.v[i], c.v[i]); return r; }
inline M4 m4_inverse_transpose(const M4 &m) {
    Q4 r0, r1, r2, r3;
    for (int i = 0; i < 4; ++i) { r0.v[i] = m.m[0][i]; r1.v[i] = m.m[1][i]; r2.v[i] = m.m[2][i]; r3.v[i] = m.m[3][i]; }
    r1 = q_swap_halves(r1); r3 = q_swap_halves(r3);
    Q4 t, c0, c1, c2, c3;
    t = q_swap_pairs(q_mul(r2, r3));
    c0 = q_mul(r1, t); c1 = q_mul(r0, t);
    t = q_swap_halves(t);
    c0 = q_fms(r1, t, c0); c1 = q_swap_halves(q_fms(r0, t, c1));
    t = q_swap_pairs(q_mul(r1, r2));
    c0 = q_fma(r3, t, c0); c3 = q_mul(r0, t);
    t = q_swap_halves(t);
    c0 = q_fnma(r3, t, c0); c3 = q_swap_halves(q_fms(r0, t, c3));
    t = q_swap_pairs(q_mul(q_swap_halves(r1), r3));
    r2 = q_swap_halves(r2);
    c0 = q_fma(r2, t, c0); c2 = q_mul(r0, t);
    t = q_swap_halves(t);
    c0 = q_fnma(r2, t, c0); c2 = q_swap_halves(q_fms(r0, t, c2));
    t = q_swap_pairs(q_mul(r0, r1));
    c2 = q_fma(r3, t, c2); c3 = q_fms(r2, t, c3);
    t = q_swap_halves(t);
    c2 = q_fms(r3, t, c2); c3 = q_fnma(r2, t, c3);
    t = q_swap_pairs(q_mul(r0, r3));
    c1 = q_fnma(r2, t, c1); c2 = q_fma(r1, t, c2);
    t = q_swap_halves(t);
    c1 = q_fma(r2, t, c1); c2 = q_fnma(r1, t, c2);
    t = q_swap_pairs(q_mul(r0, r2));
    c1 = q_fma(r3, t, c1); c3 = q_fnma(r1, t, c3);
    t = q_swap_halves(t);
    c1 = q_fnma(r3, t, c1); c3 = q_fma(r1, t, c3);
    float det = r0.v[0] * c0.v[0];
    for (int i = 1; i < 4; ++i) det = fmadd(r0.v[i], c0.v[i], det);
    float inv_det = rcp(det);
    M4 r;
    for (int i = 0; i < 4; ++i) { r.m[0][i] = c0.v[i] * inv_det; r.m[1][i] = c1.v[i] * inv_det; r.m[2][i] = c2.v[i] * inv_det; r.m[3][i] = c3.v[i] * inv_det; }
    return r;
}
inline M4 m4_inverse(const M4 &m) { return m4_transpose(m4_inverse_transpose(m)); }

inline Xf xf_mul(const Xf &a, const Xf &b) { return { m4_mul(a.mat, b.mat), m4_mul(a.inv_t, b.inv_t) }; }
inline Xf xf_translate(float x, float y, float z) {      // transform.h:195-200
    M4 a = m4_identity(), b = m4_identity();
    a.m[0][3] = x; a.m[1][3] = y; a.m[2][3] = z; b.m[0][3] = -x; b.m[1][3] = -y; b.m[2][3] = -z;
    return { a, m4_transpose(b) };
}
inline Xf xf_scale(float x, float y, float z) {          // transform.h:203-206
    M4 a = m4_identity(), b = m4_identity();
    a.m[0][0] = x; a.m[1][1] = y; a.m[2][2] = z; b.m[0][0] = rcp(x); b.m[1][1] = rcp(y); b.m[2][2] = rcp(z);
    return { a, b };
}
inline Xf xf_perspective(float fov, float near_, float far_) {   // transform.h:235-252
    float recip = 1.f / (far_ - near_);
    float tn = tan_((fov * .5f) * (AM_PI / 180.f)), cot = 1.f / tn;
    M4 t{}; t.m[0][0] = cot; t.m[1][1] = cot; t.m[2][2] = far_ * recip; t.m[2][3] = -near_ * far_ * recip; t.m[3][2] = 1.f;
    M4 it{}; it.m[0][0] = tn; it.m[1][1] = tn; it.m[3][3] = rcp(near_); it.m[2][3] = 1.f; it.m[3][2] = (near_ - far_) / (far_ * near_);
    return { t, m4_transpose(it) };
}
// sensor.h:320-356 (crop = full film)
inline Xf perspective_projection(int fw, int fh, float fov_x, float near_, float far_) {
    float fsx = (float) fw, fsy = (float) fh, aspect = fsx / fsy;
    float rsx = (float) fw / fsx, rsy = (float) fh / fsy, rox = 0.f / fsx, roy = 0.f / fsy;
    Xf r = xf_scale(1.f / rsx, 1.f / rsy, 1.f);
    r = xf_mul(r, xf_translate(-rox, -roy, 0.f));
    r = xf_mul(r, xf_scale(-0.5f, -0.5f * aspect, 1.f));
    r = xf_mul(r, xf_translate(-1.f, -1.f / aspect, 0.f));
    r = xf_mul(r, xf_perspective(fov_x, near_, far_));
    return r;
}
// PerspectiveCamera::update_camera_transforms (src/sensors/perspective.cpp:173-181): sample_to_camera.matrix
inline M4 sample_to_camera(int fw, int fh, float x_fov, float near_, float far_, float lens_shift) {
    Xf c2s = perspective_projection(fw, fh, x_fov, near_, far_);
    c2s.mat.m[0][2] += lens_shift;
    return m4_transpose(m4_inverse_transpose(c2s.mat));
}
// src/render/sensor.cpp:177-231
inline double parse_fov_x(double fov, int axis /*0 x,1 y,2 diagonal,3 smaller,4 larger*/, double aspect) {
    if (axis == 3) axis = aspect > 1 ? 1 : 0; else if (axis == 4) axis = aspect > 1 ? 0 : 1;
    const double d2r = 3.14159265358979323846 / 180.0, r2d = 180.0 / 3.14159265358979323846;
    if (axis == 0) return fov;
    if (axis == 1) return r2d * (2.0 * std::atan(std::tan(0.5 * (fov * d2r)) * aspect));
    double diagonal = 2.0 * std::tan(0.5 * (fov * d2r));
    double width = diagonal / std::sqrt(1.0 + 1.0 / (aspect * aspect));
    return r2d * (2.0 * std::atan(width * 0.5));
}

// ---- double-precision transforms of the scene description (Properties::Float = double)
inline M4d m4d_identity() { M4d r{}; for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0; return r; }
inline M4d m4d_mul(const M4d &a, const M4d &b) {
    M4d r;
    for (int i = 0; i < 4; ++i) for (int c = 0; c < 4; ++c) {
        double acc = a.m[i][0] * b.m[0][c];
        for (int j = 1; j < 4; ++j) acc = std::fma(a.m[i][j], b.m[j][c], acc);
        r.m[i][c] = acc;
    }
    return r;
}
inline M4 m4_from_d(const M4d &d) { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = (float) d.m[i][j]; return r; }
M4d m4d_inverse(const M4d &m);                                   // general inverse (Gauss-Jordan), scene.cpp
M4d m4d_translate(double x, double y, double z);
M4d m4d_scale(double x, double y, double z);
M4d m4d_rotate(double ax, double ay, double az, double angle_deg);    // transform.h:209-232
M4d m4d_look_at(const double *origin, const double *target, const double *up);   // transform.h:269-300

} // namespace amvpt
