// oracle/o_transform.h -- TEST INFRASTRUCTURE ONLY.  Restatement of the host-side camera set-up
// arithmetic of the reference: dr::Matrix<float,4> product / inverse (ext/drjit/include/drjit/
// matrix.h:105-120, 440-503), Transform4f builders (include/mitsuba/core/transform.h:195-300),
// dr::tan (ext/drjit/include/drjit/math.h:186-270), perspective_projection (include/mitsuba/
// render/sensor.h:320-356), PerspectiveCamera::update_camera_transforms (src/sensors/
// perspective.cpp:173-181) and the GridSensor placement loop (src/sensors/grid.cpp:182-226).
// Pinned against oracle/_ref by tests/test_oracle_vs_ref.py.
#pragma once
#include "o_math.h"

namespace orc {

struct M4 { float m[4][4]; };   // row-major: m[r][c]
struct A4 { float v[4]; };

static inline M4 m4_identity() { M4 r{}; for (int i = 0; i < 4; ++i) r.m[i][i] = 1.f; return r; }
static inline M4 m4_transpose(const M4 &a) { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a.m[j][i]; return r; }
// matrix.h:105-120: row_i = a(i,0)*b.row(0); row_i = fmadd(a(i,j), b.row(j), row_i)
static inline M4 m4_mul(const M4 &a, const M4 &b) {
    M4 r;
    for (int i = 0; i < 4; ++i)
        for (int c = 0; c < 4; ++c) {
            float row = a.m[i][0] * b.m[0][c];
            for (int j = 1; j < 4; ++j) row = fmadd(a.m[i][j], b.m[j][c], row);
            r.m[i][c] = row;
        }
    return r;
}
static inline A4 sh(const A4 &a, int i0, int i1, int i2, int i3) { return { { a.v[i0], a.v[i1], a.v[i2], a.v[i3] } }; }
static inline A4 a_mul(const A4 &a, const A4 &b) { A4 r; for (int i = 0; i < 4; ++i) r.v[i] = a.v[i] * b.v[i]; return r; }
static inline A4 a_fmadd(const A4 &a, const A4 &b, const A4 &c) { A4 r; for (int i = 0; i < 4; ++i) r.v[i] = fmadd(a.v[i], b.v[i], c.v[i]); return r; }
static inline A4 a_fmsub(const A4 &a, const A4 &b, const A4 &c) { A4 r; for (int i = 0; i < 4; ++i) r.v[i] = fmsub(a.v[i], b.v[i], c.v[i]); return r; }
static inline A4 a_fnmadd(const A4 &a, const A4 &b, const A4 &c) { A4 r; for (int i = 0; i < 4; ++i) r.v[i] = fnmadd(a.v[i], b.v[i], c.v[i]); return r; }
static inline A4 row(const M4 &m, int i) { return { { m.m[i][0], m.m[i][1], m.m[i][2], m.m[i][3] } }; }

// matrix.h:440-499 (cofactor expansion on shuffled rows)
static inline M4 m4_inverse_transpose(const M4 &m) {
    A4 row0 = row(m, 0), row1 = row(m, 1), row2 = row(m, 2), row3 = row(m, 3);
    row1 = sh(row1, 2, 3, 0, 1);
    row3 = sh(row3, 2, 3, 0, 1);
    A4 temp, col0, col1, col2, col3;

    temp = sh(a_mul(row2, row3), 1, 0, 3, 2);
    col0 = a_mul(row1, temp);
    col1 = a_mul(row0, temp);
    temp = sh(temp, 2, 3, 0, 1);
    col0 = a_fmsub(row1, temp, col0);
    col1 = sh(a_fmsub(row0, temp, col1), 2, 3, 0, 1);

    temp = sh(a_mul(row1, row2), 1, 0, 3, 2);
    col0 = a_fmadd(row3, temp, col0);
    col3 = a_mul(row0, temp);
    temp = sh(temp, 2, 3, 0, 1);
    col0 = a_fnmadd(row3, temp, col0);
    col3 = sh(a_fmsub(row0, temp, col3), 2, 3, 0, 1);

    temp = sh(a_mul(sh(row1, 2, 3, 0, 1), row3), 1, 0, 3, 2);
    row2 = sh(row2, 2, 3, 0, 1);
    col0 = a_fmadd(row2, temp, col0);
    col2 = a_mul(row0, temp);
    temp = sh(temp, 2, 3, 0, 1);
    col0 = a_fnmadd(row2, temp, col0);
    col2 = sh(a_fmsub(row0, temp, col2), 2, 3, 0, 1);

    temp = sh(a_mul(row0, row1), 1, 0, 3, 2);
    col2 = a_fmadd(row3, temp, col2);
    col3 = a_fmsub(row2, temp, col3);
    temp = sh(temp, 2, 3, 0, 1);
    col2 = a_fmsub(row3, temp, col2);
    col3 = a_fnmadd(row2, temp, col3);

    temp = sh(a_mul(row0, row3), 1, 0, 3, 2);
    col1 = a_fnmadd(row2, temp, col1);
    col2 = a_fmadd(row1, temp, col2);
    temp = sh(temp, 2, 3, 0, 1);
    col1 = a_fmadd(row2, temp, col1);
    col2 = a_fnmadd(row1, temp, col2);

    temp = sh(a_mul(row0, row2), 1, 0, 3, 2);
    col1 = a_fmadd(row3, temp, col1);
    col3 = a_fnmadd(row1, temp, col3);
    temp = sh(temp, 2, 3, 0, 1);
    col1 = a_fnmadd(row3, temp, col1);
    col3 = a_fmadd(row1, temp, col3);

    // dot(row0, col0): fmadd chain (array_base.h:597-625)
    float det = row0.v[0] * col0.v[0];
    for (int i = 1; i < 4; ++i) det = fmadd(row0.v[i], col0.v[i], det);
    float inv_det = rcp(det);
    M4 r;
    for (int i = 0; i < 4; ++i) {
        r.m[0][i] = col0.v[i] * inv_det; r.m[1][i] = col1.v[i] * inv_det;
        r.m[2][i] = col2.v[i] * inv_det; r.m[3][i] = col3.v[i] * inv_det;
    }
    return r;
}
static inline M4 m4_inverse(const M4 &m) { return m4_transpose(m4_inverse_transpose(m)); }

// Transform = (matrix, inverse_transpose), transform.h:28-77
struct Xf { M4 mat, inv_t; };
static inline Xf xf_mul(const Xf &a, const Xf &b) { return { m4_mul(a.mat, b.mat), m4_mul(a.inv_t, b.inv_t) }; }
static inline Xf xf_inverse(const Xf &a) { return { m4_transpose(a.inv_t), m4_transpose(a.mat) }; }
// transform.h:195-206 (dr::translate / dr::scale: identity with the last column / diagonal set)
static inline Xf xf_translate(float x, float y, float z) {
    M4 a = m4_identity(), b = m4_identity();
    a.m[0][3] = x; a.m[1][3] = y; a.m[2][3] = z;
    b.m[0][3] = -x; b.m[1][3] = -y; b.m[2][3] = -z;
    return { a, m4_transpose(b) };
}
static inline Xf xf_scale(float x, float y, float z) {
    M4 a = m4_identity(), b = m4_identity();
    a.m[0][0] = x; a.m[1][1] = y; a.m[2][2] = z;
    b.m[0][0] = rcp(x); b.m[1][1] = rcp(y); b.m[2][2] = rcp(z);
    return { a, b };
}
// math.h:186-270 detail::tancot<true> (Single branch)
static inline float tan_(float x) {
    float xa = fabs_(x);
    int32_t j = (int32_t) (xa * 1.2732395447351626862f);
    j = (j + 1) & ~1;
    float y = (float) j;
    y = xa - y * 0.78515625f - y * 2.4187564849853515625e-4f - y * 3.77489497744594108e-8f;
    float z = y * y;
    if (xa == Inf) z = u2f(0xffffffffu);
    static const float c[6] = { 3.33331568548e-1f, 1.33387994085e-1f, 5.34112807005e-2f,
                                2.44301354525e-2f, 3.11992232697e-3f, 9.38540185543e-3f };
    float r = estrin(z, c, 6);
    r = fmadd(r, z * y, y);
    if (xa < 1e-4f) r = y;
    if ((j & 2) != 0) r = rcp(r);
    uint32_t sign = ((uint32_t) j << 30) ^ f2u(x);
    return mulsign(r, u2f(sign));
}
// transform.h:235-252
static inline Xf xf_perspective(float fov, float near_, float far_) {
    float recip = 1.f / (far_ - near_);
    float tn = tan_((fov * .5f) * (Pi / 180.f)), cot = 1.f / tn;
    M4 t{}; t.m[0][0] = cot; t.m[1][1] = cot; t.m[2][2] = far_ * recip; t.m[3][3] = 0.f;
    t.m[2][3] = -near_ * far_ * recip; t.m[3][2] = 1.f;
    M4 it{}; it.m[0][0] = tn; it.m[1][1] = tn; it.m[2][2] = 0.f; it.m[3][3] = rcp(near_);
    it.m[2][3] = 1.f; it.m[3][2] = (near_ - far_) / (far_ * near_);
    return { t, m4_transpose(it) };
}
// sensor.h:320-356
static inline Xf perspective_projection(int fw, int fh, int cw, int ch, int cox, int coy, float fov_x, float near_, float far_) {
    float fsx = (float) fw, fsy = (float) fh;
    float rsx = (float) cw / fsx, rsy = (float) ch / fsy, rox = (float) cox / fsx, roy = (float) coy / fsy;
    float aspect = fsx / fsy;
    Xf r = xf_scale(1.f / rsx, 1.f / rsy, 1.f);
    r = xf_mul(r, xf_translate(-rox, -roy, 0.f));
    r = xf_mul(r, xf_scale(-0.5f, -0.5f * aspect, 1.f));
    r = xf_mul(r, xf_translate(-1.f, -1.f / aspect, 0.f));
    r = xf_mul(r, xf_perspective(fov_x, near_, far_));
    return r;
}
// perspective.cpp:173-181 -> sample_to_camera.matrix (only the forward matrix is used by sample_ray)
static inline M4 sample_to_camera(int fw, int fh, float x_fov, float near_, float far_, float lens_shift) {
    Xf c2s = perspective_projection(fw, fh, fw, fh, 0, 0, x_fov, near_, far_);
    c2s.mat.m[0][2] += lens_shift;
    c2s.inv_t = m4_inverse_transpose(c2s.mat);
    return m4_transpose(c2s.inv_t);   // = xf_inverse(c2s).mat
}

// ---- double-precision scene-description transforms (src/core/xml.cpp:838-926; Properties::Float = double)
struct M4d { double m[4][4]; };
static inline M4d m4d_identity() { M4d r{}; for (int i = 0; i < 4; ++i) r.m[i][i] = 1.0; return r; }
static inline M4d m4d_mul(const M4d &a, const M4d &b) {
    M4d r;
    for (int i = 0; i < 4; ++i) for (int c = 0; c < 4; ++c) {
        double row = a.m[i][0] * b.m[0][c];
        for (int j = 1; j < 4; ++j) row = std::fma(a.m[i][j], b.m[j][c], row);
        r.m[i][c] = row;
    }
    return r;
}

// sensor.cpp:177-231
static inline double parse_fov(double fov, int axis /*0 x,1 y,2 diagonal,3 smaller,4 larger*/, double aspect) {
    if (axis == 3) axis = aspect > 1 ? 1 : 0;
    else if (axis == 4) axis = aspect > 1 ? 0 : 1;
    const double d2r = 3.14159265358979323846 / 180.0, r2d = 180.0 / 3.14159265358979323846;
    if (axis == 0) return fov;
    if (axis == 1) return r2d * (2.0 * std::atan(std::tan(0.5 * (fov * d2r)) * aspect));
    double diagonal = 2.0 * std::tan(0.5 * (fov * d2r));
    double width = diagonal / std::sqrt(1.0 + 1.0 / (aspect * aspect));
    return r2d * (2.0 * std::atan(width * 0.5));
}

// grid.cpp:182-226: to_world of sub-camera i (line placement, or cone placement when use_cone)
struct GridParams {
    uint32_t grid_x = 1, grid_y = 1; int reverse_x = 0, reverse_y = 1;
    int use_cone = 0; float cone_deg = 0.f;
    float cam_dir[3] = { 1, 0, 0 }; float cam_dist = 0.1f; int cam_center = 0;
    float cam_off[3] = { 0, 0, 0 };   // as given by the user (y,z negated inside, grid.cpp:130-131)
    float foc_dist = 1.f;
};
static inline void grid_sub_camera(const GridParams &g, const M4 &to_world_inv_t, double fov_x, uint32_t i,
                                   M4 *to_world_out, float *lens_shift_out) {
    uint32_t n = g.grid_x * g.grid_y;
    float off_y = -g.cam_off[1], off_z = -g.cam_off[2];
    float dt = (float) i / (float) (n - 1);
    M4 itrafo = to_world_inv_t;
    *lens_shift_out = 0.f;
    if (g.use_cone) {
        float tan_off = tan_((dt - 0.5f) * (g.cone_deg * (Pi / 180.f)));
        float offset = g.foc_dist * tan_off;
        double shift = 0.5 * (double) tan_off / std::tan((fov_x * (3.14159265358979323846 / 180.0)) * 0.5);
        itrafo.m[3][0] += offset + g.cam_off[0];
        itrafo.m[3][1] += off_y;
        itrafo.m[3][2] += off_z;
        *lens_shift_out = (float) shift;
    } else {
        float s = g.cam_dist * (dt - 0.5f * (float) g.cam_center);
        itrafo.m[3][0] += g.cam_off[0] + g.cam_dir[0] * s;
        itrafo.m[3][1] += off_y + g.cam_dir[1] * s;
        itrafo.m[3][2] += off_z + g.cam_dir[2] * s;
    }
    *to_world_out = m4_inverse(m4_transpose(itrafo));
}

} // namespace orc
