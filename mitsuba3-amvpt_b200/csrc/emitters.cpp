// emitters.cpp -- host side of the emitters beyond area lights (src/emitters/{point,constant,envmap}.cpp) and of the colour handling
// the scene loader needs for them: wavelength spectra -> linear sRGB (src/core/spectrum.cpp:85-130 + src/core/xml.cpp:1170-1268), the
// environment map's radiance array and its Hierarchical2D sampling hierarchy (envmap.cpp:111-214, include/mitsuba/core/distr_2d.h:376-482),
// the scene bounding sphere the environment emitters receive in set_scene().
#include "host.h"
#include "cie_tables.h"
#include <algorithm>
#include <cmath>

namespace amvpt {

// ---------------------------------------------------------------- spectra (host, double like the reference's xml.cpp: Properties::Float)
namespace {
const double kCieMin = 360.0, kCieMax = 830.0; const int kCieSamples = 95;
const double kCieYNormalization = 1.0 / 106.7502593994140625;       // include/mitsuba/core/spectrum.h:132
const double kCieD65Normalization = 1.0 / 98.99741751876255;        // :157
// cie1931_xyz / cie_d65 (include/mitsuba/core/spectrum.h:236-262, 302-328) for a scalar double wavelength: linear interpolation of the table
void cie_lookup(double w, double *xyz, double *d65) {
    const double t = (w - kCieMin) * ((kCieSamples - 1) / (kCieMax - kCieMin));
    const bool active = w >= kCieMin && w <= kCieMax;
    int i0 = (int) (uint32_t) t; i0 = std::min(std::max(i0, 0), kCieSamples - 2); const int i1 = i0 + 1;
    const double w1 = t - (double) i0, w0 = 1.0 - w1;
    for (int c = 0; c < 3; ++c) xyz[c] = active ? std::fma(w0, (double) kCie1931[c][i0], w1 * (double) kCie1931[c][i1]) : 0.0;
    *d65 = active ? std::fma(w0, (double) kCieD65[i0], w1 * (double) kCieD65[i1]) * kCieD65Normalization : 0.0;
}
void xyz_to_srgb(const double *xyz, double *rgb) {                  // spectrum.h:405-412 (float literals widened)
    const float M[3][3] = { { 3.240479f, -1.537150f, -0.498535f }, { -0.969256f, 1.875991f, 0.041556f }, { 0.055648f, -0.204043f, 1.057311f } };
    // Matrix * vector in Dr.Jit accumulates column by column: M[:,0] * x, then fmadd with the next columns (drjit/matrix.h)
    for (int r = 0; r < 3; ++r) rgb[r] = std::fma((double) M[r][2], xyz[2], std::fma((double) M[r][1], xyz[1], (double) M[r][0] * xyz[0]));
}
} // namespace

void uniform_emitter_spectrum_to_srgb(double value, double *rgb) {   // xml.cpp:1180-1190
    const double one[3] = { 1.0, 1.0, 1.0 }; double c[3]; xyz_to_srgb(one, c);
    for (int k = 0; k < 3; ++k) rgb[k] = value * c[k];
}
void spectrum_to_srgb(std::vector<double> wavelengths, std::vector<double> values, bool within_emitter, bool unbounded_name, double *rgb) {
    if (wavelengths.size() != values.size() || wavelengths.size() < 2) AM_THROW("spectrum: need at least two wavelength:value pairs");
    for (size_t n = 0; n < wavelengths.size(); ++n) {                 // xml.cpp:1206-1220
        values[n] *= kCieYNormalization;
        if (n > 0 && wavelengths[n] - wavelengths[n - 1] < 0) AM_THROW("Wavelengths must be specified in increasing order!");
    }
    const bool bounded = !(within_emitter || unbounded_name), d65 = !(within_emitter && !unbounded_name);      // xml.cpp:1249-1252
    spectrum_list_to_srgb(wavelengths, values, bounded, d65, rgb);
}
// spectrum_list_to_srgb (src/core/spectrum.cpp:85-130); `values` already carry the unit conversion
void spectrum_list_to_srgb(const std::vector<double> &wavelengths, const std::vector<double> &values, bool bounded, bool d65, double *rgb) {
    if (wavelengths.size() != values.size() || wavelengths.size() < 2) AM_THROW("spectrum: need at least two wavelength:value pairs");
    double xyz[3] = { 0, 0, 0 }; const int steps = 1000;
    for (int i = 0; i < steps; ++i) {                                 // spectrum.cpp:92-114
        const double w = kCieMin + (i / (double) (steps - 1)) * (kCieMax - kCieMin);
        if (w < wavelengths.front() || w > wavelengths.back()) continue;
        // math::find_interval: last index with wavelengths[idx] <= w, clamped to [0, size - 2]
        size_t index = 0; while (index + 2 < wavelengths.size() && wavelengths[index + 1] <= w) ++index;
        const double w0 = wavelengths[index], w1 = wavelengths[index + 1], v0 = values[index], v1 = values[index + 1];
        const double v = ((w - w1) * v0 + (w0 - w) * v1) / (w0 - w1);
        double c[3], dd; cie_lookup(w, c, &dd);
        for (int k = 0; k < 3; ++k) xyz[k] += c[k] * v * (d65 ? dd : 1.0);
    }
    for (int k = 0; k < 3; ++k) xyz[k] *= (kCieMax - kCieMin) / (double) steps;
    xyz_to_srgb(xyz, rgb);
    for (int k = 0; k < 3; ++k) { if (rgb[k] < 0.0) rgb[k] = 0.0; if (bounded && rgb[k] > 1.0) rgb[k] = 1.0; }      // out-of-gamut clamp, spectrum.cpp:120-126
}

// ---------------------------------------------------------------- environment map
// Hierarchical2D<Float, 0>::Level::index (distr_2d.h:757-761): 2 x 2 blocks are stored contiguously
static inline uint32_t level_index(uint32_t x, uint32_t y, uint32_t width) { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + (y & ~1u) * width; }

void EnvMap::build(const float *rgb, uint32_t w, uint32_t h, float scale_, const M4d &tw) {
    if (w < 2 || h < 3) AM_THROW("the environment map resolution must be at least 2x3 pixels");      // envmap.cpp:135-137
    present = true; src_w = w; src_h = h; src_rgb.assign(rgb, rgb + (size_t) w * h * 3);
    scale = scale_; to_world = m4_from_d(tw); from_world = m4_inverse(to_world);
    res_x = w + 1; res_y = h; data.assign((size_t) res_x * res_y * 3, 0.f);
    std::vector<float> lum((size_t) res_x * res_y);
    const float theta_scale = 1.f / (float) (h - 1) * AM_PI;                                       // envmap.cpp:157
    for (uint32_t y = 0; y < h; ++y) {
        float sin_theta, cos_theta; sincos_((float) y * theta_scale, &sin_theta, &cos_theta);      // dr::sin of a scalar float = Dr.Jit's polynomial
        for (uint32_t x = 0; x < w; ++x) {
            const float *p = rgb + ((size_t) y * w + x) * 3; float *o = data.data() + ((size_t) y * res_x + x) * 3;
            float l = p[0] * 0.212671f + p[1] * 0.715160f + p[2] * 0.072169f;                       // mitsuba::luminance, spectrum.h:431-434
            l = max_(l - 0.f, 0.f);                                                                 // mis_compensation = false: offset 0
            lum[(size_t) y * res_x + x] = l * sin_theta; o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
        }
        lum[(size_t) y * res_x + w] = lum[(size_t) y * res_x];                                      // last column mirrors the first (:199-205)
        for (int c = 0; c < 3; ++c) data[((size_t) y * res_x + w) * 3 + c] = data[((size_t) y * res_x) * 3 + c];
    }
    // Hierarchical2D(data = lum, size = (res_x, res_y)), normalize = true, enable_sampling = true
    const uint32_t npx = res_x - 1, npy = res_y - 1, nmax = std::max(npx, npy);
    uint32_t max_level = 31u - (uint32_t) __builtin_clz(nmax); if (nmax & (nmax - 1u)) max_level += 1;     // math::log2i_ceil
    patch_size[0] = 1.f / (float) npx; patch_size[1] = 1.f / (float) npy; inv_patch_size[0] = (float) npx; inv_patch_size[1] = (float) npy;
    max_patch[0] = npx - 1; max_patch[1] = npy - 1;
    struct L { uint32_t w, h; size_t off; }; std::vector<L> lv; size_t total = 0;
    lv.push_back({ res_x, res_y, 0 }); total += (size_t) res_x * res_y;
    { uint32_t lx = npx, ly = npy;
      for (int level = (int) max_level; level >= 0; --level) { lx += lx & 1u; ly += ly & 1u; lv.push_back({ lx, ly, total }); total += (size_t) lx * ly; lx >>= 1; ly >>= 1; } }
    if (lv.size() > 12) AM_THROW("environment map too large for the device-side level table (%zu levels)", lv.size());
    warp.assign(total, 0.f); level_off.clear(); level_width.clear();
    for (auto &l : lv) { level_off.push_back((uint32_t) l.off); level_width.push_back(l.w); }
    float *l0 = warp.data(), *l1 = warp.data() + lv[1].off; double sum = 0.0;
    for (uint32_t y = 0; y < npy; ++y) for (uint32_t x = 0; x < npx; ++x) {
        const float *in = lum.data() + (size_t) y * res_x + x;
        const float avg = .25f * (in[0] + in[1] + in[res_x] + in[res_x + 1]);
        sum += (double) avg; l1[level_index(x, y, lv[1].w)] = avg;
    }
    const float norm = (float) ((double) (npx * npy) / sum);
    for (size_t i = 0; i < (size_t) res_x * res_y; ++i) l0[i] = lum[i] * norm;
    for (size_t i = 0; i < (size_t) lv[1].w * lv[1].h; ++i) l1[i] *= norm;
    uint32_t lx = npx, ly = npy;
    for (uint32_t level = 2; level <= max_level + 1; ++level) {
        const L &a = lv[level - 1], &b = lv[level]; lx = (lx + 1u) >> 1; ly = (ly + 1u) >> 1;
        const float *pa = warp.data() + a.off; float *pb = warp.data() + b.off;
        for (uint32_t y = 0; y < ly; ++y) for (uint32_t x = 0; x < lx; ++x) {
            const float *d0 = pa + level_index(2 * x, 2 * y, a.w);
            pb[level_index(x, y, b.w)] = d0[0] + d0[1] + d0[2] + d0[3];
        }
    }
}

// ---------------------------------------------------------------- Scene: emitters that are not attached to a shape
int Scene::add_emitter_point(const float *position, const float *intensity) {
    DEmitter e{}; e.type = EMITTER_POINT; e.mesh = -1;
    for (int k = 0; k < 3; ++k) { e.position[k] = position[k]; e.radiance[k] = intensity[k]; }
    emitters.push_back(e); committed = false; return (int) emitters.size() - 1;
}
int Scene::add_emitter_constant(const float *radiance) {
    if (env_emitter >= 0) AM_THROW("Only one environment emitter can be specified per scene.");      // scene.cpp:52-54
    DEmitter e{}; e.type = EMITTER_CONSTANT; e.mesh = -1; for (int k = 0; k < 3; ++k) e.radiance[k] = radiance[k];
    emitters.push_back(e); env_emitter = (int) emitters.size() - 1; committed = false; return env_emitter;
}
int Scene::add_emitter_envmap(const float *rgb, uint32_t w, uint32_t h, float scale, const M4d &to_world) {
    if (env_emitter >= 0) AM_THROW("Only one environment emitter can be specified per scene.");
    env.build(rgb, w, h, scale, to_world);
    DEmitter e{}; e.type = EMITTER_ENVMAP; e.mesh = -1; e.radiance[0] = e.radiance[1] = e.radiance[2] = scale;
    emitters.push_back(e); env_emitter = (int) emitters.size() - 1; committed = false; return env_emitter;
}
// Scene::bbox() over the shapes (Mesh::recompute_bbox: every vertex position, src/render/mesh.cpp) -> BoundingBox::bounding_sphere
// (include/mitsuba/core/bbox.h:343-346) -> radius = max(RayEpsilon, r * (1 + RayEpsilon)) (constant.cpp:73-88)
void Scene::bounding_sphere(float *c, float *radius) const {
    float lo[3] = { AM_INF, AM_INF, AM_INF }, hi[3] = { -AM_INF, -AM_INF, -AM_INF }; bool any = false;
    for (auto &m : meshes) for (size_t i = 0; i + 2 < m.pos.size(); i += 3) for (int k = 0; k < 3; ++k) { lo[k] = min_(lo[k], m.pos[i + k]); hi[k] = max_(hi[k], m.pos[i + k]); any = true; }
    if (!any) { c[0] = c[1] = c[2] = 0.f; *radius = AM_RAY_EPSILON; return; }
    for (int k = 0; k < 3; ++k) c[k] = (hi[k] + lo[k]) * .5f;                        // bbox.h center(): (max + min) * 0.5
    const float r = norm(v3(c[0] - hi[0], c[1] - hi[1], c[2] - hi[2]));
    *radius = max_(AM_RAY_EPSILON, r * (1.f + AM_RAY_EPSILON));
}

} // namespace amvpt
