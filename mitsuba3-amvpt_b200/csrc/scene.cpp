// scene.cpp -- host scene objects of the AMVPT path and their upload into HBM.
#include "host.h"
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <atomic>
#include <thread>
#include <tuple>
#include <cstring>

namespace amvpt {

std::string strfmt(const char *fmt, ...) {
    char buf[2048]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap); return buf;
}
void cuda_check(cudaError_t e, const char *what) {
    if (e != cudaSuccess) AM_THROW("CUDA error %s: %s", what, cudaGetErrorString(e));
}

// ---------------------------------------------------------------- double-precision description transforms
M4d m4d_translate(double x, double y, double z) { M4d r = m4d_identity(); r.m[0][3] = x; r.m[1][3] = y; r.m[2][3] = z; return r; }
M4d m4d_scale(double x, double y, double z) { M4d r = m4d_identity(); r.m[0][0] = x; r.m[1][1] = y; r.m[2][2] = z; return r; }
// dr::rotate<Matrix4>(axis, angle) (ext/drjit/include/drjit/transform.h:43-65); axis is normalised by the XML loader
M4d m4d_rotate(double ax, double ay, double az, double angle_deg) {
    double len = std::sqrt(ax * ax + ay * ay + az * az); ax /= len; ay /= len; az /= len;
    double th = angle_deg * (3.14159265358979323846 / 180.0), s = std::sin(th), c = std::cos(th), cm = 1.0 - c;
    double a[3] = { ax, ay, az }, s1[3] = { ay, az, ax }, s2[3] = { az, ax, ay }, t0[3], t1[3], t2[3];
    for (int i = 0; i < 3; ++i) {
        t0[i] = std::fma(a[i] * a[i], cm, c);
        t1[i] = std::fma(a[i] * s1[i], cm, s2[i] * s);
        t2[i] = std::fma(a[i] * s2[i], cm, -(s1[i] * s));
    }
    M4d r = m4d_identity();
    r.m[0][0] = t0[0]; r.m[0][1] = t2[1]; r.m[0][2] = t1[2];
    r.m[1][0] = t1[0]; r.m[1][1] = t0[1]; r.m[1][2] = t2[2];
    r.m[2][0] = t2[0]; r.m[2][1] = t1[1]; r.m[2][2] = t0[2];
    return r;
}
// Transform::look_at (include/mitsuba/core/transform.h:269-300)
M4d m4d_look_at(const double *o, const double *t, const double *up) {
    auto nrm = [](double *v) { double l = 1.0 / std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); v[0] *= l; v[1] *= l; v[2] *= l; };
    auto crs = [](const double *a, const double *b, double *r) { r[0] = a[1] * b[2] - a[2] * b[1]; r[1] = a[2] * b[0] - a[0] * b[2]; r[2] = a[0] * b[1] - a[1] * b[0]; };
    double dir[3] = { t[0] - o[0], t[1] - o[1], t[2] - o[2] }, left[3], nup[3];
    nrm(dir); crs(up, dir, left); nrm(left); crs(dir, left, nup);
    M4d r = m4d_identity();
    for (int i = 0; i < 3; ++i) { r.m[i][0] = left[i]; r.m[i][1] = nup[i]; r.m[i][2] = dir[i]; r.m[i][3] = o[i]; }
    return r;
}
M4d m4d_inverse(const M4d &m) {
    double a[4][8];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { a[i][j] = m.m[i][j]; a[i][4 + j] = i == j ? 1.0 : 0.0; }
    for (int c = 0; c < 4; ++c) {
        int p = c; for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
        if (a[p][c] == 0.0) AM_THROW("singular transform");
        if (p != c) for (int j = 0; j < 8; ++j) std::swap(a[p][j], a[c][j]);
        double inv = 1.0 / a[c][c]; for (int j = 0; j < 8; ++j) a[c][j] *= inv;
        for (int r = 0; r < 4; ++r) if (r != c) { double f = a[r][c]; if (f != 0.0) for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j]; }
    }
    M4d r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = a[i][4 + j];
    return r;
}

// ---------------------------------------------------------------- meshes
// Mesh::build_pmf (src/render/mesh.cpp:444-486): table[i] = 0.5 * |cross(p1 - p0, p2 - p0)|; DiscreteDistribution::update
// (include/mitsuba/core/distr_1d.h:233-264): cdf accumulated in double, stored as float; normalization = 1 / sum
void HMesh::build_area_distr() {
    uint32_t nf = nfaces(); pmf.resize(nf); cdf.resize(nf); double acc = 0.0;
    for (uint32_t i = 0; i < nf; ++i) {
        const float *a = &pos[3 * faces[3 * i]], *b = &pos[3 * faces[3 * i + 1]], *c = &pos[3 * faces[3 * i + 2]];
        V3 p0 = v3(a[0], a[1], a[2]), p1 = v3(b[0], b[1], b[2]), p2 = v3(c[0], c[1], c[2]);
        pmf[i] = .5f * norm(cross(p1 - p0, p2 - p0)); acc += (double) pmf[i]; cdf[i] = (float) acc;
    }
    sum = nf ? cdf[nf - 1] : 0.f; normalization = rcp(sum);
}

// ---------------------------------------------------------------- BSDF preparation
// include/mitsuba/core/quad.h:27-86 gauss_legendre (double recurrences, float nodes / weights)
static void legendre_pd(int l, double x, double *p, double *d) {
    if (l == 0) { *p = 1; *d = 0; return; }
    if (l == 1) { *p = x; *d = 1; return; }
    double l_ppred = 1, l_pred = x, l_cur = 0, d_ppred = 0, d_pred = 1, d_cur = 0, k0 = 3, k1 = 2, k2 = 1;
    for (int k = 2; k <= l; ++k) {
        l_cur = (k0 * x * l_pred - k2 * l_ppred) / k1; d_cur = d_ppred + k0 * l_pred;
        l_ppred = l_pred; l_pred = l_cur; d_ppred = d_pred; d_pred = d_cur; k2 = k1; k0 += 2; k1 += 1;
    }
    *p = l_cur; *d = d_cur;
}
static void gauss_legendre(int n, std::vector<float> &nodes, std::vector<float> &weights) {
    nodes.assign(n, 0.f); weights.assign(n, 0.f); n--;
    if (n == 0) { nodes[0] = 0; weights[0] = 2; }
    else if (n == 1) { nodes[0] = (float) -std::sqrt(1.0 / 3.0); nodes[1] = -nodes[0]; weights[0] = weights[1] = 1; }
    int m = (n + 1) / 2;
    for (int i = 0; i < m; ++i) {
        double x = -std::cos((double) (2 * i + 1) / (double) (2 * n + 2) * 3.14159265358979323846);
        int it = 0;
        while (true) {
            if (++it > 20) break;
            double p, d; legendre_pd(n + 1, x, &p, &d);
            double step = p / d; x -= step;
            if (std::fabs(step) <= 4 * std::fabs(x) * 0x1p-53) break;
        }
        double p, d; legendre_pd(n + 1, x, &p, &d);
        weights[i] = weights[n - i] = (float) (2 / ((1 - x * x) * (d * d)));
        nodes[i] = (float) x; nodes[n - i] = (float) -x;
    }
    if ((n % 2) == 0) { double p, d; legendre_pd(n + 1, 0.0, &p, &d); weights[n / 2] = (float) (2.0 / (d * d)); nodes[n / 2] = 0; }
}
// MicrofacetDistribution eval_reflectance / eval_transmittance (include/mitsuba/render/microfacet.h:464-560)
static float mf_quadrature(const Microfacet &d, V3 wi, float eta, bool transmit) {
    int res = eta > 1 ? 32 : 128;
    std::vector<float> nodes, weights; gauss_legendre(res, nodes, weights);
    float result = 0.f;
    for (int jy = 0; jy < res; ++jy) for (int jx = 0; jx < res; ++jx) {
        V2 node = v2(fmadd(nodes[jx], 0.5f, 0.5f), fmadd(nodes[jy], 0.5f, 0.5f));
        float pdf; V3 m = d.sample(wi, node, &pdf);
        float f, ct, eit, eti; fresnel(dot(wi, m), eta, &f, &ct, &eit, &eti);
        float smith;
        if (transmit) { V3 wo = refract(wi, m, ct, eti); smith = d.smith_g1(wo, m) * (1.f - f); if (wo.z * wi.z >= 0.f) smith = 0.f; }
        else { V3 wo = reflect(wi, m); smith = d.smith_g1(wo, m) * f; if (wo.z <= 0.f || wi.z <= 0.f) smith = 0.f; }
        result += smith * (weights[jx] * weights[jy]) * 0.25f;
    }
    return result;
}
void bsdf_prepare(DBsdf &b) {
    b.flags = bsdf_flags(b.type);
    if (b.type != BSDF_ROUGHPLASTIC) return;
    b.inv_eta_2 = 1.f / (b.ior_eta * b.ior_eta);
    float d_mean = (b.reflectance[0] + b.reflectance[1] + b.reflectance[2]) / 3.f;
    float s_mean = (b.spec_refl[0] + b.spec_refl[1] + b.spec_refl[2]) / 3.f;
    b.spec_sampling_weight = s_mean / (d_mean + s_mean);
    // The two tables cost 128 quadratures of 32 x 32 nodes (0.4 s on one core) and depend only on (distribution, alpha, eta): they are computed
    // on several threads (each entry independently, the sum taken in index order afterwards: same bits) and kept per process, so that a frame
    // loop that rebuilds its scene does not recompute them.
    struct Key { int distr; uint32_t alpha, eta; bool operator<(const Key &o) const { return std::tie(distr, alpha, eta) < std::tie(o.distr, o.alpha, o.eta); } };
    struct Val { float ext[AM_ROUGH_TRANSMITTANCE_RES]; float internal; };
    static std::mutex cache_mutex; static std::map<Key, Val> cache;
    Key key{ (int) b.distr, 0u, 0u }; memcpy(&key.alpha, &b.alpha_u, 4); memcpy(&key.eta, &b.ior_eta, 4);
    { std::lock_guard<std::mutex> l(cache_mutex); auto it = cache.find(key);
      if (it != cache.end()) { memcpy(b.ext_transmittance, it->second.ext, sizeof(it->second.ext)); b.internal_reflectance = it->second.internal; return; } }
    Microfacet d(b.distr, b.alpha_u, b.alpha_u);
    float term[AM_ROUGH_TRANSMITTANCE_RES];
    auto entry = [&](int i) {
        float mu = max_(1e-6f, fmadd((float) i, 1.f / (float) (AM_ROUGH_TRANSMITTANCE_RES - 1), 0.f));
        V3 wi = v3(fsqrt(1.f - mu * mu), 0.f, mu);
        b.ext_transmittance[i] = mf_quadrature(d, wi, b.ior_eta, true);
        term[i] = mf_quadrature(d, wi, 1.f / b.ior_eta, false) * wi.z;
    };
    const unsigned n_threads = std::min(8u, std::max(1u, std::thread::hardware_concurrency()));
    std::atomic<int> next{ 0 };
    auto worker = [&]() { for (int i; (i = next.fetch_add(1)) < AM_ROUGH_TRANSMITTANCE_RES;) entry(i); };
    std::vector<std::thread> pool; for (unsigned t = 1; t < n_threads; ++t) pool.emplace_back(worker);
    worker(); for (auto &t : pool) t.join();
    float acc = 0.f;
    for (int i = 0; i < AM_ROUGH_TRANSMITTANCE_RES; ++i) acc += term[i];
    b.internal_reflectance = acc / (float) AM_ROUGH_TRANSMITTANCE_RES * 2.f;
    { Val v; memcpy(v.ext, b.ext_transmittance, sizeof(v.ext)); v.internal = b.internal_reflectance; std::lock_guard<std::mutex> l(cache_mutex); cache[key] = v; }
}

// ---------------------------------------------------------------- Scene
// Device block cache.  A frame loop that builds a scene per frame (bench.py's end-to-end step, the reference's viewer) allocates and frees the
// same film-sized blocks every frame; cudaMalloc / cudaFree of a 268 MB block cost about a millisecond each and cudaFree synchronises the
// device -- and a commit makes about twenty small allocations besides.  Blocks released by a scene are kept per (device, size) and handed to the next scene that asks for exactly that size;
// at most AMVPT_BLOCK_CACHE_MB (default 8192) are kept per process, AMVPT_BLOCK_CACHE_MB=0 switches the cache off.  Contents are stale:
// every consumer initialises what it reads (film_clear, uploads), exactly as with cudaMalloc, which does not zero either.
namespace {
struct BlockCache {
    std::mutex m; std::multimap<std::pair<int, size_t>, void *> blocks; size_t held = 0, limit;
    BlockCache() { const char *e = getenv("AMVPT_BLOCK_CACHE_MB"); limit = (size_t) (e ? atoll(e) : 8192) << 20; }
    ~BlockCache() { /* process exit: the driver reclaims device memory; no CUDA calls during static destruction */ }
    void *take(int dev, size_t bytes) {
        std::lock_guard<std::mutex> l(m); auto it = blocks.find({ dev, bytes });
        if (it == blocks.end()) return nullptr; void *p = it->second; blocks.erase(it); held -= bytes; return p;
    }
    bool give(int dev, size_t bytes, void *p) {
        std::lock_guard<std::mutex> l(m); if (held + bytes > limit || blocks.size() >= 65536) return false;
        blocks.insert({ { dev, bytes }, p }); held += bytes; return true;
    }
};
BlockCache &block_cache() { static BlockCache *c = new BlockCache(); return *c; }
} // namespace

void Scene::release_device_blocks() {
    if (device < 0) { allocs.clear(); alloc_bytes.clear(); return; }
    cudaSetDevice(device);
    bool synced = false;
    for (size_t i = 0; i < allocs.size(); ++i) {
        const size_t bytes = i < alloc_bytes.size() ? alloc_bytes[i] : 0;
        if (bytes && block_cache().limit) {
            if (!synced) { cudaDeviceSynchronize(); synced = true; }       // no kernel of this scene may still touch a block that changes hands
            if (block_cache().give(device, bytes, allocs[i])) continue;
        }
        cudaFree(allocs[i]);
    }
    allocs.clear(); alloc_bytes.clear();
}
Scene::~Scene() {
    release_device_blocks();
    if (device >= 0 && adapt_compact) cudaFree(adapt_compact);
}
void *Scene::dalloc(size_t bytes) {
    if (!bytes) bytes = 16;
    void *p = device >= 0 ? block_cache().take(device, bytes) : nullptr;
    if (!p) {
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess && device >= 0) {          // out of memory with blocks parked in the cache: give them back and retry once
            cudaGetLastError();
            for (;;) { void *q = nullptr; { std::lock_guard<std::mutex> l(block_cache().m); auto &b = block_cache().blocks; if (b.empty()) break; q = b.begin()->second; block_cache().held -= b.begin()->first.second; b.erase(b.begin()); } cudaFree(q); }
            e = cudaMalloc(&p, bytes);
        }
        AM_CUDA(e);
    }
    allocs.push_back(p); alloc_bytes.push_back(bytes); return p;
}
int Scene::add_bsdf(const DBsdf &b_, const std::string &id) {
    DBsdf b = b_; bsdf_prepare(b); bsdfs.push_back(b); bsdf_ids.push_back(id); committed = false;
    return (int) bsdfs.size() - 1;
}
int Scene::add_mesh(HMesh &&m, const float *radiance) {
    if (m.bsdf < 0 || m.bsdf >= (int) bsdfs.size()) AM_THROW("add_mesh: invalid BSDF index %d", m.bsdf);
    uint32_t nv = (uint32_t) (m.pos.size() / 3);
    for (uint32_t f : m.faces) if (f >= nv) AM_THROW("add_mesh: face index %u out of range (%u vertices)", f, nv);
    if (!m.nrm.empty() && m.nrm.size() != m.pos.size()) AM_THROW("add_mesh: normal count does not match vertex count");
    m.prim_offset = nprims();
    int id = (int) meshes.size();
    if (radiance) { m.emitter = (int) emitters.size(); DEmitter e{}; e.type = EMITTER_AREA; e.mesh = id; e.radiance[0] = radiance[0]; e.radiance[1] = radiance[1]; e.radiance[2] = radiance[2]; emitters.push_back(e); }
    else m.emitter = -1;
    m.build_area_distr();
    meshes.push_back(std::move(m)); committed = false; host_ready = false;
    return id;
}
// GridSensor ctor placement loop (src/sensors/grid.cpp:182-226) + PerspectiveCamera / ThinLensCamera ctor and
// update_camera_transforms (src/sensors/perspective.cpp:139-181, src/sensors/thinlens.cpp:152-187)
void Scene::set_sensor_grid(const GridSensorParams &g) {
    if (g.grid_x == 0 || g.grid_y == 0) AM_THROW("grid: grid_x / grid_y must be positive");
    if (g.res_x == 0 || g.res_y == 0 || g.res_x % g.grid_x || g.res_y % g.grid_y) AM_THROW("Film size must be divisible by grid dimensions !");
    grid_x = g.grid_x; grid_y = g.grid_y; reverse_x = g.reverse_x; reverse_y = g.reverse_y; thinlens = g.thinlens; multi_sensor = true;
    film_w = g.res_x; film_h = g.res_y; res_x = g.res_x / g.grid_x; res_y = g.res_y / g.grid_y;
    double sub_asp = res_x / (double) res_y;
    double fov_x = parse_fov_x(g.fov, g.fov_axis, sub_asp);
    M4d inv_d = m4d_inverse(g.to_world);
    M4 inv_t = m4_transpose(m4_from_d(inv_d));          // to_world.inverse_transpose, cast to float
    uint32_t n = grid_x * grid_y;
    float off_y = -g.cam_off[1], off_z = -g.cam_off[2];
    sensors.clear();
    for (uint32_t i = 0; i < n; ++i) {
        float dt = (float) i / (float) (n - 1), lens_shift = 0.f;
        M4 itrafo = inv_t;
        if (g.single) {
        } else if (g.use_cone) {
            float tan_off = tan_((dt - 0.5f) * (g.cone_deg * (AM_PI / 180.f)));
            float offset = g.focus_distance * tan_off;
            double shift = 0.5 * (double) tan_off / std::tan((fov_x * (3.14159265358979323846 / 180.0)) * 0.5);
            itrafo.m[3][0] += offset + g.cam_off[0]; itrafo.m[3][1] += off_y; itrafo.m[3][2] += off_z;
            lens_shift = (float) shift;
        } else {
            float sc = g.cam_dist * (dt - 0.5f * (float) (g.cam_center ? 1 : 0));
            itrafo.m[3][0] += g.cam_off[0] + g.cam_dir[0] * sc;
            itrafo.m[3][1] += off_y + g.cam_dir[1] * sc;
            itrafo.m[3][2] += off_z + g.cam_dir[2] * sc;
        }
        M4 tw = g.single ? m4_from_d(g.to_world) : m4_inverse(m4_transpose(itrafo));
        M4 s2c = sample_to_camera((int) res_x, (int) res_y, (float) fov_x, g.near_clip, g.far_clip, lens_shift);
        DSensor c;
        for (int r = 0; r < 4; ++r) for (int k = 0; k < 4; ++k) c.s2c[4 * r + k] = s2c.m[r][k];
        for (int r = 0; r < 3; ++r) for (int k = 0; k < 4; ++k) c.to_world[4 * r + k] = tw.m[r][k];
        c.near_clip = g.near_clip; c.far_clip = g.far_clip; c.aperture_radius = g.aperture_radius; c.focus_distance = g.focus_distance;
        sensors.push_back(c);
    }
    committed = false; host_ready = false;
}
// BatchSensor (src/sensors/batch.cpp:94-130): n explicit sub-cameras side by side
void Scene::set_sensor_explicit(uint32_t n, uint32_t rx, uint32_t ry, const float *s2c, const float *to_world,
                                float near_clip, float far_clip, bool rev_x) {
    if (!n || !rx || !ry) AM_THROW("batch: need at least one sub-sensor and a non-empty film");
    grid_x = n; grid_y = 1; reverse_x = rev_x; reverse_y = false; thinlens = false; multi_sensor = true;
    res_x = rx; res_y = ry; film_w = rx * n; film_h = ry; sensors.clear();
    for (uint32_t i = 0; i < n; ++i) {
        DSensor c; memcpy(c.s2c, s2c + 16 * i, 64); memcpy(c.to_world, to_world + 16 * i, 48);
        c.near_clip = near_clip; c.far_clip = far_clip; c.aperture_radius = 0.f; c.focus_distance = far_clip; sensors.push_back(c);
    }
    committed = false;
}
// Any MultiSensor as a table of sub-cameras (MultiSensor::sensors(), grid_dim(), reverse_axis(); src/sensors/grid.cpp:435-457): what a plugin
// shim inside the reference hands over.  scalars4[i] = near_clip, far_clip, aperture_radius, focus_distance of sub-camera i
void Scene::set_sensor_table(uint32_t gx, uint32_t gy, uint32_t rx, uint32_t ry, bool rev_x, bool rev_y, bool thin, const float *s2c, const float *to_world, const float *scalars4) {
    if (!gx || !gy || !rx || !ry) AM_THROW("sensor table: need at least one sub-sensor and a non-empty film");
    grid_x = gx; grid_y = gy; reverse_x = rev_x; reverse_y = rev_y; thinlens = thin; multi_sensor = true;
    res_x = rx; res_y = ry; film_w = rx * gx; film_h = ry * gy; sensors.clear();
    for (uint32_t i = 0; i < gx * gy; ++i) {
        DSensor c; memcpy(c.s2c, s2c + 16 * i, 64); memcpy(c.to_world, to_world + 16 * i, 48);
        c.near_clip = scalars4[4 * i]; c.far_clip = scalars4[4 * i + 1]; c.aperture_radius = scalars4[4 * i + 2]; c.focus_distance = scalars4[4 * i + 3]; sensors.push_back(c);
    }
    committed = false;
}
// ReconstructionFilter plugins of the film (src/rfilters/{box,gaussian,tent,mitchell,catmullrom,lanczos}.cpp)
void Scene::set_rfilter(int type, float param, float param2) {
    if (type < 0 || type > 5) AM_THROW("rfilter: unknown filter type %d (0 box, 1 gaussian, 2 tent, 3 mitchell, 4 catmullrom, 5 lanczos)", type);
    if ((type == 1 || type == 2) && !(param > 0.f)) AM_THROW("rfilter: the %s must be positive", type == 1 ? "standard deviation" : "radius");
    if (type == 3 && !(std::isfinite(param) && std::isfinite(param2))) AM_THROW("rfilter: mitchell B and C must be finite");
    if (type == 5 && !(param >= 1.f && param == std::floor(param))) AM_THROW("rfilter: lanczos lobes must be a positive integer");
    float radius = type == 1 ? 4.f * param : (type == 2 || type == 5 ? param : (type == 3 || type == 4 ? 2.f : .5f));
    if (std::ceil(radius - .5f) > 7.f) AM_THROW("rfilter: radius %g is too large (footprint limited to 15 x 15 pixels)", radius);
    rfilter_type = type; rfilter_param = param; rfilter_param2 = param2; committed = false;
}
uint32_t Scene::reuse_samples() const {
    const uint32_t n_sensors = (uint32_t) sensors.size();
    if (!(integrator.sa_reuse && n_sensors > 1 && integrator.reuse_count != 1)) return 0;
    uint32_t n = std::min(integrator.reuse_count, n_sensors);
    if (n == 0 || n_sensors % n) {        // "The number of sensors should be divisible by reused samples count!" (mvpath.cpp:196-217)
        n = 0;
        for (uint32_t p = 8; p < n_sensors; ++p) if (n_sensors % p == 0) { n = p; break; }
        if (!n) { for (uint32_t p = 8; p > 1; --p) if (n_sensors % p == 0) { n = p; break; } n = n ? n : n_sensors; }
    }
    return n;
}
void Scene::set_reuse(bool sa_reuse, bool sa_mis, bool fast_mis, uint32_t reuse_count, uint32_t adaptive) {
    integrator.sa_reuse = sa_reuse; integrator.sa_mis = sa_mis; integrator.fast_mis = fast_mis; integrator.reuse_count = reuse_count;
    integrator.adaptive = adaptive;
    if (reuse_samples() > 64) AM_THROW("mvpath: reuse_count %u exceeds the 64 cameras per group this build keeps per lane", reuse_samples());
    if (committed) { d.n_reuse = reuse_samples(); d.sa_mis = sa_mis; d.fast_mis = fast_mis; d.n_adapt = d.n_reuse ? std::min(adaptive, d.n_reuse - 1u) : 0u; d.adapt_w = 1.f / (float) (d.n_adapt + 1u); d.reuse_debug = integrator.debug ? 1 : 0; }
}
// View tiles of the quilt owned by one rank: view v = ix + grid_x * iy (sub-sensor index), tile position follows
// GridSensor::sample_ray's axis reversal (grid.cpp:246-255).
std::vector<TileDesc> Scene::view_tiles(const Shard &sh) const {
    std::vector<TileDesc> out;
    for (uint32_t ty = 0; ty < grid_y; ++ty) for (uint32_t tx = 0; tx < grid_x; ++tx) {
        uint32_t ix = reverse_x ? grid_x - 1 - tx : tx, iy = reverse_y ? grid_y - 1 - ty : ty, view = ix + grid_x * iy;
        if (view % sh.world != sh.rank) continue;
        TileDesc t{}; t.x0 = tx * res_x; t.y0 = ty * res_y; t.w = res_x; t.h = res_y; t.spp = 0; t.pad = view; t.lane_begin = 0;
        out.push_back(t);
    }
    std::sort(out.begin(), out.end(), [](const TileDesc &a, const TileDesc &b) { return a.pad < b.pad; });
    return out;
}
void Scene::film_clear(cudaStream_t st) {
    if (!committed) AM_THROW("film_clear: scene not committed");
    AM_CUDA(cudaMemsetAsync(film, 0, sizeof(float4) * (size_t) film_w * film_h, st));
    AM_CUDA(cudaMemsetAsync(moments, 0, sizeof(float2) * (size_t) film_w * film_h, st));
}
void Scene::prepare_host() {
    if (host_ready && !bvh.nodes.empty()) return;
    // scene scale for the conservative box padding: largest absolute coordinate of geometry and camera origins
    float scale = 0.f;
    for (auto &m : meshes) for (float v : m.pos) scale = std::max(scale, std::fabs(v));
    for (auto &c : sensors) for (int r = 0; r < 3; ++r) scale = std::max(scale, std::fabs(c.to_world[4 * r + 3]));
    bvh = BvhOut(); build_bvh(*this, scale, bvh); host_ready = true;
}
void Scene::commit(int dev) {
    if (sensors.empty()) AM_THROW("commit: no sensor");
    if ((unsigned long long) film_w * film_h * 4ull >= (1ull << 32)) AM_THROW("commit: film index exceeds 32 bits");
    if (device >= 0 && device != dev) AM_THROW("commit: scene already lives on device %d", device);
    AM_CUDA(cudaSetDevice(dev)); device = dev;
    release_device_blocks();
    wf = WfPool{}; reuse_buf = nullptr; reuse_cap = 0; split_buf = nullptr; split_cap = 0; srgb8 = nullptr;
    cudaDeviceProp prop; AM_CUDA(cudaGetDeviceProperties(&prop, dev)); sm_count = prop.multiProcessorCount;
    prepare_host();
    n_nodes = (uint32_t) (bvh.nodes.size() / AM_NODE_F4); n_tris = (uint32_t) (bvh.tris.size() / 3);
    std::vector<DMesh> dm; std::vector<float> cdf_all, pmf_all; bool any_nrm = false;
    for (auto &m : meshes) {
        DMesh q{}; q.bsdf = m.bsdf; q.emitter = m.emitter; q.flip_normals = m.flip_normals; q.has_normals = m.nrm.empty() ? 0 : 1;
        q.prim_offset = m.prim_offset; q.nfaces = m.nfaces(); q.normalization = m.normalization; q.sum = m.sum; q.cdf_offset = 0;
        if (m.emitter >= 0) { q.cdf_offset = (uint32_t) cdf_all.size(); cdf_all.insert(cdf_all.end(), m.cdf.begin(), m.cdf.end()); pmf_all.insert(pmf_all.end(), m.pmf.begin(), m.pmf.end()); }
        any_nrm |= !m.nrm.empty(); dm.push_back(q);
    }
    d = DScene{};
    // nodes | triangles | compressed nodes in ONE allocation: the wavefront trace stage asks L2 to keep this range resident
    // (cudaAccessPropertyPersisting window, wavefront.cu) while the path-state pool streams through
    { const size_t nb = bvh.nodes.size() * sizeof(float4), tb = bvh.tris.size() * sizeof(float4), qb = bvh.nodes_q4.size() * sizeof(uint4);
      const size_t o_t = (nb + 255) / 256 * 256, o_q = o_t + (tb + 255) / 256 * 256; bvh_bytes = o_q + qb; bvh_dev = (char *) dalloc(bvh_bytes);
      if (nb) AM_CUDA(cudaMemcpy(bvh_dev, bvh.nodes.data(), nb, cudaMemcpyHostToDevice));
      if (tb) AM_CUDA(cudaMemcpy(bvh_dev + o_t, bvh.tris.data(), tb, cudaMemcpyHostToDevice));
      if (qb) AM_CUDA(cudaMemcpy(bvh_dev + o_q, bvh.nodes_q4.data(), qb, cudaMemcpyHostToDevice));
      d.nodes = (const float4 *) bvh_dev; d.tris = (const float4 *) (bvh_dev + o_t); d.nodes_q4 = qb ? (const uint4 *) (bvh_dev + o_q) : nullptr; bvh_hot_bytes = o_q; }
    d.tri_nrm = any_nrm ? upload(bvh.tri_nrm) : nullptr;
    { int v = 0; cudaDeviceGetAttribute(&v, cudaDevAttrMaxPersistingL2CacheSize, dev); l2_persist_max = (size_t) std::max(v, 0);
      cudaDeviceGetAttribute(&v, cudaDevAttrMaxAccessPolicyWindowSize, dev); l2_window_max = (size_t) std::max(v, 0); }
    d.flat = bvh.flat.empty() ? nullptr : upload(bvh.flat); d.n_flat = (uint32_t) (bvh.flat.size() / 2);
    d.prim_to_tri = upload(bvh.prim_to_tri); d.meshes = upload(dm); d.bsdfs = upload(bsdfs); d.emitters = upload(emitters);
    // PerspectiveCamera::update_camera_transforms (perspective.cpp:173-199) members read by sample_surface: camera_to_sample
    // (inverse of sample_to_camera), Transform4f::inverse of to_world, normalization = 1 / area of the near-plane image rectangle
    std::vector<DSensorAux> aux;
    for (auto &c : sensors) {
        DSensorAux a{}; M4 s2c, tw = m4_identity();
        std::memcpy(s2c.m, c.s2c, 64); std::memcpy(tw.m, c.to_world, 48);
        M4 c2s = m4_inverse(s2c), w2c = m4_inverse(tw);
        std::memcpy(a.c2s, c2s.m, 64); std::memcpy(a.w2c, w2c.m, 48);
        auto proj = [&](float x, float y, float z, float *o) {
            float r[4]; for (int i = 0; i < 4; ++i) r[i] = s2c.m[i][3];
            for (int i = 0; i < 4; ++i) r[i] = fmadd(s2c.m[i][0], x, r[i]);
            for (int i = 0; i < 4; ++i) r[i] = fmadd(s2c.m[i][1], y, r[i]);
            for (int i = 0; i < 4; ++i) r[i] = fmadd(s2c.m[i][2], z, r[i]);
            float iw = 1.f / r[3]; o[0] = r[0] * iw; o[1] = r[1] * iw; o[2] = r[2] * iw; };
        float pmin[3], pmax[3]; proj(0.f, 0.f, 0.f, pmin); proj(1.f, 1.f, 0.f, pmax);
        float ia = 1.f / pmin[2], ib = 1.f / pmax[2], ax = pmin[0] * ia, ay = pmin[1] * ia, bx = pmax[0] * ib, by = pmax[1] * ib;
        float lox = std::min(ax, bx), hix = std::max(ax, bx), loy = std::min(ay, by), hiy = std::max(ay, by);
        a.normalization = 1.f / ((hix - lox) * (hiy - loy));
        aux.push_back(a);
    }
    d.sensor_aux = upload(aux); d.n_reuse = reuse_samples(); d.sa_mis = integrator.sa_mis; d.fast_mis = integrator.fast_mis;
    d.n_adapt = d.n_reuse ? std::min(integrator.adaptive, d.n_reuse - 1u) : 0u; d.adapt_w = 1.f / (float) (d.n_adapt + 1u); d.reuse_debug = integrator.debug ? 1 : 0;
    adapt_mask = nullptr; adapt_scratch = nullptr; adapt_cap = 0; if (adapt_compact) { cudaFree(adapt_compact); adapt_compact = nullptr; adapt_compact_cap = 0; }
    d.sensors = upload(sensors); d.area_cdf = upload(cdf_all); d.area_pmf = upload(pmf_all);
    d.n_tris = n_tris; d.n_nodes = n_nodes; d.n_emitters = (uint32_t) emitters.size(); d.n_sensors = (uint32_t) sensors.size();
    d.grid_x = grid_x; d.grid_y = grid_y; d.reverse_x = reverse_x; d.reverse_y = reverse_y; d.thinlens = thinlens; d.shutter_draw = shutter_draw ? 1 : 0;
    d.film_w = film_w; d.film_h = film_h;
    d.max_depth = integrator.max_depth < 0 ? 0xffffffffu : (uint32_t) integrator.max_depth;   // integrator.cpp:503-520
    d.rr_depth = integrator.rr_depth;
    // emitters that are not area lights: environment data, scene bounding sphere (Emitter::set_scene)
    d.gen_emitters = general_emitters() ? 1 : 0; d.env_emitter = env_emitter; d.hide_emitters = integrator.hide_emitters ? 1 : 0;
    d.env_is_map = env_emitter >= 0 && emitters[(size_t) env_emitter].type == EMITTER_ENVMAP;
    bounding_sphere(d.bsphere, d.bsphere + 3);
    { float lo[3] = { AM_INF, AM_INF, AM_INF }, hi[3] = { -AM_INF, -AM_INF, -AM_INF };
      for (auto &m : meshes) for (size_t i = 0; i + 2 < m.pos.size(); i += 3) for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], m.pos[i + k]); hi[k] = std::max(hi[k], m.pos[i + k]); }
      for (int k = 0; k < 3; ++k) { d.box_lo[k] = lo[k] <= hi[k] ? lo[k] : 0.f; const float e = hi[k] - lo[k]; d.box_inv[k] = e > 0.f ? 16.f / e : 0.f; } }
    if (d.env_is_map) {
        d.env_data = upload(env.data); d.env_warp = upload(env.warp);
        d.env_res_x = env.res_x; d.env_res_y = env.res_y; d.env_n_levels = (uint32_t) env.level_off.size(); d.env_scale = env.scale;
        for (size_t i = 0; i < env.level_off.size(); ++i) { d.env_level_off[i] = env.level_off[i]; d.env_level_width[i] = env.level_width[i]; }
        for (int k = 0; k < 2; ++k) { d.env_patch_size[k] = env.patch_size[k]; d.env_inv_patch_size[k] = env.inv_patch_size[k]; d.env_max_patch[k] = env.max_patch[k]; }
        std::memcpy(d.env_to_world, env.to_world.m, 48); std::memcpy(d.env_from_world, env.from_world.m, 48);
    }
    d.filter_type = rfilter_type; d.filter_radius = .5f; d.filter_inv_radius = 1.f;
    if (rfilter_type == 1) {          // GaussianFilter ctor (gaussian.cpp:41-88): Remez fit to exp(-x / 2), rescaled by 1 / stddev^2 per degree
        static const double coeff[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                                          -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
        float stddev = rfilter_param; d.filter_radius = 4 * stddev; double scale = 1;
        for (int i = 0; i < 10; ++i) { d.filter_coeff[i] = (float) (coeff[i] * scale); scale /= (double) stddev * (double) stddev; }
        d.filter_coeff[0] -= estrin10(d.filter_radius * d.filter_radius, d.filter_coeff);
    } else if (rfilter_type == 2) { d.filter_radius = rfilter_param; d.filter_inv_radius = 1.f / rfilter_param; }
    else if (rfilter_type == 3) {     // MitchellNetravaliFilter::eval's scalar coefficients (mitchell.cpp:63-69)
        const float B = rfilter_param, C = rfilter_param2; d.filter_radius = 2.f;
        const float c[7] = { 12.f - 9.f * B - 6.f * C, -18.f + 12.f * B + 6.f * C, 6.f - 2.f * B,
                             -B - 6.f * C, 6.f * B + 30.f * C, -12.f * B - 48.f * C, 8.f * B + 24.f * C };
        for (int i = 0; i < 7; ++i) d.filter_coeff[i] = c[i];
    } else if (rfilter_type == 4) d.filter_radius = 2.f;
    else if (rfilter_type == 5) d.filter_radius = rfilter_param;
    size_t npix = (size_t) film_w * film_h;
    film = (float4 *) dalloc(npix * sizeof(float4)); moments = (float2 *) dalloc(npix * sizeof(float2)); rgb = nullptr;
    counters = (unsigned long long *) dalloc(sizeof(unsigned long long) * CTR_COUNT);
    tiles_cap = std::max<uint32_t>(grid_x * grid_y, 1u) + 1; tiles_dev = (TileDesc *) dalloc(sizeof(TileDesc) * tiles_cap);
    view_stats_dev = (float *) dalloc(sizeof(float) * 4 * tiles_cap);
    committed = true;
    film_clear(nullptr);
    AM_CUDA(cudaMemset(counters, 0, sizeof(unsigned long long) * CTR_COUNT));
    AM_CUDA(cudaDeviceSynchronize());
}

} // namespace amvpt
