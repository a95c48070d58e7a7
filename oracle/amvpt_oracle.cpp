// oracle/amvpt_oracle.cpp -- TEST INFRASTRUCTURE ONLY (parity checker + CPU baseline; never linked
// into or called from the product).  CPU restatement of the AMVPT render loop of the reference
// (xacond00/mitsuba3-amvpt) with sa_reuse = false, JIT (wavefront) semantics and IEEE-754 binary32
// leaves ("canonical arithmetic").  Each function cites the reference file:line
// it follows (paths relative to /root/reference).
//
// Pinning: leaf arithmetic (RNG, math, warps, Fresnel, microfacet, matrices, camera rays,
// Moeller-Trumbore, frames) is checked bit-for-bit against the reference's own headers through
// oracle/_ref (tests/test_oracle_vs_ref.py) and against the golden values of the reference's unit
// tests (tests/golden/).  The render loop itself (mvpath) has no test in the reference: PARITY
// UNPINNED for mvpath as a whole, the grid sensor and the adaptive allocator.
//
// Build: make -C oracle oracle   (g++ -O2 -mfma -ffp-contract=off)
#include "o_bsdf.h"
#include "o_transform.h"
#include "o_display.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <functional>
#include <thread>
#include <vector>

namespace orc {

// ---------------------------------------------------------------- RNG
// include/mitsuba/core/random.h:76-91
static inline void tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *o0, uint32_t *o1) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    *o0 = v0; *o1 = v1;
}
// ext/drjit/include/drjit/random.h:59-78, 143-150
struct Pcg32 {
    uint64_t state, inc;
    void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u;
        next_u32(); state += initstate; next_u32();
    }
    uint32_t next_u32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dull + inc;
        uint32_t xs = (uint32_t) (((old >> 18) ^ old) >> 27), rot = (uint32_t) (old >> 59);
        return (xs >> rot) | (xs << ((-(int32_t) rot) & 31));
    }
    float next_f32() { return u2f((next_u32() >> 9) | 0x3f800000u) - 1.f; }
};
// src/render/sampler.cpp:125-144 (array branch): per-lane TEA -> (state, stream)
static inline Pcg32 sampler_seed(uint32_t seed_value, uint32_t lane) {
    uint32_t v0, v1; tea32(seed_value, lane, 4, &v0, &v1);
    Pcg32 r; r.seed(v0, v1); return r;
}

// ---------------------------------------------------------------- scene
struct Mesh {
    std::vector<float> pos, nrm;        // 3 floats per vertex; nrm empty = no vertex normals
    std::vector<uint32_t> faces;        // 3 indices per face
    int bsdf = 0, emitter = -1, flip_normals = 0;
    uint32_t prim_offset = 0;
    // area distribution (src/render/mesh.cpp:444-486 + include/mitsuba/core/distr_1d.h:233-264)
    std::vector<float> pmf, cdf; float sum = 0, normalization = 0;
    uint32_t nfaces() const { return (uint32_t) faces.size() / 3; }
    V3 vp(uint32_t i) const { return V3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]); }
    V3 vn(uint32_t i) const { return V3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]); }
};
#include "o_emitter.h"
// emitters in the order Scene::Scene collects them (scene.cpp:22-63): type 0 area (mesh), 1 point, 2 constant, 3 envmap
struct Emitter { int mesh; V3 radiance; int type = 0; V3 position; };
struct SubSensor {
    M4 s2c, to_world; float near_clip, far_clip, aperture_radius, focus_distance;
    // sample reuse (sample_surface): camera_to_sample, inverse of to_world, 1 / area of the image rectangle (perspective.cpp:173-199)
    M4 c2s, w2c; float normalization = 0.f;
};

struct BvhNode { float lo[3], hi[3]; int32_t left, right; uint32_t first, count; };

struct Scene {
    std::vector<Mesh> meshes; std::vector<Bsdf> bsdfs; std::vector<Emitter> emitters;
    EnvMap env; int env_emitter = -1; V3 bs_center; float bs_radius = RayEpsilon;     // environment emitter, Emitter::set_scene bounding sphere
    std::vector<SubSensor> sensors;
    uint32_t grid_x = 1, grid_y = 1; int reverse_x = 0, reverse_y = 1, thinlens = 0, shutter_draw = 0;
    uint32_t film_w = 0, film_h = 0;    // quilt size
    uint32_t max_depth = 0xffffffffu, rr_depth = 5; int hide_emitters = 0;
    uint32_t base_seed = 0;
    // cross-view sample reuse (mvpath.h:120-129): n_reuse = samples per primary sample after mvpath.cpp:192-218 (0 = off)
    int sa_mis = 0, fast_mis = 0; uint32_t n_reuse = 0; uint32_t adaptive = 0; int debug = 0;     // adaptive: fill-in samples of the reuse branch (mvpath.h:125)
    // reconstruction filter of the film: 0 box, 1 gaussian (src/rfilters/gaussian.cpp), 2 tent (src/rfilters/tent.cpp)
    int filter_type = 0; float filter_radius = 0.5f, filter_coeff[10] = { 0 }, filter_inv_radius = 1.f;
    // flattened primitives
    std::vector<uint32_t> prim_mesh;    // global prim -> mesh
    std::vector<BvhNode> nodes; std::vector<uint32_t> bvh_prims;
    uint32_t nprims() const { return (uint32_t) prim_mesh.size(); }
};

struct Ray { V3 o, d; float maxt; };
struct Hit { float t = Inf, u = 0, v = 0; uint32_t prim = 0xffffffffu; bool valid() const { return prim != 0xffffffffu; } };

// include/mitsuba/render/mesh.h:467-488
static inline bool moeller_trumbore(const Ray &ray, V3 p0, V3 p1, V3 p2, float maxt, float *t_o, float *u_o, float *v_o) {
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = cross(ray.d, e2);
    float inv_det = rcp(dot(e1, pvec));
    V3 tvec = ray.o - p0;
    float u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    float v = dot(ray.d, qvec) * inv_det;
    active = active && v >= 0.f && u + v <= 1.f;
    float t = dot(e2, qvec) * inv_det;
    active = active && t >= 0.f && t <= maxt;
    *t_o = t; *u_o = u; *v_o = v;
    return active;
}
static inline void prim_verts(const Scene &s, uint32_t prim, const Mesh **m_o, uint32_t *local, V3 *p0, V3 *p1, V3 *p2) {
    const Mesh &m = s.meshes[s.prim_mesh[prim]];
    uint32_t f = prim - m.prim_offset;
    *p0 = m.vp(m.faces[3 * f]); *p1 = m.vp(m.faces[3 * f + 1]); *p2 = m.vp(m.faces[3 * f + 2]);
    *m_o = &m; *local = f;
}
// Closest hit with the semantics of ShapeKDTree::ray_intersect_naive (include/mitsuba/render/
// kdtree.h:2445-2468): primitives in global order, `t <= maxt` inclusive, a hit shrinks maxt, so
// among equal t the LATER primitive wins.
static Hit intersect_brute(const Scene &s, const Ray &ray) {
    Hit h; float maxt = ray.maxt;
    for (uint32_t prim = 0; prim < s.nprims(); ++prim) {
        const Mesh *m; uint32_t f; V3 p0, p1, p2; prim_verts(s, prim, &m, &f, &p0, &p1, &p2);
        float t, u, v;
        if (moeller_trumbore(ray, p0, p1, p2, maxt, &t, &u, &v)) { maxt = t; h.t = t; h.u = u; h.v = v; h.prim = prim; }
    }
    return h;
}
static bool ray_test_brute(const Scene &s, const Ray &ray) {
    for (uint32_t prim = 0; prim < s.nprims(); ++prim) {
        const Mesh *m; uint32_t f; V3 p0, p1, p2; prim_verts(s, prim, &m, &f, &p0, &p1, &p2);
        float t, u, v;
        if (moeller_trumbore(ray, p0, p1, p2, ray.maxt, &t, &u, &v)) return true;
    }
    return false;
}

// ---- the oracle's own acceleration structure (NOT part of the reference semantics: a conservative
// BVH whose result is, by construction and by test, identical to intersect_brute; used so that the
// 1M-triangle configuration and the timed CPU baseline finish).
struct BuildPrim { float lo[3], hi[3], c[3]; uint32_t prim; };
static void bvh_build_rec(Scene &s, std::vector<BuildPrim> &bp, uint32_t begin, uint32_t end, int depth) {
    BvhNode n; for (int a = 0; a < 3; ++a) { n.lo[a] = Inf; n.hi[a] = -Inf; }
    float clo[3] = { Inf, Inf, Inf }, chi[3] = { -Inf, -Inf, -Inf };
    for (uint32_t i = begin; i < end; ++i) for (int a = 0; a < 3; ++a) {
        n.lo[a] = std::min(n.lo[a], bp[i].lo[a]); n.hi[a] = std::max(n.hi[a], bp[i].hi[a]);
        clo[a] = std::min(clo[a], bp[i].c[a]); chi[a] = std::max(chi[a], bp[i].c[a]);
    }
    n.left = n.right = -1; n.first = begin; n.count = end - begin;
    uint32_t self = (uint32_t) s.nodes.size(); s.nodes.push_back(n);
    if (end - begin <= 4 || depth > 60) return;
    int axis = 0; float ext = chi[0] - clo[0];
    for (int a = 1; a < 3; ++a) if (chi[a] - clo[a] > ext) { ext = chi[a] - clo[a]; axis = a; }
    uint32_t mid;
    if (!(ext > 0.f)) mid = (begin + end) / 2;
    else {
        // binned SAH, 16 bins
        const int NB = 16; struct Bin { float lo[3], hi[3]; uint32_t n; } bins[NB];
        for (auto &b : bins) { for (int a = 0; a < 3; ++a) { b.lo[a] = Inf; b.hi[a] = -Inf; } b.n = 0; }
        float k = NB * (1.f - 1e-6f) / ext;
        auto binof = [&](const BuildPrim &p) { int b = (int) ((p.c[axis] - clo[axis]) * k); return std::min(std::max(b, 0), NB - 1); };
        for (uint32_t i = begin; i < end; ++i) { Bin &b = bins[binof(bp[i])]; b.n++; for (int a = 0; a < 3; ++a) { b.lo[a] = std::min(b.lo[a], bp[i].lo[a]); b.hi[a] = std::max(b.hi[a], bp[i].hi[a]); } }
        auto area = [](const float *lo, const float *hi) { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return dx * dy + dy * dz + dz * dx; };
        float la[NB], ra[NB]; uint32_t ln[NB], rn[NB];
        { float lo[3] = { Inf, Inf, Inf }, hi[3] = { -Inf, -Inf, -Inf }; uint32_t c = 0;
          for (int i = 0; i < NB; ++i) { c += bins[i].n; for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], bins[i].lo[a]); hi[a] = std::max(hi[a], bins[i].hi[a]); } la[i] = c ? area(lo, hi) : 0.f; ln[i] = c; } }
        { float lo[3] = { Inf, Inf, Inf }, hi[3] = { -Inf, -Inf, -Inf }; uint32_t c = 0;
          for (int i = NB - 1; i >= 0; --i) { c += bins[i].n; for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], bins[i].lo[a]); hi[a] = std::max(hi[a], bins[i].hi[a]); } ra[i] = c ? area(lo, hi) : 0.f; rn[i] = c; } }
        int best = -1; float best_cost = Inf;
        for (int i = 0; i < NB - 1; ++i) { if (!ln[i] || !rn[i + 1]) continue; float c = la[i] * ln[i] + ra[i + 1] * rn[i + 1]; if (c < best_cost) { best_cost = c; best = i; } }
        if (best < 0) mid = (begin + end) / 2;
        else mid = (uint32_t) (std::partition(bp.begin() + begin, bp.begin() + end, [&](const BuildPrim &p) { return binof(p) <= best; }) - bp.begin());
        if (mid == begin || mid == end) mid = (begin + end) / 2;
    }
    int32_t l = (int32_t) s.nodes.size(); bvh_build_rec(s, bp, begin, mid, depth + 1);
    int32_t r = (int32_t) s.nodes.size(); bvh_build_rec(s, bp, mid, end, depth + 1);
    s.nodes[self].left = l; s.nodes[self].right = r;
}
static void bvh_build(Scene &s) {
    std::vector<BuildPrim> bp(s.nprims());
    for (uint32_t prim = 0; prim < s.nprims(); ++prim) {
        const Mesh *m; uint32_t f; V3 p[3]; prim_verts(s, prim, &m, &f, &p[0], &p[1], &p[2]);
        BuildPrim &b = bp[prim]; b.prim = prim;
        const float *pp[3] = { &p[0].x, &p[1].x, &p[2].x };
        for (int a = 0; a < 3; ++a) {
            float lo = std::min(pp[0][a], std::min(pp[1][a], pp[2][a])), hi = std::max(pp[0][a], std::max(pp[1][a], pp[2][a]));
            float pad = 1e-5f * std::max(std::fabs(lo), std::fabs(hi)) + 1e-7f * 0.f;
            float e = hi - lo; pad += 1e-5f * e + 1e-30f;
            b.lo[a] = lo - pad; b.hi[a] = hi + pad; b.c[a] = 0.5f * (lo + hi);
        }
    }
    s.nodes.clear(); s.nodes.reserve(2 * bp.size() + 1);
    if (!bp.empty()) bvh_build_rec(s, bp, 0, (uint32_t) bp.size(), 0);
    s.bvh_prims.resize(bp.size());
    for (size_t i = 0; i < bp.size(); ++i) s.bvh_prims[i] = bp[i].prim;
}
static inline bool slab(const BvhNode &n, const Ray &r, const float *inv_d, float tbest) {
    float tmin = 0.f, tmax = tbest;
    const float *o = &r.o.x;
    for (int a = 0; a < 3; ++a) {
        float t0 = (n.lo[a] - o[a]) * inv_d[a], t1 = (n.hi[a] - o[a]) * inv_d[a];
        if (t0 > t1) std::swap(t0, t1);
        if (t0 != t0 || t1 != t1) continue;      // 0 * inf: origin on a slab plane of an axis-parallel ray -> no constraint
        tmin = std::max(tmin, t0); tmax = std::min(tmax, t1);
    }
    return tmin * (1.f - 4e-7f) <= tmax * (1.f + 4e-7f);
}
struct TraverseStats { uint64_t nodes = 0, tris = 0; };
template <bool Any> static Hit intersect_bvh(const Scene &s, const Ray &ray, TraverseStats *st = nullptr) {
    Hit h; if (s.nodes.empty()) return h;
    float best = ray.maxt;
    float inv_d[3] = { 1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z };
    int32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const BvhNode &n = s.nodes[stack[--sp]];
        if (st) st->nodes++;
        if (!slab(n, ray, inv_d, best)) continue;
        if (n.left < 0) {
            for (uint32_t i = n.first; i < n.first + n.count; ++i) {
                uint32_t prim = s.bvh_prims[i];
                const Mesh *m; uint32_t f; V3 p0, p1, p2; prim_verts(s, prim, &m, &f, &p0, &p1, &p2);
                float t, u, v; if (st) st->tris++;
                if (moeller_trumbore(ray, p0, p1, p2, ray.maxt, &t, &u, &v)) {
                    if (Any) { h.prim = prim; h.t = t; return h; }
                    if (!h.valid() || t < h.t || (t == h.t && prim > h.prim)) { h.t = t; h.u = u; h.v = v; h.prim = prim; best = t; }
                }
            }
        } else { stack[sp++] = n.left; stack[sp++] = n.right; }
    }
    return h;
}
static inline Hit scene_intersect(const Scene &s, const Ray &r, bool brute, TraverseStats *st = nullptr) { return brute ? intersect_brute(s, r) : intersect_bvh<false>(s, r, st); }
static inline bool scene_ray_test(const Scene &s, const Ray &r, bool brute, TraverseStats *st = nullptr) { return brute ? ray_test_brute(s, r) : intersect_bvh<true>(s, r, st).valid(); }

// ---------------------------------------------------------------- surface interaction
struct SI {
    bool valid = false; float t = Inf; V3 p, n, sh_n, sh_s, sh_t, wi; const Mesh *mesh = nullptr; uint32_t prim = 0;
    V3 to_local(V3 v) const { return V3(dot(v, sh_s), dot(v, sh_t), dot(v, sh_n)); }                     // frame.h:34-36
    V3 to_world(V3 v) const { return fmadd(sh_n, v.z, fmadd(sh_t, v.y, sh_s * v.x)); }                    // frame.h:39-41
};
// include/mitsuba/render/interaction.h:165-169
static inline V3 offset_p(V3 p, V3 n, V3 d) {
    float mag = (1.f + fmax_(fmax_(fabs_(p.x), fabs_(p.y)), fabs_(p.z))) * RayEpsilon;
    mag = mulsign(mag, dot(n, d));
    return fmadd(n, mag, p);
}
// interaction.h:140-142
static inline Ray spawn_ray(V3 p, V3 n, V3 d) { return { offset_p(p, n, d), d, Largest }; }
// interaction.h:145-152
static inline Ray spawn_ray_to(V3 p, V3 n, V3 t) {
    V3 o = offset_p(p, n, t - p);
    V3 d = t - o; float dist = norm(d); d = d / dist;
    return { o, d, dist * (1.f - ShadowEpsilon) };
}
// Mesh::compute_surface_interaction (src/render/mesh.cpp:1392-1565, no texcoords) +
// SurfaceInteraction::finalize_surface_interaction / initialize_sh_frame (interaction.h:278-289, 497-517)
static SI make_si(const Scene &s, const Ray &ray, const Hit &h) {
    SI si; if (!h.valid()) { si.wi = -ray.d; return si; }
    const Mesh *m; uint32_t f; V3 p0, p1, p2; prim_verts(s, h.prim, &m, &f, &p0, &p1, &p2);
    float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
    si.valid = true; si.t = h.t; si.mesh = m; si.prim = h.prim;
    si.p = fmadd(p0, b0, fmadd(p1, b1, p2 * b2));
    si.n = normalize(cross(p1 - p0, p2 - p0));                  // mesh.h:156-164 face_normal
    V3 dp_du, dp_dv; coordinate_system(si.n, &dp_du, &dp_dv);
    if (!m->nrm.empty()) {
        V3 n0 = m->vn(m->faces[3 * f]), n1 = m->vn(m->faces[3 * f + 1]), n2 = m->vn(m->faces[3 * f + 2]);
        V3 n = fmadd(n2, b2, fmadd(n1, b1, n0 * b0));
        float il = rsqrt(squared_norm(n));
        si.sh_n = n * il;
    } else si.sh_n = si.n;
    if (m->flip_normals) { si.n = -si.n; si.sh_n = -si.sh_n; }
    // Gram-Schmidt (interaction.h:278-289)
    si.sh_s = normalize(fmadd(si.sh_n, -dot(si.sh_n, dp_du), dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { V3 a, b; coordinate_system(si.sh_n, &a, &b); si.sh_s = a; }
    si.sh_t = cross(si.sh_n, si.sh_s);
    si.wi = si.to_local(-ray.d);
    return si;
}

// ---------------------------------------------------------------- emitters
// include/mitsuba/core/distr_1d.h:116-134 (JIT predicate) + drjit/util.h:134-185 binary_search
static inline uint32_t distr_sample(const Mesh &m, float sample) {
    sample *= m.sum;
    uint32_t start = 0, end = m.nfaces() - 1;
    uint32_t iterations = 0;
    if (start < end) { uint32_t d = end - start, l = 0; while (d >>= 1) ++l; iterations = l + 1; }   // log2i(end-start)+1
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        float value = m.cdf[middle];
        bool cond = ((value < sample) || value == 0.f) && value != m.sum;
        if (cond) start = std::min(middle + 1, end); else end = middle;
    }
    return start;
}
struct DirectionSample { V3 p, n, d; float pdf = 0, dist = 0; bool valid = false, delta = false; };
// Mesh::sample_position (mesh.cpp:764-818) + Shape::sample_direction (shape.cpp:360-377)
// + AreaLight::sample_direction (area.cpp:116-167) + Scene::sample_emitter_direction (scene.cpp:295-346)
static DirectionSample sample_emitter_direction(const Scene &s, const SI &ref, V2 sample, V3 *spec_out, bool brute, TraverseStats *st) {
    DirectionSample ds; *spec_out = V3(0.f);
    size_t ne = s.emitters.size(); if (ne == 0) return ds;
    // scene.cpp:223-246 sample_emitter (uniform)
    uint32_t index = 0; float emitter_weight = 1.f;
    if (ne >= 2) {
        float ecf = (float) (uint32_t) ne, scaled = sample.x * ecf;
        index = std::min((uint32_t) scaled, (uint32_t) ne - 1u);
        emitter_weight = ecf; sample.x = scaled - (float) index;
    }
    const Emitter &em = s.emitters[index];
    if (em.type != 0) {
        V3 spec(0.f);
        if (em.type == 1) {                     // PointLight::sample_direction (point.cpp:117-148)
            ds.p = em.position; ds.pdf = 1.f; ds.delta = true; ds.d = ds.p - ref.p;
            const float dist2 = squared_norm(ds.d), inv_dist = rsqrt(dist2);
            ds.dist = std::sqrt(dist2); ds.d = ds.d * inv_dist;
            spec = em.radiance * sqr(inv_dist);
        } else if (em.type == 2) {              // ConstantBackgroundEmitter::sample_direction (constant.cpp:120-150)
            const V3 d = square_to_uniform_sphere(sample);
            const float radius = fmax_(s.bs_radius, norm(ref.p - s.bs_center)), dist = 2.f * radius;
            ds.p = fmadd(d, dist, ref.p); ds.n = -d; ds.pdf = InvFourPi; ds.d = d; ds.dist = dist;
            spec = em.radiance / ds.pdf;
        } else {                                // EnvironmentMapEmitter::sample_direction (envmap.cpp:419-462)
            float pdf; V2 uv = s.env.warp.sample(sample, &pdf);
            uv.x += .5f / (float) (s.env.res_x - 1u);
            const bool act = pdf > 0.f;
            const float theta = uv.y * Pi, phi = uv.x * (2.f * Pi);
            float st_, ct_, sp_, cp_; sincos(theta, &st_, &ct_); sincos(phi, &sp_, &cp_);
            V3 d(cp_ * st_, sp_ * st_, ct_); d = V3(d.y, d.z, -d.x);
            const float radius = fmax_(s.bs_radius, norm(ref.p - s.bs_center)), dist = 2.f * radius;
            const float inv_sin_theta = rsqrt(fmax_(fmax_(sqr(d.x) + sqr(d.z), sqr(Epsilon)), 0.f));
            d = xf_vec3(s.env.to_world, d);
            ds.p = ref.p + d * dist; ds.n = -d; ds.d = d; ds.dist = dist;
            ds.pdf = act ? pdf * inv_sin_theta * (1.f / (2.f * (Pi * Pi))) : 0.f;
            spec = mask3(s.env.eval_uv(uv) / ds.pdf, act);
        }
        ds.pdf *= 1.f / (float) ne;
        spec = spec * emitter_weight;
        if (ds.pdf != 0.f) {
            bool occluded = scene_ray_test(s, spawn_ray_to(ref.p, ref.n, ds.p), brute, st);
            if (occluded) { spec = V3(0.f); ds.pdf = 0.f; }
        }
        ds.valid = true; *spec_out = spec;
        return ds;
    }
    const Mesh &m = s.meshes[em.mesh];
    // Mesh::sample_position: face by area on sample.y (sample_reuse, distr_1d.h:173-182)
    uint32_t face = distr_sample(m, sample.y);
    float pmf = m.pmf[face] * m.normalization, cdf = face > 0 ? m.cdf[face - 1] * m.normalization : 0.f;
    sample.y = (sample.y - cdf) / pmf;
    V3 p0 = m.vp(m.faces[3 * face]), p1 = m.vp(m.faces[3 * face + 1]), p2 = m.vp(m.faces[3 * face + 2]);
    V3 e0 = p1 - p0, e1 = p2 - p0;
    V2 b = square_to_uniform_triangle(sample);
    ds.p = fmadd(e0, b.x, fmadd(e1, b.y, p0));
    ds.pdf = m.normalization;
    if (!m.nrm.empty()) {
        V3 n0 = m.vn(m.faces[3 * face]), n1 = m.vn(m.faces[3 * face + 1]), n2 = m.vn(m.faces[3 * face + 2]);
        ds.n = fmadd(n0, 1.f - b.x - b.y, fmadd(n1, b.x, n2 * b.y));
    } else ds.n = cross(e0, e1);
    ds.n = normalize(ds.n);
    if (m.flip_normals) ds.n = -ds.n;
    // Shape::sample_direction
    ds.d = ds.p - ref.p;
    float dist_squared = squared_norm(ds.d);
    ds.dist = std::sqrt(dist_squared);
    ds.d = ds.d / ds.dist;
    float dp = fabs_(dot(ds.d, ds.n)), x = dist_squared / dp;
    ds.pdf *= std::isfinite(x) ? x : 0.f;
    // AreaLight::sample_direction
    bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
    V3 spec = em.radiance / ds.pdf;
    spec = mask3(spec, active);
    // Scene: pdf_emitter, emitter_weight, visibility
    ds.pdf *= (ne == 0 ? 0.f : 1.f / (float) ne);
    spec = spec * emitter_weight;
    active = ds.pdf != 0.f;           // scene.cpp:318 (active &= ds.pdf != 0 on the scene-level mask)
    if (active) {
        bool occluded = scene_ray_test(s, spawn_ray_to(ref.p, ref.n, ds.p), brute, st);
        if (occluded) { spec = V3(0.f); ds.pdf = 0.f; }
    }
    ds.valid = true; *spec_out = spec;
    return ds;
}
// Scene::pdf_emitter_direction (scene.cpp:351-360) + AreaLight::pdf_direction (area.cpp:169-195)
// + Shape::pdf_direction (shape.cpp:379-390) + Mesh::pdf_position (mesh.cpp:846-849)
static float pdf_emitter_direction(const Scene &s, const Mesh &em_mesh, V3 d, V3 n, float dist, bool active) {
    float dp = dot(d, n);
    active = active && dp < 0.f;
    float pdf = em_mesh.normalization, adp = fabs_(dp);
    pdf *= (adp != 0.f) ? (dist * dist) / adp : 0.f;
    float value = active ? pdf : 0.f;
    return value * (s.emitters.empty() ? 0.f : 1.f / (float) s.emitters.size());
}
// src/integrators/mvpath.h:306-311
static inline float mis_weight(float a, float b) { a *= a; b *= b; float w = a / (a + b); return std::isfinite(w) ? w : 0.f; }

// Emitter::eval / pdf_direction of the scene's environment emitter (constant.cpp:92-96, 152-157; envmap.cpp:308-320, 464-481) for an escaped
// ray of direction d (= -si.wi of the invalid interaction, interaction.h:514)
static inline V2 env_dir_to_uv(const Scene &s, V3 d, V3 *local) {
    const V3 v = xf_vec3(s.env.from_world, d); *local = v;
    return { atan2_(v.x, -v.z) * InvTwoPi, safe_acos(v.y) * InvPi };
}
static inline V3 env_eval(const Scene &s, V3 d) {
    const Emitter &em = s.emitters[(size_t) s.env_emitter];
    if (em.type == 2) return em.radiance;
    V3 l; return s.env.eval_uv(env_dir_to_uv(s, d, &l));
}
static inline float env_pdf_direction(const Scene &s, V3 d) {
    if (s.emitters[(size_t) s.env_emitter].type == 2) return InvFourPi;
    V3 v; V2 uv = env_dir_to_uv(s, d, &v);
    uv.x -= .5f / (float) (s.env.res_x - 1u);
    uv.x -= std::floor(uv.x); uv.y -= std::floor(uv.y);
    const float inv_sin_theta = rsqrt(fmax_(fmax_(sqr(v.x) + sqr(v.z), sqr(Epsilon)), 0.f));
    return s.env.warp.eval(uv) * inv_sin_theta * (1.f / (2.f * (Pi * Pi)));
}

// ---------------------------------------------------------------- sensors
// GridSensor::sample_ray (src/sensors/grid.cpp:238-267) -> PerspectiveCamera::sample_ray
// (src/sensors/perspective.cpp:205-241) / ThinLensCamera::sample_ray (src/sensors/thinlens.cpp:220-257)
static inline V3 xf_point(const M4 &m, V3 p) {       // transform.h:134-144 (perspective divide via rcp)
    float r[4];
    for (int i = 0; i < 4; ++i) r[i] = m.m[i][3];
    for (int i = 0; i < 4; ++i) r[i] = fmadd(m.m[i][0], p.x, r[i]);
    for (int i = 0; i < 4; ++i) r[i] = fmadd(m.m[i][1], p.y, r[i]);
    for (int i = 0; i < 4; ++i) r[i] = fmadd(m.m[i][2], p.z, r[i]);
    float iw = rcp(r[3]);
    return V3(r[0] * iw, r[1] * iw, r[2] * iw);
}
static inline V3 xf_point_affine(const M4 &m, V3 p) { // transform.h:121-129
    float r[3];
    for (int i = 0; i < 3; ++i) r[i] = fmadd(m.m[i][2], p.z, fmadd(m.m[i][1], p.y, fmadd(m.m[i][0], p.x, m.m[i][3])));
    return V3(r[0], r[1], r[2]);
}
static inline V3 xf_vector(const M4 &m, V3 v) {       // transform.h:150-160
    float r[3];
    for (int i = 0; i < 3; ++i) r[i] = fmadd(m.m[i][2], v.z, fmadd(m.m[i][1], v.y, m.m[i][0] * v.x));
    return V3(r[0], r[1], r[2]);
}
static Ray sensor_sample_ray(const Scene &s, V2 pos, V2 aperture, uint32_t *view_out) {
    float fx = pos.x * (float) s.grid_x, fy = pos.y * (float) s.grid_y;
    uint32_t ux = (uint32_t) fx, uy = (uint32_t) fy, ix = ux, iy = uy;
    if (s.reverse_x) ix = (s.grid_x - 1) - ix;
    if (s.reverse_y) iy = (s.grid_y - 1) - iy;
    uint32_t index = ix + s.grid_x * iy;
    index = std::min(index, (uint32_t) s.sensors.size() - 1u);
    V2 p2 = { fx - (float) ux, fy - (float) uy };
    const SubSensor &c = s.sensors[index];
    if (view_out) *view_out = index;
    V3 near_p = xf_point(c.s2c, V3(p2.x + 0.f, p2.y + 0.f, 0.f));
    Ray ray;
    if (!s.thinlens) {
        V3 d = normalize(near_p);
        ray.o = V3(c.to_world.m[0][3], c.to_world.m[1][3], c.to_world.m[2][3]);
        ray.d = xf_vector(c.to_world, d);
        float inv_z = rcp(d.z), near_t = c.near_clip * inv_z, far_t = c.far_clip * inv_z;
        ray.o = ray.o + ray.d * near_t; ray.maxt = far_t - near_t;
    } else {
        V2 tmp = square_to_uniform_disk_concentric(aperture);
        V3 ap(c.aperture_radius * tmp.x, c.aperture_radius * tmp.y, 0.f);
        V3 focus_p = near_p * (c.focus_distance / near_p.z);
        V3 d = normalize(focus_p - ap);
        ray.o = xf_point_affine(c.to_world, ap);
        ray.d = xf_vector(c.to_world, d);
        float inv_z = rcp(d.z), near_t = c.near_clip * inv_z, far_t = c.far_clip * inv_z;
        ray.o = ray.o + ray.d * near_t; ray.maxt = far_t - near_t;
    }
    return ray;
}

// ---------------------------------------------------------------- integrator
struct Tap {            // optional per-lane records for the parity tests
    uint32_t max_hits = 0; uint32_t *hit_prim = nullptr; float *hit_t = nullptr;   // [max_hits] per lane, 0xffffffff = miss / not reached
    uint32_t n_iter = 0;
};
struct PathResult { V3 L; bool valid; };
// MVPT::sample_single (src/integrators/mvpath_single.h:82-278), JIT semantics
static PathResult sample_single(const Scene &s, Pcg32 &rng, Ray ray, bool brute, Tap *tap, TraverseStats *st) {
    if (s.max_depth == 0) return { V3(0.f), false };
    V3 throughput(1.f), result(0.f); float eta = 1.f; uint32_t depth = 0;
    bool valid_ray = !s.hide_emitters && s.env_emitter >= 0;      // :98
    V3 prev_p(0.f); float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true, active = true;
    while (active) {
        Hit h = scene_intersect(s, ray, brute, st);
        if (tap && tap->n_iter < tap->max_hits) { tap->hit_prim[tap->n_iter] = h.prim; tap->hit_t[tap->n_iter] = h.t; }
        if (tap) tap->n_iter++;
        SI si = make_si(s, ray, h);
        // direct emission (:148-163)
        if (si.valid && si.mesh->emitter >= 0) {
            V3 rel = si.p - prev_p; float dist = norm(rel); V3 d = rel / dist;
            float em_pdf = 0.f;
            if (!prev_bsdf_delta) em_pdf = pdf_emitter_direction(s, *si.mesh, d, si.sh_n, dist, true);
            float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
            V3 le = mask3(s.emitters[si.mesh->emitter].radiance, si.wi.z > 0.f && prev_bsdf_pdf > 0.f);   // area.cpp:82-86, vcall mask :162
            result = fmadd(throughput, le * mis_bsdf, result);
        } else if (!si.valid && s.env_emitter >= 0) {      // si.emitter(scene) = the environment (scene.h:647-657); ds.d = -si.wi (records.h:172-179)
            float em_pdf = 0.f;
            if (!prev_bsdf_delta) em_pdf = env_pdf_direction(s, ray.d) * (1.f / (float) s.emitters.size());
            float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
            V3 le = mask3(env_eval(s, ray.d), prev_bsdf_pdf > 0.f);
            result = fmadd(throughput, le * mis_bsdf, result);
        }
        bool active_next = (depth + 1 < s.max_depth) && si.valid;
        if (!active_next) {              // remaining statements cannot change `result`/`valid_ray` except :252
            valid_ray = valid_ray || (active && si.valid);   // JIT: :252 executes with bsdf_sample from a valid BSDF (never Null here)
            // the six RNG draws of this iteration are consumed by a lane that is about to die: unobservable
            break;
        }
        const Bsdf &bsdf = s.bsdfs[si.mesh->bsdf];
        // emitter sampling (:179-201)
        bool active_em = (bsdf.flags() & L_SMOOTH) != 0;
        V2 es = { 0, 0 }; es.x = rng.next_f32(); es.y = rng.next_f32();
        DirectionSample ds; V3 em_weight(0.f), wo(0.f);
        if (active_em) {
            ds = sample_emitter_direction(s, si, es, &em_weight, brute, st);
            active_em = active_em && ds.pdf != 0.f;
            wo = si.to_local(ds.d);
        }
        float s1 = rng.next_f32(); V2 s2; s2.x = rng.next_f32(); s2.y = rng.next_f32();
        BsdfResult br = bsdf_eval_pdf_sample(bsdf, si.wi, wo, s1, s2);
        if (active_em) {
            float mis_em = ds.delta ? 1.f : mis_weight(ds.pdf, br.pdf);       // :217
            result = fmadd(throughput, br.f * em_weight * mis_em, result);
        }
        ray = spawn_ray(si.p, si.n, si.to_world(br.bs.wo));
        throughput = throughput * br.weight;
        eta *= br.bs.eta;
        valid_ray = valid_ray || true;   // active && si.valid && !Null
        prev_p = si.p; prev_bsdf_pdf = br.bs.pdf; prev_bsdf_delta = (br.bs.type & L_DELTA) != 0;
        depth += 1;
        float tmax = max3(throughput);
        float rr_prob = fmin_(tmax * sqr(eta), .95f);
        bool rr_active = depth >= s.rr_depth, rr_continue = rng.next_f32() < rr_prob;
        if (rr_active) throughput = throughput * rcp(rr_prob);
        active = active_next && (!rr_active || rr_continue) && (tmax != 0.f);
    }
    return { valid_ray ? result : V3(0.f), valid_ray };
}

#include "o_reuse.h"

struct LaneOut { float sample_pos[2]; Ray ray; uint32_t view; PathResult r; uint32_t pixel; bool in_film; };
// MVPT::render JIT branch lane mapping (src/integrators/mvpath.cpp:174-190) + render_sample
// (src/integrators/mvpath_single.h:50-80)
static LaneOut render_lane(const Scene &s, uint32_t seed_value, uint32_t lane, uint32_t spp_per_pass, bool brute, Tap *tap, TraverseStats *st, bool trace = true) {
    LaneOut o;
    Pcg32 rng = sampler_seed(seed_value, lane);
    uint32_t idx = lane / spp_per_pass;
    uint32_t py = idx / s.film_w, px = idx - s.film_w * py;
    float sx = 1.f / (float) s.film_w, sy = 1.f / (float) s.film_h;
    float jx = rng.next_f32(), jy = rng.next_f32();
    float spx = (float) (int32_t) px + jx, spy = (float) (int32_t) py + jy;
    V2 adjusted = { fmadd(spx, sx, -0.f * sx), fmadd(spy, sy, -0.f * sy) };
    V2 aperture = { .5f, .5f };
    if (s.thinlens) { aperture.x = rng.next_f32(); aperture.y = rng.next_f32(); }
    if (s.shutter_draw) (void) rng.next_f32();       // mvpath_single.h:66-68: time += next_1d() * shutter_open_time
    o.sample_pos[0] = spx; o.sample_pos[1] = spy;
    o.ray = sensor_sample_ray(s, adjusted, aperture, &o.view);
    if (trace) o.r = sample_single(s, rng, o.ray, brute, tap, st); else o.r = { V3(0.f), false };
    // ImageBlock::put box-filter path (src/render/imageblock.cpp:210-233)
    int32_t fx = (int32_t) std::floor(spx), fy = (int32_t) std::floor(spy);
    uint32_t ux = (uint32_t) fx, uy = (uint32_t) fy;
    o.in_film = ux < s.film_w && uy < s.film_h;
    o.pixel = uy * s.film_w + ux;
    return o;
}

// ReconstructionFilter::eval: GaussianFilter (gaussian.cpp:94-103, the non-CUDA branch: a degree-9 polynomial in x^2 fitted
// to exp(-x^2 / (2 stddev^2)), shifted so that it reaches zero at the radius, :52-88), TentFilter (tent.cpp:54-56),
// MitchellNetravaliFilter (mitchell.cpp:50-78; param = B, param2 = C), CatmullRomFilter (catmullrom.cpp:34-53) and
// LanczosSincFilter (lanczos.cpp:52-67; param = lobes)
static void filter_setup(Scene &s, int type, float param, float param2) {
    s.filter_type = type;
    if (type == 1) {
        static const double coeff[10] = { 9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                                          -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11 };
        float stddev = param; s.filter_radius = 4 * stddev;
        double scale = 1;
        for (int i = 0; i < 10; ++i) { s.filter_coeff[i] = (float) (coeff[i] * scale); scale /= (double) stddev * (double) stddev; }
        s.filter_coeff[0] -= estrin(s.filter_radius * s.filter_radius, s.filter_coeff, 10);
    } else if (type == 2) { s.filter_radius = param; s.filter_inv_radius = 1.f / param; }
    else if (type == 3) {
        float B = param, C = param2; s.filter_radius = 2.f;
        float c[7] = { 12.f - 9.f * B - 6.f * C, -18.f + 12.f * B + 6.f * C, 6.f - 2.f * B,
                       -B - 6.f * C, 6.f * B + 30.f * C, -12.f * B - 48.f * C, 8.f * B + 24.f * C };
        for (int i = 0; i < 7; ++i) s.filter_coeff[i] = c[i];
    } else if (type == 4) s.filter_radius = 2.f;
    else if (type == 5) s.filter_radius = param;
    else s.filter_radius = 0.5f;
}
static inline float filter_eval(const Scene &s, float x) {
    if (s.filter_type == 1) return std::max(estrin(x * x, s.filter_coeff, 10), 0.f);
    if (s.filter_type == 2) return std::max(0.f, 1.f - std::fabs(x * s.filter_inv_radius));
    x = std::fabs(x);
    if (s.filter_type == 5) {
        float x1 = Pi * x, x2 = x1 / s.filter_radius, s1, s2, c;
        sincos(x1, &s1, &c); sincos(x2, &s2, &c);
        float result = (s1 * s2) / (x1 * x2);
        return x < Epsilon ? 1.f : (x > s.filter_radius ? 0.f : result);
    }
    float x2 = x * x, x3 = x2 * x, r;
    const float *k = s.filter_coeff;
    if (s.filter_type == 3) r = x < 1.f ? fmadd(k[0], x3, fmadd(k[1], x2, k[2])) : fmadd(k[3], x3, fmadd(k[4], x2, fmadd(k[5], x, k[6])));
    else r = x < 1.f ? (9.f * x3 + -15.f * x2) + 6.f : ((-3.f * x3 + 15.f * x2) + -24.f * x) + 12.f;
    r = (1.f / 6.f) * r;
    return x < 2.f ? r : 0.f;
}
// ImageBlock::put, coalesced accumulation (src/render/imageblock.cpp:433-505) into a block without border (film->create_block()
// defaults, hdrfilm.cpp:276-291), normalize = false: the footprint (2n+1)^2, n = ceil(radius - 1/2), around floor(pos) - n
static void filter_put(const Scene &s, float *film, float spx, float spy, V3 L) {
    uint32_t n = (uint32_t) (int32_t) std::ceil(s.filter_radius - .5f), count = 2 * n + 1;
    int32_t pix = (int32_t) std::floor(spx) - (int32_t) n, piy = (int32_t) std::floor(spy) - (int32_t) n;
    uint32_t x = (uint32_t) pix, y = (uint32_t) piy;
    float rx = ((float) pix + .5f) - spx, ry = ((float) piy + .5f) - spy;
    float wx[16], wy[16];
    for (uint32_t i = 0; i < count; ++i) { wx[i] = filter_eval(s, rx); wy[i] = filter_eval(s, ry); rx += 1.f; ry += 1.f; }
    float values[4] = { L.x, L.y, L.z, 1.f };
    for (uint32_t ys = 0; ys < count; ++ys, ++y) {
        if (!(y < s.film_h)) continue;
        for (uint32_t xs = 0; xs < count; ++xs) {
            uint32_t xx = x + xs; if (!(xx < s.film_w)) continue;
            float weight = wy[ys] * wx[xs];
            float *px = film + 4 * ((size_t) y * s.film_w + xx);
            for (int k = 0; k < 4; ++k) px[k] += values[k] * weight;
        }
    }
}

} // namespace orc

// ====================================================================== C interface (ctypes)
using namespace orc;
namespace orc {
// PerspectiveCamera::update_camera_transforms (perspective.cpp:173-199): camera_to_sample = inverse of sample_to_camera,
// image rectangle of the near plane -> normalization; Transform4f::inverse of to_world for sample_surface
static void sensor_finish(SubSensor &c) {
    c.c2s = m4_inverse(c.s2c); c.w2c = m4_inverse(c.to_world);
    V3 pmin = xf_point(c.s2c, V3(0.f, 0.f, 0.f)), pmax = xf_point(c.s2c, V3(1.f, 1.f, 0.f));
    V3 a = pmin / pmin.z, b = pmax / pmax.z;
    float lox = fmin_(a.x, b.x), hix = fmax_(a.x, b.x), loy = fmin_(a.y, b.y), hiy = fmax_(a.y, b.y);
    c.normalization = 1.f / ((hix - lox) * (hiy - loy));
}
} // namespace orc
using namespace orc;

extern "C" {

struct OrcBsdfDesc {        // mirrors the XML parameters of the reference BSDF plugins
    int32_t type, twosided, distr, nonlinear;
    float reflectance[3], eta[3], k[3], spec_refl[3], spec_trans[3];
    float int_ior, ext_ior, alpha_u, alpha_v;
};
struct OrcSensorDesc {      // grid sensor over perspective / thinlens sub-cameras
    int32_t thinlens, fov_axis; uint32_t grid_x, grid_y, res_x, res_y; int32_t reverse_x, reverse_y;
    int32_t use_cone, cam_center;
    double to_world[16], to_world_inv_t[16], fov;
    float near_clip, far_clip, focus_distance, aperture_radius, cone_deg, cam_dist;
    float cam_dir[3], cam_off[3];
};

void *orc_scene_new() { return new Scene(); }
void orc_scene_free(void *s) { delete (Scene *) s; }
int orc_add_bsdf(void *sp, const OrcBsdfDesc *d) {
    Scene &s = *(Scene *) sp; Bsdf b;
    b.type = d->type; b.twosided = d->twosided; b.distr = d->distr; b.nonlinear = d->nonlinear;
    b.reflectance = V3(d->reflectance[0], d->reflectance[1], d->reflectance[2]);
    b.eta = V3(d->eta[0], d->eta[1], d->eta[2]); b.k = V3(d->k[0], d->k[1], d->k[2]);
    b.spec_refl = V3(d->spec_refl[0], d->spec_refl[1], d->spec_refl[2]);
    b.spec_trans = V3(d->spec_trans[0], d->spec_trans[1], d->spec_trans[2]);
    b.ior_eta = d->int_ior / d->ext_ior; b.alpha_u = d->alpha_u; b.alpha_v = d->alpha_v;
    b.prepare();
    s.bsdfs.push_back(b); return (int) s.bsdfs.size() - 1;
}
int orc_add_mesh(void *sp, const float *pos, uint32_t nv, const uint32_t *faces, uint32_t nf, const float *normals,
                 int bsdf, const float *radiance, int flip_normals) {
    Scene &s = *(Scene *) sp; Mesh m;
    m.pos.assign(pos, pos + 3 * (size_t) nv); m.faces.assign(faces, faces + 3 * (size_t) nf);
    if (normals) m.nrm.assign(normals, normals + 3 * (size_t) nv);
    m.bsdf = bsdf; m.flip_normals = flip_normals;
    m.prim_offset = s.nprims();
    int id = (int) s.meshes.size();
    if (radiance) { m.emitter = (int) s.emitters.size(); s.emitters.push_back({ id, V3(radiance[0], radiance[1], radiance[2]) }); }
    // area pmf/cdf: table[i] = .5f * norm(cross(p1-p0, p2-p0)) (mesh.cpp:466-474); cdf accumulated in double (distr_1d.h:240-247)
    m.pmf.resize(nf); m.cdf.resize(nf); double sum = 0.0;
    for (uint32_t i = 0; i < nf; ++i) {
        V3 p0 = m.vp(m.faces[3 * i]), p1 = m.vp(m.faces[3 * i + 1]), p2 = m.vp(m.faces[3 * i + 2]);
        m.pmf[i] = .5f * norm(cross(p1 - p0, p2 - p0)); sum += (double) m.pmf[i]; m.cdf[i] = (float) sum;
    }
    m.sum = nf ? m.cdf[nf - 1] : 0.f; m.normalization = rcp(m.sum);
    for (uint32_t i = 0; i < nf; ++i) s.prim_mesh.push_back((uint32_t) id);
    s.meshes.push_back(std::move(m)); return id;
}
static M4 m4_from_d(const double *d) { M4 r; for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = (float) d[4 * i + j]; return r; }
// GridSensor ctor (grid.cpp:84-236) + PerspectiveCamera ctor (perspective.cpp:139-154, 173-181)
void orc_set_sensor_grid(void *sp, const OrcSensorDesc *d) {
    Scene &s = *(Scene *) sp;
    s.grid_x = d->grid_x; s.grid_y = d->grid_y; s.reverse_x = d->reverse_x; s.reverse_y = d->reverse_y; s.thinlens = d->thinlens;
    uint32_t sub_w = d->res_x, sub_h = d->res_y;
    s.film_w = sub_w * d->grid_x; s.film_h = sub_h * d->grid_y;
    double sub_asp = sub_w / (double) sub_h;
    double fov_x = parse_fov(d->fov, d->fov_axis, sub_asp);
    GridParams g; g.grid_x = d->grid_x; g.grid_y = d->grid_y; g.use_cone = d->use_cone; g.cone_deg = d->cone_deg;
    for (int i = 0; i < 3; ++i) { g.cam_dir[i] = d->cam_dir[i]; g.cam_off[i] = d->cam_off[i]; }
    g.cam_dist = d->cam_dist; g.cam_center = d->cam_center; g.foc_dist = d->focus_distance;
    M4 inv_t = m4_from_d(d->to_world_inv_t);
    s.sensors.clear();
    for (uint32_t i = 0; i < d->grid_x * d->grid_y; ++i) {
        SubSensor c; float lens_shift;
        grid_sub_camera(g, inv_t, fov_x, i, &c.to_world, &lens_shift);
        c.s2c = sample_to_camera((int) sub_w, (int) sub_h, (float) fov_x, d->near_clip, d->far_clip, lens_shift);
        c.near_clip = d->near_clip; c.far_clip = d->far_clip; c.aperture_radius = d->aperture_radius; c.focus_distance = d->focus_distance;
        sensor_finish(c); s.sensors.push_back(c);
    }
}
// BatchSensor-style explicit sub-cameras (src/sensors/batch.cpp:94-130): n views side by side
void orc_set_sensor_explicit(void *sp, uint32_t n, uint32_t res_x, uint32_t res_y, const float *s2c, const float *to_world,
                             float near_clip, float far_clip) {
    Scene &s = *(Scene *) sp; s.grid_x = n; s.grid_y = 1; s.reverse_x = 0; s.reverse_y = 0; s.thinlens = 0;
    s.film_w = res_x * n; s.film_h = res_y; s.sensors.clear();
    for (uint32_t i = 0; i < n; ++i) {
        SubSensor c; std::memcpy(c.s2c.m, s2c + 16 * i, 64); std::memcpy(c.to_world.m, to_world + 16 * i, 64);
        c.near_clip = near_clip; c.far_clip = far_clip; c.aperture_radius = 0; c.focus_distance = far_clip; sensor_finish(c); s.sensors.push_back(c);
    }
}
// the sub-sensor table as a scene loader produced it (for scenes read from XML by the product): n = grid_x * grid_y records
void orc_set_sensor_raw(void *sp, uint32_t grid_x, uint32_t grid_y, uint32_t view_w, uint32_t view_h, int reverse_x, int reverse_y, int thinlens,
                        const float *s2c, const float *to_world, const float *scalars4) {
    Scene &s = *(Scene *) sp; s.grid_x = grid_x; s.grid_y = grid_y; s.reverse_x = reverse_x; s.reverse_y = reverse_y; s.thinlens = thinlens;
    s.film_w = view_w * grid_x; s.film_h = view_h * grid_y; s.sensors.clear();
    for (uint32_t i = 0; i < grid_x * grid_y; ++i) {
        SubSensor c; std::memcpy(c.s2c.m, s2c + 16 * i, 64); std::memcpy(c.to_world.m, to_world + 16 * i, 64);
        c.near_clip = scalars4[4 * i]; c.far_clip = scalars4[4 * i + 1]; c.aperture_radius = scalars4[4 * i + 2]; c.focus_distance = scalars4[4 * i + 3];
        sensor_finish(c); s.sensors.push_back(c);
    }
}
void orc_get_sensor(void *sp, uint32_t i, float *s2c, float *to_world) {
    Scene &s = *(Scene *) sp; std::memcpy(s2c, s.sensors[i].s2c.m, 64); std::memcpy(to_world, s.sensors[i].to_world.m, 64);
}
void orc_set_integrator(void *sp, int64_t max_depth, uint32_t rr_depth, int hide_emitters, uint32_t sampler_seed) {
    Scene &s = *(Scene *) sp;
    s.max_depth = max_depth < 0 ? 0xffffffffu : (uint32_t) max_depth;        // integrator.cpp:503-520
    s.rr_depth = rr_depth; s.hide_emitters = hide_emitters; s.base_seed = sampler_seed;
}
void orc_set_reuse_adaptive(void *sp, uint32_t adaptive, int debug) { Scene &s = *(Scene *) sp; s.adaptive = adaptive; s.debug = debug; }
void orc_set_shutter_draw(void *sp, int draw) { ((Scene *) sp)->shutter_draw = draw; }
// scene-level emitters (src/emitters/{point,constant,envmap}.cpp); the environment map is rebuilt here from the raw RGB pixels
int orc_add_emitter_point(void *sp, const float *pos, const float *intensity) {
    Scene &s = *(Scene *) sp; Emitter e; e.mesh = -1; e.type = 1; e.position = V3(pos[0], pos[1], pos[2]); e.radiance = V3(intensity[0], intensity[1], intensity[2]); s.emitters.push_back(e); return (int) s.emitters.size() - 1;
}
int orc_add_emitter_constant(void *sp, const float *radiance) {
    Scene &s = *(Scene *) sp; Emitter e; e.mesh = -1; e.type = 2; e.radiance = V3(radiance[0], radiance[1], radiance[2]); s.emitters.push_back(e); s.env_emitter = (int) s.emitters.size() - 1; return s.env_emitter;
}
int orc_add_emitter_envmap(void *sp, const float *rgb, uint32_t w, uint32_t h, float scale, const float *to_world16) {
    Scene &s = *(Scene *) sp; M4 tw; std::memcpy(tw.m, to_world16, 64); s.env.build(rgb, w, h, scale, tw);
    Emitter e; e.mesh = -1; e.type = 3; e.radiance = V3(scale); s.emitters.push_back(e); s.env_emitter = (int) s.emitters.size() - 1; return s.env_emitter;
}
// puts the emitter list into the order given by (type, mesh) pairs -- the order of the scene description the product loaded
void orc_order_emitters(void *sp, const int32_t *type, const int32_t *mesh, uint32_t n) {
    Scene &s = *(Scene *) sp; std::vector<Emitter> out; std::vector<bool> used(s.emitters.size(), false);
    for (uint32_t i = 0; i < n; ++i) for (size_t j = 0; j < s.emitters.size(); ++j)
        if (!used[j] && s.emitters[j].type == type[i] && (type[i] != 0 || s.emitters[j].mesh == mesh[i])) { used[j] = true; out.push_back(s.emitters[j]); break; }
    if (out.size() != s.emitters.size()) return;
    s.emitters = out; s.env_emitter = -1;
    for (size_t j = 0; j < s.emitters.size(); ++j) { if (s.emitters[j].type == 0) s.meshes[(size_t) s.emitters[j].mesh].emitter = (int) j; else if (s.emitters[j].type >= 2) s.env_emitter = (int) j; }
}
void orc_square_to_uniform_sphere(const float *u, int n, float *o) { for (int i = 0; i < n; ++i) { V3 v = square_to_uniform_sphere({ u[2 * i], u[2 * i + 1] }); o[3 * i] = v.x; o[3 * i + 1] = v.y; o[3 * i + 2] = v.z; } }
void orc_square_to_bilinear(const float *v4, const float *u, int n, float *out2, float *pdf, float *pdf_eval) {
    for (int i = 0; i < n; ++i) { const float *v = v4 + 4 * i; V2 p = square_to_bilinear(v[0], v[1], v[2], v[3], { u[2 * i], u[2 * i + 1] }, pdf + i);
        out2[2 * i] = p.x; out2[2 * i + 1] = p.y; pdf_eval[i] = square_to_bilinear_pdf(v[0], v[1], v[2], v[3], p); }
}
// DiscreteDistribution over an unnormalised pmf exactly as the area CDF of a mesh is built and sampled (orc_add_mesh, sample_emitter_direction)
void orc_discrete_sample_reuse(const float *pmf, uint32_t n_pmf, const float *u, int n, uint32_t *index, float *reused, float *pmf_out, float *sum_o, float *norm_o) {
    Mesh m; m.faces.assign(3 * (size_t) n_pmf, 0u); m.pmf.assign(pmf, pmf + n_pmf); m.cdf.resize(n_pmf); double sum = 0.0;
    for (uint32_t i = 0; i < n_pmf; ++i) { sum += (double) m.pmf[i]; m.cdf[i] = (float) sum; }
    m.sum = m.cdf[n_pmf - 1]; m.normalization = rcp(m.sum); *sum_o = m.sum; *norm_o = m.normalization;
    for (int i = 0; i < n; ++i) {
        const uint32_t face = distr_sample(m, u[i]);
        const float p = m.pmf[face] * m.normalization, c = face > 0 ? m.cdf[face - 1] * m.normalization : 0.f;
        index[i] = face; pmf_out[i] = p; reused[i] = (u[i] - c) / p;
    }
}
void orc_env_warp(void *sp, const float *u, int n, float *uv, float *pdf, float *eval_back) {      // taps: Hierarchical2D sample + eval
    Scene &s = *(Scene *) sp;
    for (int i = 0; i < n; ++i) { V2 r = s.env.warp.sample({ u[2 * i], u[2 * i + 1] }, pdf + i); uv[2 * i] = r.x; uv[2 * i + 1] = r.y; eval_back[i] = s.env.warp.eval(r); }
}
void orc_hier2d(const float *data, uint32_t sx, uint32_t sy, const float *u, int n, float *uv, float *pdf, float *eval_back) {
    Hier2D h; h.build(data, sx, sy);
    for (int i = 0; i < n; ++i) { V2 r = h.sample({ u[2 * i], u[2 * i + 1] }, pdf + i); uv[2 * i] = r.x; uv[2 * i + 1] = r.y; eval_back[i] = h.eval(r); }
}
// MVPathIntegrator sample reuse (mvpath.h:121-128) + the sample-size rule of MVPT::render (mvpath.cpp:17-22, 192-218).
// Returns the number of samples per primary sample (0 = reuse off).
uint32_t orc_set_reuse(void *sp, int sa_reuse, int sa_mis, int fast_mis, uint32_t reuse_count) {
    Scene &s = *(Scene *) sp; uint32_t n_sensors = (uint32_t) s.sensors.size();
    s.sa_mis = sa_mis; s.fast_mis = fast_mis; s.n_reuse = 0;
    bool reuse = sa_reuse && n_sensors > 1 && reuse_count != 1;
    if (!reuse) return 0;
    uint32_t n = std::min(reuse_count, n_sensors);
    if (n == 0 || n_sensors % n) {
        n = 0;
        for (uint32_t p = 8; p < n_sensors; ++p) if (n_sensors % p == 0) { n = p; break; }
        if (!n) { for (uint32_t p = 8; p > 1; --p) if (n_sensors % p == 0) { n = p; break; } n = n ? n : n_sensors; }
    }
    s.n_reuse = n; return n;
}
// film reconstruction filter: type 0 box, 1 gaussian (param = stddev), 2 tent (param = radius), 3 mitchell (param = B, param2 = C),
// 4 catmullrom, 5 lanczos (param = lobes)
void orc_set_rfilter(void *sp, int type, float param, float param2) { filter_setup(*(Scene *) sp, type, param, param2); }
void orc_filter_eval(int type, float param, float param2, const float *x, int n, float *out) { Scene s; filter_setup(s, type, param, param2); for (int i = 0; i < n; ++i) out[i] = filter_eval(s, x[i]); }
void orc_prepare(void *sp) {
    Scene &s = *(Scene *) sp; bvh_build(s);
    // Scene::bbox().bounding_sphere() (bbox.h:113-115, 343-346) enlarged as in Emitter::set_scene (constant.cpp:73-88)
    float lo[3] = { Inf, Inf, Inf }, hi[3] = { -Inf, -Inf, -Inf }; bool any = false;
    for (auto &m : s.meshes) for (size_t i = 0; i + 2 < m.pos.size(); i += 3) for (int k = 0; k < 3; ++k) { lo[k] = fmin_(lo[k], m.pos[i + k]); hi[k] = fmax_(hi[k], m.pos[i + k]); any = true; }
    if (!any) { s.bs_center = V3(0.f); s.bs_radius = RayEpsilon; return; }
    s.bs_center = V3((hi[0] + lo[0]) * .5f, (hi[1] + lo[1]) * .5f, (hi[2] + lo[2]) * .5f);
    s.bs_radius = fmax_(RayEpsilon, norm(s.bs_center - V3(hi[0], hi[1], hi[2])) * (1.f + RayEpsilon));
}
void orc_film_size(void *sp, uint32_t *w, uint32_t *h) { Scene &s = *(Scene *) sp; *w = s.film_w; *h = s.film_h; }
uint32_t orc_num_prims(void *sp) { return ((Scene *) sp)->nprims(); }
void orc_bvh_info(void *sp, uint32_t *n_nodes) { *n_nodes = (uint32_t) ((Scene *) sp)->nodes.size(); }

// ---- leaf taps
void orc_sampler_stream(uint32_t seed_value, uint32_t lane, int n, float *out) {
    Pcg32 r = sampler_seed(seed_value, lane); for (int i = 0; i < n; ++i) out[i] = r.next_f32();
}
void orc_tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t *out) { tea32(v0, v1, rounds, out, out + 1); }
void orc_pcg32_f32(uint64_t initstate, uint64_t initseq, int n, float *out) { Pcg32 r; r.seed(initstate, initseq); for (int i = 0; i < n; ++i) out[i] = r.next_f32(); }
void orc_pcg32_u32(uint64_t initstate, uint64_t initseq, int n, uint32_t *out) { Pcg32 r; r.seed(initstate, initseq); for (int i = 0; i < n; ++i) out[i] = r.next_u32(); }
void orc_math1(int op, const float *in, int n, float *out) {
    for (int i = 0; i < n; ++i) {
        float x = in[i], r = x, s, c;
        switch (op) {
            case 0: sincos(x, &s, &c); r = s; break; case 1: sincos(x, &s, &c); r = c; break;
            case 2: r = exp_(x); break; case 3: r = log_(x); break; case 4: r = erf_(x); break; case 5: r = erfinv_(x); break;
            case 6: r = exp2_(x); break; case 7: r = std::sqrt(x); break; case 8: r = rcp(x); break; case 9: r = rsqrt(x); break;
            case 10: r = safe_sqrt(x); break; case 11: r = tan_(x); break;
            case 12: r = acos_(x); break; case 13: r = atan2_(x, 0.3f); break; case 14: r = atan2_(0.3f, x); break; case 15: r = atan2_(x, -0.3f); break; case 16: r = safe_acos(x); break;
        }
        out[i] = r;
    }
}
static V3 ld3(const float *p) { return V3(p[0], p[1], p[2]); }
static void st3(float *p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
void orc_dot(const float *a, const float *b, int n, float *o) { for (int i = 0; i < n; ++i) o[i] = dot(ld3(a + 3 * i), ld3(b + 3 * i)); }
void orc_cross(const float *a, const float *b, int n, float *o) { for (int i = 0; i < n; ++i) st3(o + 3 * i, cross(ld3(a + 3 * i), ld3(b + 3 * i))); }
void orc_normalize(const float *a, int n, float *o) { for (int i = 0; i < n; ++i) st3(o + 3 * i, normalize(ld3(a + 3 * i))); }
void orc_coordinate_system(const float *nn, int n, float *s, float *t) { for (int i = 0; i < n; ++i) { V3 a, b; coordinate_system(ld3(nn + 3 * i), &a, &b); st3(s + 3 * i, a); st3(t + 3 * i, b); } }
void orc_square_to_uniform_disk_concentric(const float *u, int n, float *o) { for (int i = 0; i < n; ++i) { V2 p = square_to_uniform_disk_concentric({ u[2 * i], u[2 * i + 1] }); o[2 * i] = p.x; o[2 * i + 1] = p.y; } }
void orc_square_to_cosine_hemisphere(const float *u, int n, float *o, float *pdf) { for (int i = 0; i < n; ++i) { V3 v = square_to_cosine_hemisphere({ u[2 * i], u[2 * i + 1] }); st3(o + 3 * i, v); pdf[i] = square_to_cosine_hemisphere_pdf(v); } }
void orc_square_to_uniform_triangle(const float *u, int n, float *o) { for (int i = 0; i < n; ++i) { V2 p = square_to_uniform_triangle({ u[2 * i], u[2 * i + 1] }); o[2 * i] = p.x; o[2 * i + 1] = p.y; } }
void orc_fresnel(const float *ci, float eta, int n, float *o4) { for (int i = 0; i < n; ++i) fresnel(ci[i], eta, o4 + 4 * i, o4 + 4 * i + 1, o4 + 4 * i + 2, o4 + 4 * i + 3); }
void orc_fresnel_conductor(const float *ci, float eta, float k, int n, float *o) { for (int i = 0; i < n; ++i) o[i] = fresnel_conductor(ci[i], eta, k); }
void orc_reflect_m(const float *wi, const float *m, int n, float *o) { for (int i = 0; i < n; ++i) st3(o + 3 * i, reflect(ld3(wi + 3 * i), ld3(m + 3 * i))); }
void orc_refract_m(const float *wi, const float *m, const float *ct, const float *eti, int n, float *o) { for (int i = 0; i < n; ++i) st3(o + 3 * i, refract(ld3(wi + 3 * i), ld3(m + 3 * i), ct[i], eti[i])); }
void orc_mf_eval(int type, float au, float av, const float *m, int n, float *o) { Microfacet d((MfType) type, au, av); for (int i = 0; i < n; ++i) o[i] = d.eval(ld3(m + 3 * i)); }
void orc_mf_smith_g1(int type, float au, float av, const float *v, const float *m, int n, float *o) { Microfacet d((MfType) type, au, av); for (int i = 0; i < n; ++i) o[i] = d.smith_g1(ld3(v + 3 * i), ld3(m + 3 * i)); }
void orc_mf_pdf(int type, float au, float av, const float *wi, const float *m, int n, float *o) { Microfacet d((MfType) type, au, av); for (int i = 0; i < n; ++i) o[i] = d.pdf(ld3(wi + 3 * i), ld3(m + 3 * i)); }
void orc_mf_sample(int type, float au, float av, const float *wi, const float *u, int n, float *mo, float *pdf) { Microfacet d((MfType) type, au, av); for (int i = 0; i < n; ++i) st3(mo + 3 * i, d.sample(ld3(wi + 3 * i), { u[2 * i], u[2 * i + 1] }, pdf + i)); }
void orc_rough_tables(int type, float alpha, float eta, int res, float *ext_t, float *int_r) {
    Bsdf b; b.type = BSDF_ROUGHPLASTIC; b.distr = type; b.alpha_u = b.alpha_v = alpha; b.ior_eta = eta; b.prepare();
    for (int i = 0; i < res && i < MI_ROUGH_TRANSMITTANCE_RES; ++i) ext_t[i] = b.ext_transmittance[i];
    *int_r = b.internal_reflectance;
}
void orc_gauss_legendre(int n, float *nodes, float *weights) { std::vector<float> a, b; gauss_legendre(n, a, b); for (int i = 0; i < n; ++i) { nodes[i] = a[i]; weights[i] = b[i]; } }
void orc_mat_inverse(const float *m, float *o) { M4 a; std::memcpy(a.m, m, 64); M4 r = m4_inverse(a); std::memcpy(o, r.m, 64); }
void orc_mat_inverse_transpose(const float *m, float *o) { M4 a; std::memcpy(a.m, m, 64); M4 r = m4_inverse_transpose(a); std::memcpy(o, r.m, 64); }
void orc_mat_mul(const float *a_, const float *b_, float *o) { M4 a, b; std::memcpy(a.m, a_, 64); std::memcpy(b.m, b_, 64); M4 r = m4_mul(a, b); std::memcpy(o, r.m, 64); }
void orc_sample_to_camera(int fw, int fh, float x_fov, float near_, float far_, float lens_shift, float *c2s_o, float *s2c_o) {
    Xf c2s = perspective_projection(fw, fh, fw, fh, 0, 0, x_fov, near_, far_); c2s.mat.m[0][2] += lens_shift;
    std::memcpy(c2s_o, c2s.mat.m, 64); M4 r = sample_to_camera(fw, fh, x_fov, near_, far_, lens_shift); std::memcpy(s2c_o, r.m, 64);
}
void orc_grid_to_world(const float *inv_t, const float *offset, float *o) {
    M4 it; std::memcpy(it.m, inv_t, 64); it.m[3][0] += offset[0]; it.m[3][1] += offset[1]; it.m[3][2] += offset[2];
    M4 r = m4_inverse(m4_transpose(it)); std::memcpy(o, r.m, 64);
}
void orc_perspective_sample_ray(const float *s2c, const float *tw, float near_, float far_, const float *pos, int n, float *o, float *d, float *maxt) {
    Scene s; s.grid_x = s.grid_y = 1; s.reverse_x = s.reverse_y = 0; SubSensor c; std::memcpy(c.s2c.m, s2c, 64); std::memcpy(c.to_world.m, tw, 64);
    c.near_clip = near_; c.far_clip = far_; c.aperture_radius = 0; c.focus_distance = far_; s.sensors.push_back(c);
    for (int i = 0; i < n; ++i) {
        // bypass the grid split (pos in [0,1) maps to tile 0 with identical fractional part only if < 1): call the sub-sensor directly
        V3 near_p = xf_point(c.s2c, V3(pos[2 * i], pos[2 * i + 1], 0.f)); V3 dd = normalize(near_p);
        V3 oo(c.to_world.m[0][3], c.to_world.m[1][3], c.to_world.m[2][3]); V3 dw = xf_vector(c.to_world, dd);
        float inv_z = rcp(dd.z), nt = near_ * inv_z, ft = far_ * inv_z; oo = oo + dw * nt;
        st3(o + 3 * i, oo); st3(d + 3 * i, dw); maxt[i] = ft - nt;
    }
}
void orc_moeller_trumbore(const float *o, const float *d, const float *maxt, const float *p0, const float *p1, const float *p2, int n, float *tuv, int *hit) {
    for (int i = 0; i < n; ++i) { Ray r{ ld3(o + 3 * i), ld3(d + 3 * i), maxt[i] }; hit[i] = moeller_trumbore(r, ld3(p0 + 3 * i), ld3(p1 + 3 * i), ld3(p2 + 3 * i), maxt[i], tuv + 3 * i, tuv + 3 * i + 1, tuv + 3 * i + 2) ? 1 : 0; }
}
// BSDF::eval_pdf_sample on arrays (bsdf id of this scene): outputs f[3], pdf, wo[3], bs_pdf, bs_eta, bs_type, weight[3]
void orc_bsdf_eval_pdf_sample(void *sp, int bsdf, const float *wi, const float *wo, const float *s1, const float *s2, int n,
                              float *f, float *pdf, float *bs_wo, float *bs_pdf, float *bs_eta, uint32_t *bs_type, float *weight) {
    Scene &s = *(Scene *) sp; const Bsdf &b = s.bsdfs[bsdf];
    for (int i = 0; i < n; ++i) {
        BsdfResult r = bsdf_eval_pdf_sample(b, ld3(wi + 3 * i), ld3(wo + 3 * i), s1[i], { s2[2 * i], s2[2 * i + 1] });
        st3(f + 3 * i, r.f); pdf[i] = r.pdf; st3(bs_wo + 3 * i, r.bs.wo); bs_pdf[i] = r.bs.pdf; bs_eta[i] = r.bs.eta; bs_type[i] = r.bs.type; st3(weight + 3 * i, r.weight);
    }
}
void orc_bsdf_tables(void *sp, int bsdf, float *ext_t, float *scalars /* inv_eta_2, spec_sampling_weight, internal_reflectance */) {
    const Bsdf &b = ((Scene *) sp)->bsdfs[bsdf];
    std::memcpy(ext_t, b.ext_transmittance, sizeof(b.ext_transmittance)); scalars[0] = b.inv_eta_2; scalars[1] = b.spec_sampling_weight; scalars[2] = b.internal_reflectance;
}

// ---- geometry taps
void orc_intersect(void *sp, const float *o, const float *d, const float *maxt, int n, int brute, uint32_t *prim, float *tuv, uint64_t *stats) {
    Scene &s = *(Scene *) sp; TraverseStats st;
    for (int i = 0; i < n; ++i) {
        Ray r{ ld3(o + 3 * i), ld3(d + 3 * i), maxt[i] }; Hit h = scene_intersect(s, r, brute != 0, &st);
        prim[i] = h.prim; tuv[3 * i] = h.t; tuv[3 * i + 1] = h.u; tuv[3 * i + 2] = h.v;
    }
    if (stats) { stats[0] = st.nodes; stats[1] = st.tris; }
}
void orc_ray_test(void *sp, const float *o, const float *d, const float *maxt, int n, int brute, int *occluded) {
    Scene &s = *(Scene *) sp;
    for (int i = 0; i < n; ++i) { Ray r{ ld3(o + 3 * i), ld3(d + 3 * i), maxt[i] }; occluded[i] = scene_ray_test(s, r, brute != 0) ? 1 : 0; }
}

// ---- lane taps: everything observable about lanes [lane_begin, lane_begin + n) of one pass
// seed_value = sampler base seed + (spp_per_pass * pass + render seed) (mvpath.cpp:226-231, sampler.cpp:129)
void orc_trace_lanes(void *sp, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, int brute, int trace,
                     float *sample_pos /*2n*/, float *ray_o /*3n*/, float *ray_d /*3n*/, float *ray_maxt /*n*/, uint32_t *view /*n*/,
                     float *radiance /*4n: rgb + valid*/, uint32_t *pixel /*n, 0xffffffff if outside*/,
                     uint32_t max_hits, uint32_t *hit_prim /*n*max_hits*/, float *hit_t, uint64_t *stats) {
    Scene &s = *(Scene *) sp; TraverseStats st;
    for (uint32_t i = 0; i < n; ++i) {
        Tap tap; tap.max_hits = max_hits;
        if (max_hits) { tap.hit_prim = hit_prim + (size_t) i * max_hits; tap.hit_t = hit_t + (size_t) i * max_hits;
            for (uint32_t k = 0; k < max_hits; ++k) { tap.hit_prim[k] = 0xfffffffeu; tap.hit_t[k] = 0.f; } }
        LaneOut o = render_lane(s, seed_value, lane_begin + i, spp_per_pass, brute != 0, max_hits ? &tap : nullptr, &st, trace != 0);
        if (sample_pos) { sample_pos[2 * i] = o.sample_pos[0]; sample_pos[2 * i + 1] = o.sample_pos[1]; }
        if (ray_o) { st3(ray_o + 3 * i, o.ray.o); st3(ray_d + 3 * i, o.ray.d); ray_maxt[i] = o.ray.maxt; }
        if (view) view[i] = o.view;
        if (radiance) { radiance[4 * i] = o.r.L.x; radiance[4 * i + 1] = o.r.L.y; radiance[4 * i + 2] = o.r.L.z; radiance[4 * i + 3] = o.r.valid ? 1.f : 0.f; }
        if (pixel) pixel[i] = o.in_film ? o.pixel : 0xffffffffu;
    }
    if (stats) { stats[0] = st.nodes; stats[1] = st.tris; }
}

// ---- full render (MVPT::render JIT branch, mvpath.cpp:132-272): film = RGBW float32 [H][W][4]
// Returns seconds spent in the pass loop.  Rows of the quilt are distributed over `nthreads` threads; each pixel's
// spp_per_pass lanes are accumulated in lane order, passes in pass order (deterministic).
double orc_render_rows(void *sp, uint32_t seed, uint32_t spp, uint32_t spp_pass_lim, int brute, int nthreads, float *film,
                       uint32_t row_begin, uint32_t row_end, uint32_t row_stride, uint64_t *stats);
double orc_render(void *sp, uint32_t seed, uint32_t spp, uint32_t spp_pass_lim, int brute, int nthreads, float *film,
                  uint32_t row_begin, uint32_t row_end, uint64_t *stats) {
    return orc_render_rows(sp, seed, spp, spp_pass_lim, brute, nthreads, film, row_begin, row_end, 1, stats);
}
// Same, restricted to rows row_begin, row_begin + row_stride, ... < row_end (bounded samples of a large quilt for the timed CPU baseline)
double orc_render_rows(void *sp, uint32_t seed, uint32_t spp, uint32_t spp_pass_lim, int brute, int nthreads, float *film,
                       uint32_t row_begin, uint32_t row_end, uint32_t row_stride, uint64_t *stats) {
    Scene &s = *(Scene *) sp;
    uint32_t spp_per_pass = spp_pass_lim ? std::min(spp_pass_lim, spp) : spp;
    uint32_t n_passes = spp / spp_per_pass;
    if (row_end > s.film_h) row_end = s.film_h;
    if (!row_stride) row_stride = 1;
    std::atomic<uint32_t> next_row{ 0 };
    std::vector<TraverseStats> tstats(std::max(nthreads, 1));
    // a wider filter splats across rows: every thread then accumulates into its own film, summed in thread order at the end
    size_t film_floats = 4 * (size_t) s.film_w * s.film_h;
    std::vector<std::vector<float>> tfilm; if (s.filter_type != 0 && nthreads > 1) tfilm.assign(nthreads, std::vector<float>(film_floats, 0.f));
    auto t0 = std::chrono::steady_clock::now();
    const uint32_t n_adapt = s.n_reuse ? std::min(s.adaptive, s.n_reuse - 1u) : 0u;
    if (n_adapt && !s.debug && row_stride == 1 && row_begin == 0 && row_end == s.film_h) {
        // render_multisample with the fill-in pass (mvpath_multi.h:79-115), pass by pass: (1) every lane of the wavefront, recording
        // adapt_mask; (2) idx = compress(mask) in lane order, each index repeated n_adapt times; (3) a forked sampler seeded
        // (wavefront2, wavefront2) traces one plain path per entry through the primary sample's own camera ray and puts it with weight adapt_w
        auto add = [](float *p, float v) { uint32_t o, n; std::memcpy(&o, p, 4);
            do { float f; std::memcpy(&f, &o, 4); f += v; std::memcpy(&n, &f, 4); } while (!__atomic_compare_exchange_n((uint32_t *) p, &o, n, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)); };
        auto put = [&](uint32_t ux, uint32_t uy, V3 v, float w) { float *px = film + 4 * ((size_t) uy * s.film_w + ux); add(px, v.x); add(px + 1, v.y); add(px + 2, v.z); add(px + 3, w); };
        const float adapt_w = 1.f / (float) (n_adapt + 1u);
        const size_t wavefront = (size_t) s.film_w * s.film_h * spp_per_pass;
        std::vector<uint8_t> mask(wavefront);
        auto parallel = [&](size_t count, const std::function<void(size_t, int)> &fn) {
            std::atomic<size_t> next{ 0 }; const size_t chunk = 4096;
            auto w = [&](int tid) { for (;;) { size_t b = next.fetch_add(chunk); if (b >= count) break; for (size_t i = b; i < std::min(count, b + chunk); ++i) fn(i, tid); } };
            if (nthreads <= 1) w(0); else { std::vector<std::thread> th; for (int i = 0; i < nthreads; ++i) th.emplace_back(w, i); for (auto &t : th) t.join(); }
        };
        for (uint32_t pass = 0; pass < n_passes; ++pass) {
            const uint32_t seed_value = s.base_seed + (spp_per_pass * pass + seed);
            parallel(wavefront, [&](size_t lane, int tid) { bool m = false;
                render_lane_multi(s, seed_value, (uint32_t) lane, spp_per_pass, s.n_reuse, brute != 0, nullptr, &tstats[tid], put, nullptr, &m); mask[lane] = m ? 1 : 0; });
            std::vector<uint32_t> idx; for (size_t lane = 0; lane < wavefront; ++lane) if (mask[lane]) idx.push_back((uint32_t) lane);
            const size_t wavefront2 = idx.size() * n_adapt; if (!wavefront2) continue;
            const uint32_t seed2 = s.base_seed + (uint32_t) wavefront2;              // sampler2->seed(wavefront, wavefront), sampler.cpp:125-144
            parallel(wavefront2, [&](size_t j, int tid) {
                const uint32_t src = idx[j / n_adapt];
                LaneOut o = render_lane(s, seed_value, src, spp_per_pass, brute != 0, nullptr, &tstats[tid], /* trace = */ false);   // sample_pos, aperture, ray of the primary sample
                Pcg32 rng2 = sampler_seed(seed2, (uint32_t) j);
                PathResult r = sample_single(s, rng2, o.ray, brute != 0, nullptr, &tstats[tid]);
                if (o.in_film) put(o.pixel % s.film_w, o.pixel / s.film_w, V3(adapt_w) * r.L, adapt_w); });
        }
        double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (stats) { stats[0] = stats[1] = 0; for (auto &t : tstats) { stats[0] += t.nodes; stats[1] += t.tris; } }
        return sec;
    }
    auto worker = [&](int tid) {
        for (;;) {
            uint64_t yy = (uint64_t) row_begin + (uint64_t) next_row.fetch_add(1) * row_stride; if (yy >= row_end) break;
            uint32_t y = (uint32_t) yy;
            for (uint32_t pass = 0; pass < n_passes; ++pass) {
                uint32_t seed_value = s.base_seed + (spp_per_pass * pass + seed);
                for (uint32_t x = 0; x < s.film_w; ++x) for (uint32_t k = 0; k < spp_per_pass; ++k) {
                    uint32_t lane = (y * s.film_w + x) * spp_per_pass + k;
                    if (s.n_reuse) {          // reused samples land in other view tiles (other threads' rows): atomic float adds
                        auto add = [](float *p, float v) { uint32_t o, n; std::memcpy(&o, p, 4);
                            do { float f; std::memcpy(&f, &o, 4); f += v; std::memcpy(&n, &f, 4); } while (!__atomic_compare_exchange_n((uint32_t *) p, &o, n, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)); };
                        render_lane_multi(s, seed_value, lane, spp_per_pass, s.n_reuse, brute != 0, nullptr, &tstats[tid],
                                          [&](uint32_t ux, uint32_t uy, V3 v, float w) { float *px = film + 4 * ((size_t) uy * s.film_w + ux);
                                              if (nthreads > 1) { add(px, v.x); add(px + 1, v.y); add(px + 2, v.z); add(px + 3, w); } else { px[0] += v.x; px[1] += v.y; px[2] += v.z; px[3] += w; } });
                        continue;
                    }
                    LaneOut o = render_lane(s, seed_value, lane, spp_per_pass, brute != 0, nullptr, &tstats[tid]);
                    // a jittered position that rounds up into the next pixel/row is extremely rare; handled with the same index rule
                    if (s.filter_type != 0) { filter_put(s, tfilm.empty() ? film : tfilm[tid].data(), o.sample_pos[0], o.sample_pos[1], o.r.L); continue; }
                    if (!o.in_film) continue;
                    float *px = film + 4 * (size_t) o.pixel;
                    px[0] += o.r.L.x; px[1] += o.r.L.y; px[2] += o.r.L.z; px[3] += 1.f;
                }
            }
        }
    };
    if (nthreads <= 1) worker(0);
    else { std::vector<std::thread> th; for (int i = 0; i < nthreads; ++i) th.emplace_back(worker, i); for (auto &t : th) t.join(); }
    for (auto &tf : tfilm) for (size_t i = 0; i < film_floats; ++i) film[i] += tf[i];
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (stats) { stats[0] = stats[1] = 0; for (auto &t : tstats) { stats[0] += t.nodes; stats[1] += t.tris; } }
    return sec;
}
// The film of an adaptive render whose allocation is given: pass p traces, for every pixel of view v, the lanes k < alloc[p][v] of the pass's
// lane map (lane = pixel * spp_per_pass + k, seed_value = base_seed + spp_per_pass * p + seed) -- exactly the samples the product claims it traced.
void orc_render_alloc(void *sp, uint32_t seed, uint32_t spp_per_pass, uint32_t n_passes, const uint32_t *alloc, int brute, int nthreads, float *film) {
    Scene &s = *(Scene *) sp; const uint32_t n_views = s.grid_x * s.grid_y, rx = s.film_w / s.grid_x, ry = s.film_h / s.grid_y;
    std::atomic<uint32_t> next_row{ 0 };
    auto worker = [&](int) {
        for (;;) {
            const uint32_t y = next_row.fetch_add(1); if (y >= s.film_h) break;
            for (uint32_t pass = 0; pass < n_passes; ++pass) {
                const uint32_t seed_value = s.base_seed + (spp_per_pass * pass + seed);
                for (uint32_t x = 0; x < s.film_w; ++x) {
                    uint32_t ix = x / rx, iy = y / ry; if (s.reverse_x) ix = s.grid_x - 1 - ix; if (s.reverse_y) iy = s.grid_y - 1 - iy;
                    const uint32_t view = ix + s.grid_x * iy, nk = alloc[(size_t) pass * n_views + view];
                    for (uint32_t k = 0; k < nk; ++k) {
                        LaneOut o = render_lane(s, seed_value, (y * s.film_w + x) * spp_per_pass + k, spp_per_pass, brute != 0, nullptr, nullptr);
                        if (!o.in_film) continue;
                        float *px = film + 4 * (size_t) o.pixel;       // strays into another row are rare; rows are owned by one thread at a time except for them
                        px[0] += o.r.L.x; px[1] += o.r.L.y; px[2] += o.r.L.z; px[3] += 1.f;
                    }
                }
            }
        }
    };
    if (nthreads <= 1) worker(0); else { std::vector<std::thread> th; for (int i = 0; i < nthreads; ++i) th.emplace_back(worker, i); for (auto &t : th) t.join(); }
}
// Per-lane tap of the reuse branch: for lanes [lane_begin, lane_begin + n) the n_reuse SampleData records after
// render_multisample's tile offset: out[lane][k] = { pos.x, pos.y, weight, result.rgb, valid, indirect, idx } (9 floats)
void orc_trace_lanes_multi(void *sp, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, int brute, float *out) {
    Scene &s = *(Scene *) sp; uint32_t nr = s.n_reuse; if (!nr) return;
    for (uint32_t i = 0; i < n; ++i) {
        SampleData sd[MaxReuse];
        bool adapt = false;
        render_lane_multi(s, seed_value, lane_begin + i, spp_per_pass, nr, brute != 0, nullptr, nullptr, [](uint32_t, uint32_t, V3, float) {}, sd, &adapt);
        for (uint32_t k = 0; k < nr; ++k) {
            float *o = out + 9 * ((size_t) i * nr + k);
            o[0] = sd[k].pos.x; o[1] = sd[k].pos.y; o[2] = sd[k].weight; o[3] = sd[k].result.x; o[4] = sd[k].result.y; o[5] = sd[k].result.z;
            o[6] = (sd[k].valid ? 1.f : 0.f) + (k == 0 && adapt ? 2.f : 0.f); o[7] = sd[k].indirect ? 1.f : 0.f; o[8] = (float) sd[k].idx;      // + 2: adapt_mask of the lane (record 0 only)
        }
    }
}
// HDRFilm::develop (src/films/hdrfilm.cpp:321-408): rgb = RGB / (W == 0 ? 1 : W)
void orc_develop(const float *rgbw, uint64_t npix, float *rgb) {
    for (uint64_t i = 0; i < npix; ++i) { float w = rgbw[4 * i + 3]; float iw = (w == 0.f ? 1.f : w);
        rgb[3 * i] = rgbw[4 * i] / iw; rgb[3 * i + 1] = rgbw[4 * i + 1] / iw; rgb[3 * i + 2] = rgbw[4 * i + 2] / iw; }
}

// display step, see o_display.h: Bitmap::convert(RGB, UInt8, srgb) and Program::display_image
void orc_linear_to_srgb(const float *in, uint64_t n, float *out) { for (uint64_t i = 0; i < n; ++i) out[i] = odisp::linear_to_srgb(in[i]); }
void orc_to_srgb8(const float *in, uint64_t n, uint8_t *out) { for (uint64_t i = 0; i < n; ++i) out[i] = odisp::to_srgb8(in[i]); }
void orc_clip_mod(const float *x, const float *y, int n, float *out_clip, float *out_mod) {
    for (int i = 0; i < n; ++i) { out_clip[i] = odisp::clip(x[i]); out_mod[i] = odisp::mod(x[i], y[i]); }
}
void orc_interpolate2d_u8(const uint8_t *data, uint32_t w, uint32_t h, uint32_t nch, const float *xy, int n, uint32_t c, int nearest, uint8_t *out) {
    for (int i = 0; i < n; ++i) out[i] = odisp::interpolate2d(data, xy[2 * i], xy[2 * i + 1], w, h, c, nch, nearest != 0);
}
// preset: 7 floats (center, focus, pitch, tilt, subp, view x, view y) + 5 ints (grid x, grid y, flip_rgb, quilt, nearest)
void orc_interlace(const uint8_t *src, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const float *pf, const int32_t *pi,
                   uint8_t *dst, uint32_t dst_w, uint32_t dst_h, uint32_t dst_ch) {
    odisp::Preset p{ pf[0], pf[1], pf[2], pf[3], pf[4], { pf[5], pf[6] }, { pi[0], pi[1] }, pi[2], pi[3], pi[4] };
    odisp::interlace(src, src_w, src_h, src_ch, p, dst, dst_w, dst_h, dst_ch);
}

} // extern "C"
