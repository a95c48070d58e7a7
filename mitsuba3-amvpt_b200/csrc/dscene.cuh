// dscene.cuh -- device-resident scene of the AMVPT path: sub-sensor table, BVH, triangles, emitters,
// and the per-ray / per-vertex device functions (camera ray, closest hit, any hit, surface
// interaction, emitter sampling).  Data layout is described in DESIGN.md "Data layout in HBM".
#pragma once
#include "dtypes.h"

namespace amvpt {

struct Ray { V3 o, d; float maxt; };
struct Hit { float t, u, v; uint32_t tri; };     // tri = position in DScene::tris, 0xffffffff = miss
#define AM_NO_HIT 0xffffffffu

AM_D float4 ldg4(const float4 *p) { return __ldg(p); }

// ---------------------------------------------------------------- sensors
// Transform4f * Point3f with perspective divide (include/mitsuba/core/transform.h:134-144)
AM_D V3 xf_point_proj(const float *m, V3 p) {
    float r[4];
    #pragma unroll
    for (int i = 0; i < 4; ++i) r[i] = m[4 * i + 3];
    #pragma unroll
    for (int i = 0; i < 4; ++i) r[i] = fmadd(m[4 * i + 0], p.x, r[i]);
    #pragma unroll
    for (int i = 0; i < 4; ++i) r[i] = fmadd(m[4 * i + 1], p.y, r[i]);
    #pragma unroll
    for (int i = 0; i < 4; ++i) r[i] = fmadd(m[4 * i + 2], p.z, r[i]);
    float iw = rcp(r[3]);
    return v3(r[0] * iw, r[1] * iw, r[2] * iw);
}
AM_D V3 xf_point_affine(const float *m, V3 p) {   // transform.h:121-129
    return v3(fmadd(m[2], p.z, fmadd(m[1], p.y, fmadd(m[0], p.x, m[3]))),
              fmadd(m[6], p.z, fmadd(m[5], p.y, fmadd(m[4], p.x, m[7]))),
              fmadd(m[10], p.z, fmadd(m[9], p.y, fmadd(m[8], p.x, m[11]))));
}
AM_D V3 xf_vector(const float *m, V3 v) {         // transform.h:150-160
    return v3(fmadd(m[2], v.z, fmadd(m[1], v.y, m[0] * v.x)),
              fmadd(m[6], v.z, fmadd(m[5], v.y, m[4] * v.x)),
              fmadd(m[10], v.z, fmadd(m[9], v.y, m[8] * v.x)));
}
// GridSensor::sample_ray (src/sensors/grid.cpp:238-267) [BatchSensor::sample_ray src/sensors/batch.cpp:132-160 is
// the grid_y = 1, reverse_y = 0 case] -> PerspectiveCamera::sample_ray / ThinLensCamera::sample_ray.
AM_D Ray sensor_sample_ray(const DScene &s, V2 pos, V2 aperture, uint32_t *view_out) {
    float fx = pos.x * (float) s.grid_x, fy = pos.y * (float) s.grid_y;
    uint32_t ux = (uint32_t) fx, uy = (uint32_t) fy, ix = ux, iy = uy;
    if (s.reverse_x) ix = (s.grid_x - 1) - ix;
    if (s.reverse_y) iy = (s.grid_y - 1) - iy;
    uint32_t index = ix + s.grid_x * iy;
    index = min(index, s.n_sensors - 1u);
    V2 p2 = v2(fx - (float) ux, fy - (float) uy);
    *view_out = index;
    const float4 *c4 = reinterpret_cast<const float4 *>(s.sensors + index);
    float m[32];
    #pragma unroll
    for (int i = 0; i < 8; ++i) { float4 q = ldg4(c4 + i); m[4 * i] = q.x; m[4 * i + 1] = q.y; m[4 * i + 2] = q.z; m[4 * i + 3] = q.w; }
    const float *s2c = m, *tw = m + 16;
    float near_clip = m[28], far_clip = m[29], aperture_radius = m[30], focus_distance = m[31];
    V3 near_p = xf_point_proj(s2c, v3(p2.x + 0.f, p2.y + 0.f, 0.f));
    Ray ray; V3 d;
    if (!s.thinlens) {
        d = normalize(near_p);
        ray.o = v3(tw[3], tw[7], tw[11]);
    } else {
        V2 tmp = square_to_uniform_disk_concentric(aperture);
        V3 ap = v3(aperture_radius * tmp.x, aperture_radius * tmp.y, 0.f);
        V3 focus_p = near_p * fdiv(focus_distance, near_p.z);
        d = normalize(focus_p - ap);
        ray.o = xf_point_affine(tw, ap);
    }
    ray.d = xf_vector(tw, d);
    float inv_z = rcp(d.z), near_t = near_clip * inv_z, far_t = far_clip * inv_z;
    ray.o = ray.o + ray.d * near_t; ray.maxt = far_t - near_t;
    return ray;
}

// ---------------------------------------------------------------- ray / triangle, BVH traversal
// Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:467-488).  The reference evaluates u, v, t unconditionally and
// combines the masks; leaving as soon as one test fails skips the remaining products without changing any value that
// is observed (t, u, v are only read when every test passed).
AM_D bool moeller_trumbore(const Ray &ray, V3 p0, V3 p1, V3 p2, float *t_o, float *u_o, float *v_o) {
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = cross(ray.d, e2);
    float inv_det = rcp(dot(e1, pvec));
    V3 tvec = ray.o - p0;
    float u = dot(tvec, pvec) * inv_det;
    if (!(u >= 0.f && u <= 1.f)) return false;
    V3 qvec = cross(tvec, e1);
    float v = dot(ray.d, qvec) * inv_det;
    if (!(v >= 0.f && u + v <= 1.f)) return false;
    float t = dot(e2, qvec) * inv_det;
    if (!(t >= 0.f && t <= ray.maxt)) return false;
    *t_o = t; *u_o = u; *v_o = v;
    return true;
}

#define AM_SLAB_EPS 2e-6f
AM_D float fmax3(float a, float b, float c) { float r; asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
AM_D float fmin3(float a, float b, float c) { float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
// Per-ray constants of the slab test: id = 1 / d (clamped to a large finite value for axis-parallel rays so that no
// inf - inf arises) and ood = o * id, so that the plane distances are one FMA each: t = plane * id - ood.
struct SlabRay { V3 id, ood; };
// 1 / d to within 1 ulp (MUFU.RCP): the slab test only prunes, its reciprocal is not part of the arithmetic contract
AM_D float slab_inv(float d) { float a; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(a) : "f"(d)); return fabs_(d) < 1e-20f ? copysign_(1e20f, d) : a; }
AM_D SlabRay slab_ray(V3 o, V3 d) {
    SlabRay r; r.id = v3(slab_inv(d.x), slab_inv(d.y), slab_inv(d.z)); r.ood = v3(o.x * r.id.x, o.y * r.id.y, o.z * r.id.z); return r;
}
// Slab test of one child box.  It only prunes, so it is conservative rather than exact: plane * id - ood carries an
// absolute error of about 2^-24 (|o| + |plane|) |id| (one rounding of ood, one of the FMA), i.e. 2^-24 * scene scale in
// space, and a relative one of 2^-23; the builder pads every box by 2e-6 * scene scale + 1e-5 |coordinate| (bvh.cpp) and the
// interval is widened by AM_SLAB_EPS.  min/max are the non-NaN-propagating PTX forms (FMNMX / FMNMX3).
AM_D bool slab(float lox, float hix, float loy, float hiy, float loz, float hiz, const SlabRay &r, float tbest, float *tnear) {
    float t0x = fmadd(lox, r.id.x, -r.ood.x), t1x = fmadd(hix, r.id.x, -r.ood.x);
    float t0y = fmadd(loy, r.id.y, -r.ood.y), t1y = fmadd(hiy, r.id.y, -r.ood.y);
    float t0z = fmadd(loz, r.id.z, -r.ood.z), t1z = fmadd(hiz, r.id.z, -r.ood.z);
    float tmin = fmaxf(fmax3(fminf(t0x, t1x), fminf(t0y, t1y), fminf(t0z, t1z)), 0.f);
    float tmax = fminf(fmin3(fmaxf(t0x, t1x), fmaxf(t0y, t1y), fmaxf(t0z, t1z)), tbest);
    *tnear = tmin;
    return tmin <= tmax * (1.f + AM_SLAB_EPS);
}

// One inner-node step of the traversal: tests the children's boxes, continues with the nearest hit child and pushes the
// others so that the nearer one is popped first; pops when nothing is hit.  Returns the next node reference.
#if AM_BVH_WIDTH == 4
AM_D int32_t node_step(const DScene &s, const SlabRay &r, float best, int32_t node, int32_t *stack, int &sp) {
    const float4 *np = s.nodes + 8 * (size_t) node;
    const float4 lx = ldg4(np), hx = ldg4(np + 1), ly = ldg4(np + 2), hy = ldg4(np + 3), lz = ldg4(np + 4), hz = ldg4(np + 5), cf = ldg4(np + 6);
    const int32_t c0 = __float_as_int(cf.x), c1 = __float_as_int(cf.y), c2 = __float_as_int(cf.z), c3 = __float_as_int(cf.w);
    float t0, t1, t2, t3;
    const bool h0 = slab(lx.x, hx.x, ly.x, hy.x, lz.x, hz.x, r, best, &t0) && c0 != 0;
    const bool h1 = slab(lx.y, hx.y, ly.y, hy.y, lz.y, hz.y, r, best, &t1) && c1 != 0;
    const bool h2 = slab(lx.z, hx.z, ly.z, hy.z, lz.z, hz.z, r, best, &t2) && c2 != 0;
    const bool h3 = slab(lx.w, hx.w, ly.w, hy.w, lz.w, hz.w, r, best, &t3) && c3 != 0;
    const int n = (int) h0 + (int) h1 + (int) h2 + (int) h3;
    if (n == 0) return stack[--sp];
    if (n == 1) return h0 ? c0 : (h1 ? c1 : (h2 ? c2 : c3));
    // rank of a hit child = number of hit children nearer than it (ties: lower slot first); rank 0 is visited next, the others
    // are pushed so that rank 1 ends on top of the stack
    t0 = h0 ? t0 : AM_INF; t1 = h1 ? t1 : AM_INF; t2 = h2 ? t2 : AM_INF; t3 = h3 ? t3 : AM_INF;
    const int r0 = (int) (t1 < t0) + (int) (t2 < t0) + (int) (t3 < t0);
    const int r1 = (int) (t0 <= t1) + (int) (t2 < t1) + (int) (t3 < t1);
    const int r2 = (int) (t0 <= t2) + (int) (t1 <= t2) + (int) (t3 < t2);
    const int r3 = (int) (t0 <= t3) + (int) (t1 <= t3) + (int) (t2 <= t3);
    int32_t next = c3; const int top = sp + n - 1;
    if (h0) { if (r0 == 0) next = c0; else stack[top - r0] = c0; }
    if (h1) { if (r1 == 0) next = c1; else stack[top - r1] = c1; }
    if (h2) { if (r2 == 0) next = c2; else stack[top - r2] = c2; }
    if (h3 && r3 != 0) stack[top - r3] = c3;
    sp = top;
    return next;
}
#else
#if AM_FFMA2
// d = a * b + c on two packed binary32 lanes (sm_100 FFMA2; each lane is the IEEE fma)
AM_D float2 fma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{ .reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mov.b64 rc, {%6, %7}; fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0, %1}, rd; }"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
// slab test of one child box from (lo.x, lo.y, hi.x, hi.y) and (lo.z, hi.z): three FFMA2 instead of six FFMA, same values
AM_D bool slab2(float4 xy, float loz, float hiz, const SlabRay &r, float tbest, float *tnear) {
    const float2 idxy = make_float2(r.id.x, r.id.y), nxy = make_float2(-r.ood.x, -r.ood.y), idzz = make_float2(r.id.z, r.id.z), nzz = make_float2(-r.ood.z, -r.ood.z);
    const float2 t0 = fma2(make_float2(xy.x, xy.y), idxy, nxy), t1 = fma2(make_float2(xy.z, xy.w), idxy, nxy), tz = fma2(make_float2(loz, hiz), idzz, nzz);
    float tmin = fmaxf(fmax3(fminf(t0.x, t1.x), fminf(t0.y, t1.y), fminf(tz.x, tz.y)), 0.f);
    float tmax = fminf(fmin3(fmaxf(t0.x, t1.x), fmaxf(t0.y, t1.y), fmaxf(tz.x, tz.y)), tbest);
    *tnear = tmin;
    return tmin <= tmax * (1.f + AM_SLAB_EPS);
}
#endif
AM_D int32_t node_step(const DScene &s, const SlabRay &r, float best, int32_t node, int32_t *stack, int &sp) {
    const float4 *np = s.nodes + 4 * (size_t) node;
    float4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3);
    float ta, tb;
#if AM_FFMA2
    bool ha = slab2(n0, n2.x, n2.y, r, best, &ta);
    bool hb = slab2(n1, n2.z, n2.w, r, best, &tb);
#else
    bool ha = slab(n0.x, n0.y, n0.z, n0.w, n2.x, n2.y, r, best, &ta);
    bool hb = slab(n1.x, n1.y, n1.z, n1.w, n2.z, n2.w, r, best, &tb);
#endif
    int32_t ca = __float_as_int(n3.x), cb = __float_as_int(n3.y);
    if (ha && hb) {
        bool swap = tb < ta; int32_t far = swap ? ca : cb;
#if AM_PREFETCH_FAR       // the postponed child's record is pulled towards L1 while the near subtree is walked
        if (far >= 0) asm volatile("prefetch.global.L1 [%0];" :: "l"(s.nodes + 4 * (size_t) far));
#endif
        stack[sp++] = far; return swap ? cb : ca;
    }
    if (ha || hb) return ha ? ca : cb;
    return stack[--sp];
}
#endif

// One step through the compressed 4-wide tree (DScene::nodes_q4).  The child planes are origin + q * 2^(e - 127); with the per-node
// products sa = 2^(e - 127) * id and ca = origin * id - ood every plane distance is one FMA: t = q * sa + ca.  Children are visited
// nearest first (rank = number of hit children nearer than this one), like the uncompressed 4-wide step above.
AM_D float q8(uint32_t w, int i) { return (float) ((w >> (8 * i)) & 0xffu); }
AM_D int32_t node_step_q4(const DScene &s, const SlabRay &r, float best, int32_t node, int32_t *stack, int &sp) {
    const uint4 *np = s.nodes_q4 + 4 * (size_t) node;
    const uint4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2); const uint2 q3 = __ldg(reinterpret_cast<const uint2 *>(np + 3));
    const float sx = __uint_as_float((q0.w & 0xffu) << 23) * r.id.x, sy = __uint_as_float(((q0.w >> 8) & 0xffu) << 23) * r.id.y, sz = __uint_as_float(((q0.w >> 16) & 0xffu) << 23) * r.id.z;
    const float cx = fmadd(__uint_as_float(q0.x), r.id.x, -r.ood.x), cy = fmadd(__uint_as_float(q0.y), r.id.y, -r.ood.y), cz = fmadd(__uint_as_float(q0.z), r.id.z, -r.ood.z);
    const int32_t c[4] = { (int32_t) q2.z, (int32_t) q2.w, (int32_t) q3.x, (int32_t) q3.y };
    float t[4]; bool h[4]; int n = 0;
    #pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float t0x = fmadd(q8(q1.x, i), sx, cx), t1x = fmadd(q8(q1.w, i), sx, cx);
        const float t0y = fmadd(q8(q1.y, i), sy, cy), t1y = fmadd(q8(q2.x, i), sy, cy);
        const float t0z = fmadd(q8(q1.z, i), sz, cz), t1z = fmadd(q8(q2.y, i), sz, cz);
        const float tmin = fmaxf(fmax3(fminf(t0x, t1x), fminf(t0y, t1y), fminf(t0z, t1z)), 0.f);
        const float tmax = fminf(fmin3(fmaxf(t0x, t1x), fmaxf(t0y, t1y), fmaxf(t0z, t1z)), best);
        h[i] = tmin <= tmax * (1.f + AM_SLAB_EPS) && c[i] != 0; t[i] = h[i] ? tmin : AM_INF; n += (int) h[i];
    }
    if (n == 0) return stack[--sp];
    if (n == 1) return h[0] ? c[0] : (h[1] ? c[1] : (h[2] ? c[2] : c[3]));
    const int r0 = (int) (t[1] < t[0]) + (int) (t[2] < t[0]) + (int) (t[3] < t[0]);
    const int r1 = (int) (t[0] <= t[1]) + (int) (t[2] < t[1]) + (int) (t[3] < t[1]);
    const int r2 = (int) (t[0] <= t[2]) + (int) (t[1] <= t[2]) + (int) (t[3] < t[2]);
    const int r3 = (int) (t[0] <= t[3]) + (int) (t[1] <= t[3]) + (int) (t[2] <= t[3]);
    int32_t next = c[3]; const int top = sp + n - 1;
    if (h[0]) { if (r0 == 0) next = c[0]; else stack[top - r0] = c[0]; }
    if (h[1]) { if (r1 == 0) next = c[1]; else stack[top - r1] = c[1]; }
    if (h[2]) { if (r2 == 0) next = c[2]; else stack[top - r2] = c[2]; }
    if (h[3] && r3 != 0) stack[top - r3] = c[3];
    sp = top;
    return next;
}

// Closest hit (ANY = false) with the tie rule of ShapeKDTree::ray_intersect_naive
// (include/mitsuba/render/kdtree.h:2445-2468): among equal t the primitive with the larger global index
// wins.  ANY = true: Scene::ray_test semantics, first accepted triangle ends the traversal.
// "while-while" traversal: every lane first descends through inner nodes until it holds a leaf, the warp reconverges,
// then the lanes intersect their leaves together -- the triangle test is the expensive, divergence-sensitive part.
#define AM_TRAVERSAL_DONE ((int32_t) 0x80000000)      /* not a valid leaf code: first would be 2^28 - 1 */
// ACCEL = 1: the caller knows the scene has a flat leaf list and asks for no counters -- only that path is compiled in (the megakernel instantiation
// for small scenes: no traversal stack, no tree code).
template <bool ANY, bool Q4 = false, int ACCEL = 0> AM_D Hit intersect(const DScene &s, const Ray &ray, uint32_t *n_nodes = nullptr, uint32_t *n_tris = nullptr) {
    Hit h; h.t = AM_INF; h.u = 0.f; h.v = 0.f; h.tri = AM_NO_HIT;
    uint32_t best_prim = 0;
    float best = ray.maxt;
    SlabRay sr = slab_ray(ray.o, ray.d);
#if AM_FLAT_LEAVES
    if (ACCEL == 1 || (!Q4 && s.n_flat && !n_nodes)) {      // passes that count tree steps (stats, the roofline leg) walk the tree
        // Small scene: every lane tests ALL leaf boxes in one loop of fixed length (no stack, no divergence: the binary tree's node loop runs
        // at 11 of 32 lanes because rays need different numbers of steps), remembering which boxes it hit and which of them is nearest;
        // then the lanes visit their leaves, nearest first, re-testing each further box against the hit distance found so far.
        uint32_t mask = 0, k_near = 0; float t_near = AM_INF;
        for (uint32_t k = 0; k < s.n_flat; ++k) {
            const float4 lo = ldg4(s.flat + 2 * k), hi = ldg4(s.flat + 2 * k + 1); float tn;
            if (slab(lo.x, hi.x, lo.y, hi.y, lo.z, hi.z, sr, best, &tn)) { mask |= 1u << k; if (tn < t_near) { t_near = tn; k_near = k; } }
        }
        bool first = true;
        while (mask) {
            uint32_t k;
            if (first) { k = k_near; first = false; } else k = (uint32_t) __ffs((int) mask) - 1u;
            mask &= ~(1u << k);
            const float4 lo = ldg4(s.flat + 2 * k);
            if (!ANY && h.tri != AM_NO_HIT) {        // a hit exists: is this box still in front of it?
                const float4 hi = ldg4(s.flat + 2 * k + 1); float tn;
                if (!slab(lo.x, hi.x, lo.y, hi.y, lo.z, hi.z, sr, best, &tn)) continue;
            }
            const uint32_t code = ~__float_as_uint(lo.w), first_tri = code >> 3, count = code & 7u;
            for (uint32_t i = 0; i < count; ++i) {
                const float4 *tp = s.tris + 3 * (size_t) (first_tri + i);
                float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
                if (n_tris) (*n_tris)++;
                float t, u, v;
                if (moeller_trumbore(ray, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &u, &v)) {
                    uint32_t prim = __float_as_uint(a.w);
                    if (ANY) { h.t = t; h.u = u; h.v = v; h.tri = first_tri + i; return h; }
                    if (h.tri == AM_NO_HIT || t < h.t || (t == h.t && prim > best_prim)) {
                        h.t = t; h.u = u; h.v = v; h.tri = first_tri + i; best_prim = prim; best = t;
                    }
                }
            }
        }
        return h;
    }
    if (ACCEL == 1) return h;
#endif
    int32_t stack[AM_STACK_SIZE]; int sp = 0;
    int32_t node = 0;
    stack[sp++] = AM_TRAVERSAL_DONE;
    for (;;) {
#if AM_SPECULATIVE
        // speculative descent (Aila & Laine): the first leaf a lane finds is postponed and the lane keeps walking until it holds a second
        // leaf or its ray is done, so that the warp alternates half as often between the node loop and the leaf loop.  The postponed
        // leaf's hit cannot prune the extra steps, so a few more nodes are visited; the reported hit is the same (closest hit and tie
        // rule do not depend on the order in which leaves are tested; for ANY only occlusion is defined).
        int32_t pend = AM_TRAVERSAL_DONE;
        for (;;) {
            while (node >= 0) {
                if (n_nodes) (*n_nodes)++;
                node = Q4 ? node_step_q4(s, sr, best, node, stack, sp) : node_step(s, sr, best, node, stack, sp);
            }
            if (node == AM_TRAVERSAL_DONE || pend != AM_TRAVERSAL_DONE) break;
            pend = node; node = stack[--sp];
        }
        if (pend == AM_TRAVERSAL_DONE) break;                       // no leaf at all: the ray is done
        for (int which = 0; which < 2; ++which) {
            const int32_t leaf = which == 0 ? pend : node;
            if (leaf == AM_TRAVERSAL_DONE) break;
            uint32_t code = ~(uint32_t) leaf, first = code >> 3, count = code & 7u;
#else
        while (node >= 0) {
            if (n_nodes) (*n_nodes)++;
            node = Q4 ? node_step_q4(s, sr, best, node, stack, sp) : node_step(s, sr, best, node, stack, sp);
        }
        if (node == AM_TRAVERSAL_DONE) break;
        {
            uint32_t code = ~(uint32_t) node, first = code >> 3, count = code & 7u;
#endif
            for (uint32_t i = 0; i < count; ++i) {
                const float4 *tp = s.tris + 3 * (size_t) (first + i);
                float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
                if (n_tris) (*n_tris)++;
                float t, u, v;
                if (moeller_trumbore(ray, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &u, &v)) {
                    uint32_t prim = __float_as_uint(a.w);
                    if (ANY) { h.t = t; h.u = u; h.v = v; h.tri = first + i; return h; }
                    if (h.tri == AM_NO_HIT || t < h.t || (t == h.t && prim > best_prim)) {
                        h.t = t; h.u = u; h.v = v; h.tri = first + i; best_prim = prim; best = t;
                    }
                }
            }
        }
#if AM_SPECULATIVE
        if (node == AM_TRAVERSAL_DONE) break;
#endif
        node = stack[--sp];
    }
    return h;
}

// ---------------------------------------------------------------- surface interaction
struct SI {
    bool valid; float t; V3 p, n, sh_n, sh_s, sh_t, wi; uint32_t mesh, prim;
    AM_D V3 to_local(V3 v) const { return v3(dot(v, sh_s), dot(v, sh_t), dot(v, sh_n)); }          // frame.h:34-36
    AM_D V3 to_world(V3 v) const { return fmadd(sh_n, v.z, fmadd(sh_t, v.y, sh_s * v.x)); }        // frame.h:39-41
};
// Interaction::offset_p / spawn_ray / spawn_ray_to (include/mitsuba/render/interaction.h:140-170)
AM_D V3 offset_p(V3 p, V3 n, V3 d) {
    float mag = (1.f + max_(max_(fabs_(p.x), fabs_(p.y)), fabs_(p.z))) * AM_RAY_EPSILON;
    mag = mulsign(mag, dot(n, d));
    return fmadd(n, mag, p);
}
AM_D Ray spawn_ray(V3 p, V3 n, V3 d) { Ray r; r.o = offset_p(p, n, d); r.d = d; r.maxt = AM_LARGEST; return r; }
AM_D Ray spawn_ray_to(V3 p, V3 n, V3 t) {
    Ray r; r.o = offset_p(p, n, t - p);
    V3 d = t - r.o; float dist = norm(d); r.d = d / dist;
    r.maxt = dist * (1.f - AM_SHADOW_EPSILON);
    return r;
}
// Mesh::compute_surface_interaction (src/render/mesh.cpp:1392-1565, meshes without texture coordinates) +
// SurfaceInteraction::finalize_surface_interaction / initialize_sh_frame (interaction.h:278-289, 497-517)
AM_D SI make_si(const DScene &s, const Ray &ray, const Hit &h, const DMesh **mesh_out) {
    SI si; si.valid = h.tri != AM_NO_HIT; si.t = h.t; si.wi = -ray.d; si.mesh = 0; si.prim = AM_NO_HIT;
    *mesh_out = nullptr;
    if (!si.valid) return si;
    const float4 *tp = s.tris + 3 * (size_t) h.tri;
    float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
    V3 p0 = v3(a.x, a.y, a.z), p1 = v3(b.x, b.y, b.z), p2 = v3(c.x, c.y, c.z);
    si.prim = __float_as_uint(a.w); si.mesh = __float_as_uint(b.w);
    const DMesh *m = s.meshes + si.mesh; *mesh_out = m;
    float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
    si.p = fmadd(p0, b0, fmadd(p1, b1, p2 * b2));
    si.n = normalize(cross(p1 - p0, p2 - p0));
    V3 dp_du, dp_dv; coordinate_system(si.n, &dp_du, &dp_dv);
    if (m->has_normals) {
        const float4 *np = s.tri_nrm + 3 * (size_t) h.tri;
        float4 na = ldg4(np), nb = ldg4(np + 1), nc = ldg4(np + 2);
        V3 n = fmadd(v3(nc.x, nc.y, nc.z), b2, fmadd(v3(nb.x, nb.y, nb.z), b1, v3(na.x, na.y, na.z) * b0));
        si.sh_n = n * rsqrt(squared_norm(n));
    } else si.sh_n = si.n;
    if (m->flip_normals) { si.n = -si.n; si.sh_n = -si.sh_n; }
    si.sh_s = normalize(fmadd(si.sh_n, -dot(si.sh_n, dp_du), dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { V3 t2; coordinate_system(si.sh_n, &si.sh_s, &t2); }
    si.sh_t = cross(si.sh_n, si.sh_s);
    si.wi = si.to_local(-ray.d);
    return si;
}

// ---------------------------------------------------------------- emitters
// DiscreteDistribution::sample (include/mitsuba/core/distr_1d.h:116-134) on dr::binary_search
// (ext/drjit/include/drjit/util.h:134-185): fixed iteration count log2i(n-1)+1
AM_D uint32_t distr_sample(const float *cdf, uint32_t n, float sum, float sample) {
    sample *= sum;
    uint32_t start = 0, end = n - 1, iterations = 0;
    if (start < end) iterations = 32u - (uint32_t) __clz(end - start);
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        float value = __ldg(cdf + middle);
        bool cond = ((value < sample) || value == 0.f) && value != sum;
        if (cond) start = min(middle + 1, end); else end = middle;
    }
    return start;
}
struct DirectionSample { V3 p, n, d; float pdf, dist; bool delta; };

// ---------------------------------------------------------------- emitters beyond area lights (GEN = true instantiations only)
// Hierarchical2D<Float, 0> (include/mitsuba/core/distr_2d.h): Level::index :757-761, sample :490-575, eval :668-698
AM_D uint32_t warp_level_index(uint32_t x, uint32_t y, uint32_t width) { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + (y & ~1u) * width; }
AM_D float warp_lookup(const DScene &s, uint32_t level, uint32_t i) { return __ldg(s.env_warp + s.env_level_off[level] + i); }
AM_D V2 env_warp_sample(const DScene &s, V2 sample, float *pdf) {
    sample.x = clip_(sample.x, 0.f, 1.f); sample.y = clip_(sample.y, 0.f, 1.f);
    uint32_t ox = 0, oy = 0;
    for (int l = (int) s.env_n_levels - 2; l > 0; --l) {
        ox <<= 1; oy <<= 1;
        const uint32_t oi = warp_level_index(ox, oy, s.env_level_width[l]);
        const float v00 = warp_lookup(s, l, oi), v10 = warp_lookup(s, l, oi + 1), v01 = warp_lookup(s, l, oi + 2), v11 = warp_lookup(s, l, oi + 3);
        sample.x = clip_(sample.x, 0.f, 1.f); sample.y = clip_(sample.y, 0.f, 1.f);
        const float r0 = v00 + v10, r1 = v01 + v11;
        sample.y *= r0 + r1;
        bool mask = sample.y > r0;
        if (mask) { oy += 1u; sample.y -= r0; }
        sample.y = fdiv(sample.y, mask ? r1 : r0);
        const float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
        sample.x *= c0 + c1;
        mask = sample.x > c0;
        if (mask) sample.x -= c0;
        sample.x = fdiv(sample.x, mask ? c1 : c0);
        if (mask) ox += 1u;
    }
    const uint32_t width = s.env_level_width[0], oi = ox + oy * width;
    const float v00 = warp_lookup(s, 0, oi), v10 = warp_lookup(s, 0, oi + 1), v01 = warp_lookup(s, 0, oi + width), v11 = warp_lookup(s, 0, oi + width + 1);
    sample = square_to_bilinear(v00, v10, v01, v11, sample, pdf);
    return v2(((float) (int32_t) ox + sample.x) * s.env_patch_size[0], ((float) (int32_t) oy + sample.y) * s.env_patch_size[1]);
}
AM_D float env_warp_eval(const DScene &s, V2 pos) {
    pos.x = clip_(pos.x, 0.f, 1.f) * s.env_inv_patch_size[0]; pos.y = clip_(pos.y, 0.f, 1.f) * s.env_inv_patch_size[1];
    const uint32_t ox = min((uint32_t) (int32_t) pos.x, s.env_max_patch[0]), oy = min((uint32_t) (int32_t) pos.y, s.env_max_patch[1]);
    pos.x -= (float) (int32_t) ox; pos.y -= (float) (int32_t) oy;
    const uint32_t width = s.env_level_width[0], oi = ox + oy * width;
    const float v00 = warp_lookup(s, 0, oi), v10 = warp_lookup(s, 0, oi + 1), v01 = warp_lookup(s, 0, oi + width), v11 = warp_lookup(s, 0, oi + width + 1);
    return square_to_bilinear_pdf(v00, v10, v01, v11, pos);
}
// EnvironmentMapEmitter::eval_spectrum (src/emitters/envmap.cpp:528-586, RGB branch): bilinear lookup, periodic in u
AM_D V3 env_eval_uv(const DScene &s, V2 uv) {
    uv.x -= fdiv(.5f, (float) (s.env_res_x - 1u));
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    uv.x *= (float) (s.env_res_x - 1u); uv.y *= (float) (s.env_res_y - 1u);
    const uint32_t px = min((uint32_t) uv.x, s.env_res_x - 2u), py = min((uint32_t) uv.y, s.env_res_y - 2u);
    const float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    const uint32_t width = s.env_res_x, index = py * width + px;
    const float *d = s.env_data;
    const V3 v00 = v3(__ldg(d + 3 * index), __ldg(d + 3 * index + 1), __ldg(d + 3 * index + 2)),
             v10 = v3(__ldg(d + 3 * (index + 1)), __ldg(d + 3 * (index + 1) + 1), __ldg(d + 3 * (index + 1) + 2)),
             v01 = v3(__ldg(d + 3 * (index + width)), __ldg(d + 3 * (index + width) + 1), __ldg(d + 3 * (index + width) + 2)),
             v11 = v3(__ldg(d + 3 * (index + width + 1)), __ldg(d + 3 * (index + width + 1) + 1), __ldg(d + 3 * (index + width + 1) + 2));
    const V3 v0 = fmadd(v00, w0x, v10 * w1x), v1 = fmadd(v01, w0x, v11 * w1x), v = fmadd(v0, w0y, v1 * w1y);
    return v * s.env_scale;
}
// direction (world space, pointing away from the scene) -> latitude-longitude coordinates (envmap.cpp:313-318, 468-473)
AM_D V2 env_dir_to_uv(const DScene &s, V3 d_world, V3 *d_local) {
    const V3 v = xf_vector(s.env_from_world, d_world); *d_local = v;
    return v2(atan2_(v.x, -v.z) * AM_INV_TWO_PI, safe_acos_(v.y) * AM_INV_PI);
}
// Emitter::eval for the environment emitter of the scene (constant.cpp:92-96, envmap.cpp:308-320); d = -si.wi = direction of the escaping ray
AM_D V3 env_eval(const DScene &s, V3 d) {
    if (!s.env_is_map) return ld3(s.emitters[s.env_emitter].radiance);
    V3 local; const V2 uv = env_dir_to_uv(s, d, &local);
    return env_eval_uv(s, uv);
}
#define AM_ENV_PDF_SCALE (1.f / (2.f * (AM_PI * AM_PI)))
// Emitter::pdf_direction of the environment emitter (constant.cpp:152-157, envmap.cpp:464-481), without the emitter pmf
AM_D float env_pdf_direction(const DScene &s, V3 d) {
    if (!s.env_is_map) return AM_INV_FOUR_PI;
    V3 v; V2 uv = env_dir_to_uv(s, d, &v);
    uv.x -= fdiv(.5f, (float) (s.env_res_x - 1u));
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    const float inv_sin_theta = rsqrt(max_(max_(sqr(v.x) + sqr(v.z), sqr(AM_EPSILON)), 0.f));
    return env_warp_eval(s, uv) * inv_sin_theta * AM_ENV_PDF_SCALE;
}

// Scene::sample_emitter_direction (src/render/scene.cpp:295-346) + sample_emitter (:223-246) + the emitter's sample_direction:
// AreaLight (src/emitters/area.cpp:116-167) + Shape::sample_direction (src/render/shape.cpp:360-377) + Mesh::sample_position
// (src/render/mesh.cpp:764-818); with GEN also PointLight (point.cpp:117-148), ConstantBackgroundEmitter (constant.cpp:120-150) and
// EnvironmentMapEmitter (envmap.cpp:419-462); then the visibility test.
// DEFER = true leaves the visibility test to the caller: *shadow_out is the shadow ray to trace when the returned
// ds.pdf != 0; an occluded sample then contributes nothing (the reference zeroes spec and ds.pdf, scene.cpp:333-342).
template <bool DEFER, bool GEN = false, int ACCEL = 0> AM_D DirectionSample sample_emitter_direction(const DScene &s, const SI &ref, V2 sample, V3 *spec_out,
                                              uint32_t *n_nodes, uint32_t *n_tris, Ray *shadow_out = nullptr) {
    DirectionSample ds; ds.p = v3(0.f); ds.n = v3(0.f); ds.d = v3(0.f); ds.pdf = 0.f; ds.dist = 0.f; ds.delta = false;
    *spec_out = v3(0.f);
    uint32_t ne = s.n_emitters; if (ne == 0) return ds;
    uint32_t index = 0; float emitter_weight = 1.f;
    if (ne >= 2) {
        float ecf = (float) ne, scaled = sample.x * ecf;
        index = min((uint32_t) scaled, ne - 1u);
        emitter_weight = ecf; sample.x = scaled - (float) index;
    }
    const DEmitter *em = s.emitters + index;
    V3 spec;
    if (!GEN || em->type == EMITTER_AREA) {
        const DMesh *m = s.meshes + em->mesh;
        const float *cdf = s.area_cdf + m->cdf_offset, *pmfp = s.area_pmf + m->cdf_offset;
        uint32_t face = distr_sample(cdf, m->nfaces, m->sum, sample.y);
        float pmf = __ldg(pmfp + face) * m->normalization, cdf0 = face > 0 ? __ldg(cdf + face - 1) * m->normalization : 0.f;
        sample.y = fdiv(sample.y - cdf0, pmf);
        uint32_t tri = __ldg(s.prim_to_tri + m->prim_offset + face);
        const float4 *tp = s.tris + 3 * (size_t) tri;
        float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
        V3 p0 = v3(a.x, a.y, a.z), e0 = v3(b.x, b.y, b.z) - p0, e1 = v3(c.x, c.y, c.z) - p0;
        V2 bc = square_to_uniform_triangle(sample);
        ds.p = fmadd(e0, bc.x, fmadd(e1, bc.y, p0));
        ds.pdf = m->normalization;
        if (m->has_normals) {
            const float4 *np = s.tri_nrm + 3 * (size_t) tri;
            float4 na = ldg4(np), nb = ldg4(np + 1), nc = ldg4(np + 2);
            ds.n = fmadd(v3(na.x, na.y, na.z), 1.f - bc.x - bc.y, fmadd(v3(nb.x, nb.y, nb.z), bc.x, v3(nc.x, nc.y, nc.z) * bc.y));
        } else ds.n = cross(e0, e1);
        ds.n = normalize(ds.n);
        if (m->flip_normals) ds.n = -ds.n;
        ds.d = ds.p - ref.p;
        float dist_squared = squared_norm(ds.d);
        ds.dist = fsqrt(dist_squared);
        ds.d = ds.d / ds.dist;
        float dp = fabs_(dot(ds.d, ds.n)), x = fdiv(dist_squared, dp);
        ds.pdf *= finite_(x) ? x : 0.f;
        bool active = dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
        spec = ld3(em->radiance) / ds.pdf;
        spec = select3(active, spec);
    } else if (em->type == EMITTER_POINT) {
        ds.p = ld3(em->position); ds.pdf = 1.f; ds.delta = true;
        ds.d = ds.p - ref.p;
        const float dist2 = squared_norm(ds.d), inv_dist = rsqrt(dist2);
        ds.dist = fsqrt(dist2); ds.d = ds.d * inv_dist;
        spec = ld3(em->radiance) * sqr(inv_dist);
    } else if (em->type == EMITTER_CONSTANT) {
        const V3 d = square_to_uniform_sphere(sample), center = ld3(s.bsphere);
        const float radius = max_(s.bsphere[3], norm(ref.p - center)), dist = 2.f * radius;
        ds.p = fmadd(d, dist, ref.p); ds.n = -d; ds.pdf = AM_INV_FOUR_PI; ds.d = d; ds.dist = dist;
        spec = ld3(em->radiance) / ds.pdf;
    } else {
        float pdf; V2 uv = env_warp_sample(s, sample, &pdf);
        uv.x += fdiv(.5f, (float) (s.env_res_x - 1u));
        const bool active = pdf > 0.f;
        const float theta = uv.y * AM_PI, phi = uv.x * (2.f * AM_PI);
        float st, ct, sp, cp; sincos_(theta, &st, &ct); sincos_(phi, &sp, &cp);
        V3 d = v3(cp * st, sp * st, ct);          // dr::sphdir
        d = v3(d.y, d.z, -d.x);
        const V3 center = ld3(s.bsphere);
        const float radius = max_(s.bsphere[3], norm(ref.p - center)), dist = 2.f * radius;
        const float inv_sin_theta = rsqrt(max_(max_(sqr(d.x) + sqr(d.z), sqr(AM_EPSILON)), 0.f));
        d = xf_vector(s.env_to_world, d);
        ds.p = ref.p + d * dist;
        ds.n = -d; ds.d = d; ds.dist = dist;
        ds.pdf = active ? pdf * inv_sin_theta * AM_ENV_PDF_SCALE : 0.f;
        spec = select3(active, env_eval_uv(s, uv) / ds.pdf);
    }
    ds.pdf *= rcp((float) ne);
    spec = spec * emitter_weight;
    if (ds.pdf != 0.f) {
        Ray sr = spawn_ray_to(ref.p, ref.n, ds.p);
        if (DEFER) *shadow_out = sr;
        else if (intersect<true, false, ACCEL>(s, sr, n_nodes, n_tris).tri != AM_NO_HIT) { spec = v3(0.f); ds.pdf = 0.f; }
    }
    *spec_out = spec;
    return ds;
}
// Scene::pdf_emitter_direction (scene.cpp:351-360) + AreaLight::pdf_direction (area.cpp:169-195) +
// Shape::pdf_direction (shape.cpp:379-390) + Mesh::pdf_position (mesh.cpp:846-849)
AM_D float pdf_emitter_direction(const DScene &s, const DMesh *em_mesh, V3 d, V3 n, float dist) {
    float dp = dot(d, n);
    float pdf = em_mesh->normalization, adp = fabs_(dp);
    pdf *= (adp != 0.f) ? fdiv(dist * dist, adp) : 0.f;
    float value = dp < 0.f ? pdf : 0.f;
    return value * (s.n_emitters == 0 ? 0.f : rcp((float) s.n_emitters));
}
// src/integrators/mvpath.h:306-311
AM_D float mis_weight(float a, float b) { a *= a; b *= b; float w = fdiv(a, a + b); return finite_(w) ? w : 0.f; }

} // namespace amvpt
