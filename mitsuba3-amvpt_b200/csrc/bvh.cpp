// bvh.cpp -- host-side builder of the device BVH (replaces the reference's three acceleration back ends:
// ShapeKDTree include/mitsuba/render/kdtree.h, Embree src/render/scene_embree.inl, OptiX src/render/scene_optix.inl).
// Binned-SAH top-down build over padded triangle boxes, flattened into 64-byte two-child nodes and 48-byte
// pre-gathered triangles in leaf order (layout: dtypes.h / DESIGN.md).  The tree only prunes: the hit that the
// traversal reports is decided by Mesh::moeller_trumbore and the global-index tie rule, so it equals the brute-force
// result as long as every box conservatively contains its triangles -- hence the padding below.
#include "host.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <functional>
#include <chrono>
#include <cstdio>
#include <atomic>
#include <thread>
#include <cstring>

namespace amvpt {

namespace {
struct Box {
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    void grow(const Box &b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void grow(const float *p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    float half_area() const { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return dx * dy + dy * dz + dz * dx; }
};
struct Prim { Box b; float c[3]; uint32_t prim; };
struct TNode { Box b; int32_t left = -1, right = -1; uint32_t first = 0, count = 0; };

// index-parallel loop over [0, n): the bodies below write disjoint elements, so the result does not depend on the thread count
static unsigned g_threads = 1;
template <class F> void parallel_for(size_t n, F body) {
    const unsigned nt = (unsigned) std::min<size_t>(g_threads, n / 16384 + 1);
    if (nt <= 1) { for (size_t i = 0; i < n; ++i) body(i); return; }
    std::atomic<size_t> next{ 0 }; const size_t chunk = std::max<size_t>(4096, n / (8 * nt));
    auto worker = [&]() { for (size_t b; (b = next.fetch_add(chunk)) < n;) for (size_t i = b, e = std::min(n, b + chunk); i < e; ++i) body(i); };
    std::vector<std::thread> pool; for (unsigned t = 1; t < nt; ++t) pool.emplace_back(worker);
    worker(); for (auto &t : pool) t.join();
}
const int kBins = 16, kSahDepth = 40;
// Leaf policy: at most kMaxLeaf triangles (<= 7 by the leaf encoding); a range that already fits becomes a leaf unless
// the best SAH split is cheaper (cost of one node step kCostNode vs. one triangle test kCostTri, measured on sm_100a:
// the IEEE-exact Moeller-Trumbore test costs about 1.5 box-pair steps).  AMVPT_MAX_LEAF / AMVPT_SAH_CT override for tuning.
static int env_int(const char *k, int d) { const char *v = getenv(k); return v ? atoi(v) : d; }
static float env_float(const char *k, float d) { const char *v = getenv(k); return v ? (float) atof(v) : d; }
static int kMaxLeaf = 4; static float kCostTri = 1.5f, kCostNode = 1.f;

struct Builder {
    std::vector<Prim> &prims; std::vector<TNode> nodes;
    explicit Builder(std::vector<Prim> &p, size_t expect = ~(size_t) 0) : prims(p) { nodes.reserve(2 * std::min(expect, p.size()) + 2); }
    int32_t build(uint32_t begin, uint32_t end, int depth) {
        TNode n; Box cb;
        for (uint32_t i = begin; i < end; ++i) { n.b.grow(prims[i].b); cb.grow(prims[i].c); }
        n.first = begin; n.count = end - begin;
        int32_t self = (int32_t) nodes.size(); nodes.push_back(n);
        if (end - begin <= 1) return self;
        bool fits = end - begin <= (uint32_t) kMaxLeaf;
        int axis = 0; float ext = cb.hi[0] - cb.lo[0];
        for (int a = 1; a < 3; ++a) if (cb.hi[a] - cb.lo[a] > ext) { ext = cb.hi[a] - cb.lo[a]; axis = a; }
        uint32_t mid = (begin + end) / 2; bool split = false;
        if (ext > 0.f && depth < kSahDepth) {
            struct Bin { Box b; uint32_t n = 0; } bins[kBins];
            float k = kBins * (1.f - 1e-6f) / ext, lo = cb.lo[axis];
            auto binof = [&](const Prim &p) { int b = (int) ((p.c[axis] - lo) * k); return std::min(std::max(b, 0), kBins - 1); };
            for (uint32_t i = begin; i < end; ++i) { Bin &b = bins[binof(prims[i])]; b.n++; b.b.grow(prims[i].b); }
            float la[kBins], ra[kBins]; uint32_t ln[kBins], rn[kBins];
            { Box acc; uint32_t c = 0; for (int i = 0; i < kBins; ++i) { c += bins[i].n; if (bins[i].n) acc.grow(bins[i].b); la[i] = c ? acc.half_area() : 0.f; ln[i] = c; } }
            { Box acc; uint32_t c = 0; for (int i = kBins - 1; i >= 0; --i) { c += bins[i].n; if (bins[i].n) acc.grow(bins[i].b); ra[i] = c ? acc.half_area() : 0.f; rn[i] = c; } }
            int best = -1; float best_cost = FLT_MAX;
            for (int i = 0; i < kBins - 1; ++i) {
                if (!ln[i] || !rn[i + 1]) continue;
                float c = la[i] * (float) ln[i] + ra[i + 1] * (float) rn[i + 1];
                if (c < best_cost) { best_cost = c; best = i; }
            }
            if (fits && (best < 0 || kCostTri * (float) (end - begin) * n.b.half_area() <= kCostNode * n.b.half_area() + kCostTri * best_cost)) return self;
            if (best >= 0) {
                uint32_t m = (uint32_t) (std::partition(prims.begin() + begin, prims.begin() + end, [&](const Prim &p) { return binof(p) <= best; }) - prims.begin());
                if (m != begin && m != end) { mid = m; split = true; }
            }
        }
        if (!split && fits) return self;
        if (!split) {   // object median along the widest axis: guarantees termination and bounded depth
            std::nth_element(prims.begin() + begin, prims.begin() + mid, prims.begin() + end,
                             [axis](const Prim &a, const Prim &b) { return a.c[axis] < b.c[axis] || (a.c[axis] == b.c[axis] && a.prim < b.prim); });
        }
        if (stop_after_split) { did_split = true; split_mid = mid; return self; }      // build_parallel: one step only
        int32_t l = build(begin, mid, depth + 1), r = build(mid, end, depth + 1);
        nodes[self].left = l; nodes[self].right = r;
        return self;
    }
    int depth_of(int32_t n) const { return nodes[n].left < 0 ? 1 : 1 + std::max(depth_of(nodes[n].left), depth_of(nodes[n].right)); }

    // The same tree, built on several host threads.  The top of the tree is built here down to ranges of <= grain primitives, whose
    // subtrees are deferred; the deferred ranges are disjoint slices of `prims`, so each is built independently by the very same
    // recursion (private node vector); finally everything is emitted in the pre-order the sequential recursion produces (node, whole
    // left subtree, whole right subtree).  Node numbering, leaf order and boxes are therefore identical to build(0, n, 0) -- tested.
    struct Deferred { uint32_t begin, end; int depth; std::vector<TNode> nodes; };
    void build_parallel(uint32_t n, unsigned n_threads) {
        const uint32_t grain = std::max<uint32_t>(n / (8 * n_threads), 4096);
        std::vector<Deferred> tasks;
        // top tree: children < 0 that are not leaves are encoded as -(2 + task index) while it is being built
        std::vector<TNode> top; std::vector<int32_t> top_left, top_right;
        std::function<int32_t(uint32_t, uint32_t, int)> top_build = [&](uint32_t begin, uint32_t end, int depth) -> int32_t {
            if (end - begin <= grain) { tasks.push_back({ begin, end, depth, {} }); return -(2 + (int32_t) tasks.size() - 1); }
            // one step of build(): reuse it on a scratch builder restricted to this node by temporarily limiting the recursion
            Builder one(prims, 1); one.stop_after_split = true;
            one.build(begin, end, depth);
            TNode nd = one.nodes[0]; const uint32_t mid = one.split_mid;
            int32_t self = (int32_t) top.size(); top.push_back(nd); top_left.push_back(-1); top_right.push_back(-1);
            if (!one.did_split) return self;                 // became a leaf
            int32_t l = top_build(begin, mid, depth + 1), r = top_build(mid, end, depth + 1);
            top_left[self] = l; top_right[self] = r;
            return self;
        };
        auto t0 = std::chrono::steady_clock::now();
        int32_t root = top_build(0, n, 0);
        auto t1 = std::chrono::steady_clock::now();
        std::atomic<size_t> next{ 0 };
        auto worker = [&]() { for (size_t i; (i = next.fetch_add(1)) < tasks.size();) { Builder b(prims, tasks[i].end - tasks[i].begin); b.build(tasks[i].begin, tasks[i].end, tasks[i].depth); tasks[i].nodes.swap(b.nodes); } };
        std::vector<std::thread> pool; for (unsigned t = 1; t < n_threads; ++t) pool.emplace_back(worker);
        worker(); for (auto &t : pool) t.join();
        auto t2 = std::chrono::steady_clock::now();
        nodes.clear();
        std::function<int32_t(int32_t)> emit = [&](int32_t ref) -> int32_t {
            const int32_t base = (int32_t) nodes.size();
            if (ref <= -2) {                                   // a deferred subtree: append with its indices shifted
                const std::vector<TNode> &sub = tasks[(size_t) (-(ref + 2))].nodes;
                for (const TNode &t : sub) { TNode c = t; if (c.left >= 0) { c.left += base; c.right += base; } nodes.push_back(c); }
                return base;
            }
            nodes.push_back(top[ref]);
            if (top_left[ref] != -1) { const int32_t l = emit(top_left[ref]), r = emit(top_right[ref]); nodes[base].left = l; nodes[base].right = r; }
            return base;
        };
        emit(root);
        if (getenv("AMVPT_BVH_TIMING")) fprintf(stderr, "[bvh] parallel build: top %.1f ms (%zu subtrees), subtrees on %u threads %.1f ms, merge %.1f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count(), tasks.size(), n_threads,
                                               std::chrono::duration<double, std::milli>(t2 - t1).count(), std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t2).count());
    }
    bool stop_after_split = false, did_split = false; uint32_t split_mid = 0;
};
} // namespace

void build_bvh(const Scene &s, float scene_scale, BvhOut &out) {
    uint32_t np = s.nprims();
    const bool timing = getenv("AMVPT_BVH_TIMING") != nullptr; auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) { if (!timing) return; auto t = std::chrono::steady_clock::now(); fprintf(stderr, "[bvh] %-22s %.1f ms\n", what, std::chrono::duration<double, std::milli>(t - t_last).count()); t_last = t; };
    std::vector<Prim> prims(np);
    std::vector<uint32_t> prim_mesh(np);
    // Padding: Moeller-Trumbore decides with a positional error of a few 2^-24 * (|origin| + |vertex| + t); boxes get
    // an absolute margin tied to the scene scale plus a relative one, far above that error and far below feature size.
    const float pad_abs = 2e-6f * scene_scale + 1e-30f;
    uint32_t g = 0;
    g_threads = (unsigned) std::max(1, env_int("AMVPT_BVH_THREADS", (int) std::min(16u, std::max(1u, std::thread::hardware_concurrency()))));
    for (size_t mi = 0; mi < s.meshes.size(); ++mi) {
        const HMesh &m = s.meshes[mi]; const uint32_t g0 = g; g += m.nfaces();
        parallel_for(m.nfaces(), [&](size_t f) {
            const uint32_t g = g0 + (uint32_t) f;
            Prim &p = prims[g]; p.prim = g; prim_mesh[g] = (uint32_t) mi;
            Box b; for (int k = 0; k < 3; ++k) b.grow(&m.pos[3 * m.faces[3 * f + k]]);
            for (int a = 0; a < 3; ++a) {
                p.c[a] = 0.5f * (b.lo[a] + b.hi[a]);
                float pad = pad_abs + 1e-5f * std::max(std::fabs(b.lo[a]), std::fabs(b.hi[a])) + 1e-5f * (b.hi[a] - b.lo[a]);
                b.lo[a] -= pad; b.hi[a] += pad;
            }
            p.b = b;
        });
    }
    kMaxLeaf = std::min(std::max(env_int("AMVPT_MAX_LEAF", 4), 1), 7); kCostTri = env_float("AMVPT_SAH_CT", 1.5f);
    lap("primitive boxes");
    Builder bl(prims);
    // AMVPT_BVH_THREADS: host threads of the build (default: up to 16 of the machine's; 1 = the plain recursion); the tree does not depend on it
    unsigned n_threads = g_threads;
    if (np >= 65536 && n_threads > 1) bl.build_parallel(np, n_threads);
    else if (np) bl.build(0, np, 0);
    out.depth = np ? (uint32_t) bl.depth_of(0) : 1u;
    {   // flat leaf list for small scenes (DScene::flat): leaves in tree (pre-)order
        std::vector<float4> fl; bool ok = np > 0;
        for (const TNode &n : bl.nodes) if (n.left < 0) {
            if (fl.size() >= 64) { ok = false; break; }
            int32_t code = (int32_t) ~((n.first << 3) | n.count); float cf; memcpy(&cf, &code, 4);
            fl.push_back(make_float4(n.b.lo[0], n.b.lo[1], n.b.lo[2], cf)); fl.push_back(make_float4(n.b.hi[0], n.b.hi[1], n.b.hi[2], 0.f));
        }
        if (ok && env_int("AMVPT_FLAT", 1)) out.flat.swap(fl);
    }
    if (AM_BVH_WIDTH == 2 && out.depth > AM_STACK_SIZE - 2) AM_THROW("BVH depth %u exceeds the traversal stack", out.depth);

    lap("tree");
    // triangles in leaf order
    out.tris.resize(3 * (size_t) np); out.prim_to_tri.resize(np);
    bool any_nrm = false; for (auto &m : s.meshes) any_nrm |= !m.nrm.empty();
    out.tri_nrm.assign(any_nrm ? 3 * (size_t) np : 0, make_float4(0.f, 0.f, 0.f, 0.f));
    parallel_for(np, [&](size_t i) {
        uint32_t prim = prims[i].prim, mi = prim_mesh[prim]; const HMesh &m = s.meshes[mi]; uint32_t f = prim - m.prim_offset;
        out.prim_to_tri[prim] = (uint32_t) i;
        for (int k = 0; k < 3; ++k) {
            const float *p = &m.pos[3 * m.faces[3 * f + k]];
            uint32_t w = k == 0 ? prim : (k == 1 ? mi : (uint32_t) s.bsdfs[m.bsdf].type); float wf; memcpy(&wf, &w, 4);   // third word: BSDF plugin type (wavefront sort key)
            out.tris[3 * (size_t) i + k] = make_float4(p[0], p[1], p[2], wf);
            if (!m.nrm.empty()) { const float *n = &m.nrm[3 * m.faces[3 * f + k]]; out.tri_nrm[3 * (size_t) i + k] = make_float4(n[0], n[1], n[2], 0.f); }
        }
    });
    auto leaf_code = [&](const TNode &n) -> int32_t { return (int32_t) ~((n.first << 3) | n.count); };
    lap("leaf-order triangles");
    // ---- compressed 4-wide tree over the same leaves (DScene::nodes_q4, the wavefront trace stage): the binary tree is collapsed (a
    // node's two children, then repeatedly the inner child with the largest surface is replaced by its own two children until four
    // slots are used), and the child boxes are quantised to 8 bits per plane on a power-of-two grid anchored at the node's own box
    // corner -- lower planes rounded down, upper planes rounded up, checked against the decoded value, so the boxes only grow.
    out.nodes_q4.clear(); out.depth_q4 = 0;
    if (np && bl.nodes[0].left >= 0) {
        struct Wide { int32_t kid[4]; int n; };
        std::vector<Wide> wide; std::vector<int32_t> wide_of(bl.nodes.size(), -1); uint32_t depth4 = 0;
        std::function<void(int32_t, uint32_t)> build_wide = [&](int32_t t, uint32_t depth) {
            depth4 = std::max(depth4, depth + 1);
            const int32_t id = (int32_t) wide.size(); wide.push_back(Wide{ { -1, -1, -1, -1 }, 0 }); wide_of[t] = id;
            int32_t kids[4]; int n = 0; kids[n++] = bl.nodes[t].left; kids[n++] = bl.nodes[t].right;
            while (n < 4) {
                int best = -1; float area = -1.f;
                for (int i = 0; i < n; ++i) if (bl.nodes[kids[i]].left >= 0 && bl.nodes[kids[i]].b.half_area() > area) { area = bl.nodes[kids[i]].b.half_area(); best = i; }
                if (best < 0) break;
                const int32_t k = kids[best]; kids[best] = bl.nodes[k].left; kids[n++] = bl.nodes[k].right;
            }
            Wide w; w.n = n; for (int i = 0; i < 4; ++i) w.kid[i] = i < n ? kids[i] : -1;
            wide[(size_t) id] = w;
            for (int i = 0; i < n; ++i) if (bl.nodes[kids[i]].left >= 0) build_wide(kids[i], depth + 1);
        };
        build_wide(0, 0);
        if (3 * depth4 + 2 <= AM_STACK_SIZE) {
            out.depth_q4 = depth4; out.nodes_q4.assign(4 * wide.size(), make_uint4(0u, 0u, 0u, 0u));
            parallel_for(wide.size(), [&](size_t id) {
                const Wide &w = wide[id]; Box nb; for (int i = 0; i < w.n; ++i) nb.grow(bl.nodes[w.kid[i]].b);
                uint32_t e[3], qlo[3][4], qhi[3][4]; int32_t ref[4] = { 0, 0, 0, 0 };
                for (int a = 0; a < 3; ++a) {
                    // smallest power of two with 255 * scale >= extent (one more step if rounding makes the last plane fall short)
                    const float ext = nb.hi[a] - nb.lo[a]; int ex = 0; std::frexp(ext / 255.f, &ex);     // ext / 255 = m * 2^ex, m in [0.5, 1)
                    if (!(ext > 0.f)) ex = -120;
                    ex = std::min(std::max(ex, -120), 120);
                    for (;;) {
                        const float scale = std::ldexp(1.f, ex); bool ok = true;
                        for (int i = 0; i < 4 && ok; ++i) {
                            if (i >= w.n) { qlo[a][i] = 255u; qhi[a][i] = 0u; continue; }      // empty slot: inverted box, never hit
                            const Box &kb = bl.nodes[w.kid[i]].b;
                            float fl = std::floor((kb.lo[a] - nb.lo[a]) / scale), fh = std::ceil((kb.hi[a] - nb.lo[a]) / scale);
                            fl = std::min(std::max(fl, 0.f), 255.f); fh = std::min(std::max(fh, 0.f), 255.f);
                            while (fl > 0.f && std::fma(fl, scale, nb.lo[a]) > kb.lo[a]) fl -= 1.f;
                            while (fh < 255.f && std::fma(fh, scale, nb.lo[a]) < kb.hi[a]) fh += 1.f;
                            if (std::fma(fl, scale, nb.lo[a]) > kb.lo[a] || std::fma(fh, scale, nb.lo[a]) < kb.hi[a]) ok = false;
                            qlo[a][i] = (uint32_t) fl; qhi[a][i] = (uint32_t) fh;
                        }
                        if (ok) break;
                        ++ex;
                    }
                    e[a] = (uint32_t) (ex + 127);
                }
                for (int i = 0; i < w.n; ++i) { const TNode &k = bl.nodes[w.kid[i]]; ref[i] = k.left < 0 ? leaf_code(k) : wide_of[w.kid[i]]; }
                auto pack = [](const uint32_t *q) { return q[0] | (q[1] << 8) | (q[2] << 16) | (q[3] << 24); };
                uint32_t ob[3]; memcpy(ob, nb.lo, 12);
                uint4 *o = &out.nodes_q4[4 * id];
                o[0] = make_uint4(ob[0], ob[1], ob[2], e[0] | (e[1] << 8) | (e[2] << 16));
                o[1] = make_uint4(pack(qlo[0]), pack(qlo[1]), pack(qlo[2]), pack(qhi[0]));
                o[2] = make_uint4(pack(qhi[1]), pack(qhi[2]), (uint32_t) ref[0], (uint32_t) ref[1]);
                o[3] = make_uint4((uint32_t) ref[2], (uint32_t) ref[3], 0u, 0u);
            });
        }
    }
#if AM_BVH_WIDTH == 4
    // Collapse the binary tree into 4-wide nodes: start from a node's two children and keep replacing the inner child with
    // the largest surface by its own two children until four slots are used (or only leaves remain).  Nodes are numbered in
    // pre-order, so the root is node 0 and a child reference 0 can mark an empty slot.
    struct Wide { int32_t kid[4]; int n; };
    std::vector<Wide> wide; uint32_t depth4 = 0;
    std::function<int32_t(int32_t, uint32_t)> widen = [&](int32_t t, uint32_t depth) -> int32_t {
        depth4 = std::max(depth4, depth + 1);
        int32_t id = (int32_t) wide.size(); wide.push_back(Wide{ { -1, -1, -1, -1 }, 0 });
        int32_t kids[4]; int n = 0;
        if (bl.nodes[t].left < 0) kids[n++] = t;             // a tree that is a single leaf: one slot
        else { kids[n++] = bl.nodes[t].left; kids[n++] = bl.nodes[t].right; }
        while (n < 4) {
            int best = -1; float area = -1.f;
            for (int i = 0; i < n; ++i) if (bl.nodes[kids[i]].left >= 0 && bl.nodes[kids[i]].b.half_area() > area) { area = bl.nodes[kids[i]].b.half_area(); best = i; }
            if (best < 0) break;
            int32_t k = kids[best]; kids[best] = bl.nodes[k].left; kids[n++] = bl.nodes[k].right;
        }
        Wide w; w.n = n; for (int i = 0; i < 4; ++i) w.kid[i] = i < n ? kids[i] : -1;
        wide[id] = w;
        return id;
    };
    // breadth of recursion = tree depth (<= kSahDepth + log2 of the largest median-split range), fine for the host stack
    std::vector<int32_t> wide_of(bl.nodes.size(), -1);
    std::function<void(int32_t, uint32_t)> build_wide = [&](int32_t t, uint32_t depth) {
        int32_t id = widen(t, depth); wide_of[t] = id;
        Wide w = wide[id];
        for (int i = 0; i < w.n; ++i) if (bl.nodes[w.kid[i]].left >= 0) build_wide(w.kid[i], depth + 1);
    };
    if (np) build_wide(0, 0); else wide.push_back(Wide{ { -1, -1, -1, -1 }, 0 });
    out.depth = std::max(depth4, 1u);
    if (3 * out.depth + 2 > AM_STACK_SIZE) AM_THROW("BVH depth %u exceeds the traversal stack", out.depth);
    out.nodes.assign(8 * wide.size(), make_float4(0.f, 0.f, 0.f, 0.f));
    for (size_t id = 0; id < wide.size(); ++id) {
        const Wide &w = wide[id]; float lo[3][4], hi[3][4]; int32_t ref[4];
        for (int i = 0; i < 4; ++i) {
            for (int a = 0; a < 3; ++a) { lo[a][i] = 0.f; hi[a][i] = 0.f; }
            ref[i] = 0;
            if (i >= w.n) continue;
            const TNode &k = bl.nodes[w.kid[i]];
            for (int a = 0; a < 3; ++a) { lo[a][i] = k.b.lo[a]; hi[a][i] = k.b.hi[a]; }
            ref[i] = k.left < 0 ? leaf_code(k) : wide_of[w.kid[i]];
        }
        float4 *o = &out.nodes[8 * id];
        for (int a = 0; a < 3; ++a) { o[2 * a] = make_float4(lo[a][0], lo[a][1], lo[a][2], lo[a][3]); o[2 * a + 1] = make_float4(hi[a][0], hi[a][1], hi[a][2], hi[a][3]); }
        float rf[4]; memcpy(rf, ref, 16); o[6] = make_float4(rf[0], rf[1], rf[2], rf[3]);
    }
#else
    lap("compressed wide tree");
    // flatten: inner nodes numbered in depth-first order, each holding its two children's boxes
    auto child_ref = [&](const std::vector<int32_t> &inner_id, int32_t c) -> int32_t {
        const TNode &n = bl.nodes[c];
        if (n.left < 0) return leaf_code(n);
        return inner_id[c];
    };
    std::vector<int32_t> inner_id(bl.nodes.size(), -1); int32_t n_inner = 0;
    for (size_t i = 0; i < bl.nodes.size(); ++i) if (bl.nodes[i].left >= 0) inner_id[i] = n_inner++;
    auto put = [&](int32_t id, const Box &a, int32_t ca, const Box &b, int32_t cb) {
        float fa, fb; memcpy(&fa, &ca, 4); memcpy(&fb, &cb, 4);
#if AM_FFMA2       // (lo.x, lo.y, hi.x, hi.y): the packed slab test multiplies register pairs by (id.x, id.y)
        out.nodes[4 * (size_t) id + 0] = make_float4(a.lo[0], a.lo[1], a.hi[0], a.hi[1]);
        out.nodes[4 * (size_t) id + 1] = make_float4(b.lo[0], b.lo[1], b.hi[0], b.hi[1]);
#else
        out.nodes[4 * (size_t) id + 0] = make_float4(a.lo[0], a.hi[0], a.lo[1], a.hi[1]);
        out.nodes[4 * (size_t) id + 1] = make_float4(b.lo[0], b.hi[0], b.lo[1], b.hi[1]);
#endif
        out.nodes[4 * (size_t) id + 2] = make_float4(a.lo[2], a.hi[2], b.lo[2], b.hi[2]);
        out.nodes[4 * (size_t) id + 3] = make_float4(fa, fb, 0.f, 0.f);
    };
    Box empty; for (int a = 0; a < 3; ++a) { empty.lo[a] = 0.f; empty.hi[a] = 0.f; }
    const int32_t empty_leaf = (int32_t) ~0u;     // leaf with count 0
    if (n_inner == 0) {
        // zero or one leaf: synthesise a root so that node 0 is always an inner node
        out.nodes.resize(4);
        if (np) put(0, bl.nodes[0].b, child_ref(inner_id, 0), empty, empty_leaf); else put(0, empty, empty_leaf, empty, empty_leaf);
        return;
    }
    out.nodes.resize(4 * (size_t) n_inner);
    for (size_t i = 0; i < bl.nodes.size(); ++i) {
        const TNode &n = bl.nodes[i]; if (n.left < 0) continue;
        put(inner_id[i], bl.nodes[n.left].b, child_ref(inner_id, n.left), bl.nodes[n.right].b, child_ref(inner_id, n.right));
    }
#endif
}

} // namespace amvpt
