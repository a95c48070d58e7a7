// dtypes.h -- plain-old-data records shared by the host library and the device code (HBM layout).
#pragma once
#include <cuda_runtime.h>
#include "dbsdf.cuh"

#ifndef AM_BVH_WIDTH
#define AM_BVH_WIDTH 2     /* children per BVH node: 2 (default) or 4 (build with -DAM_BVH_WIDTH=4; measured slower on B200); host builder and kernels share the value */
#endif
#ifndef AM_FFMA2
#define AM_FFMA2 0         /* 1: binary-tree nodes hold (lo.x, lo.y, hi.x, hi.y) and the slab test uses Blackwell's packed fma.rn.f32x2 (FFMA2) */
#endif
#ifndef AM_FLAT_LEAVES
#define AM_FLAT_LEAVES 1   /* 1: intersect() uses DScene::flat when the scene has one */
#endif
#ifndef AM_SPECULATIVE
#define AM_SPECULATIVE 0   /* 1: intersect() postpones the first leaf a lane finds and keeps descending until it holds two (dscene.cuh) */
#endif
#ifndef AM_PREFETCH_FAR
#define AM_PREFETCH_FAR 0  /* 1: prefetch.global.L1 of the postponed child's node record when both children are hit */
#endif
#define AM_STACK_SIZE (AM_BVH_WIDTH == 4 ? 96 : 64)   /* per-thread traversal stack entries; the builder checks the tree against it */
#define AM_NODE_F4 (AM_BVH_WIDTH == 4 ? 8 : 4)         /* float4 per node */

namespace amvpt {

// 128 B per sub-camera.  PerspectiveCamera / ThinLensCamera members that sample_ray reads
// (src/sensors/perspective.cpp:205-241, src/sensors/thinlens.cpp:220-257).
struct DSensor {
    float s2c[16];        // sample_to_camera.matrix, row-major
    float to_world[12];   // first three rows of to_world.matrix
    float near_clip, far_clip, aperture_radius, focus_distance;
};
// 128 B per sub-camera, read only by the sample-reuse branch (Perspective/ThinLens::sample_surface,
// src/sensors/perspective.cpp:327-385, thinlens.cpp:358-418): camera_to_sample, inverse of to_world, 1 / image-rectangle area
struct DSensorAux {
    float c2s[16];
    float w2c[12];
    float normalization, pad[3];
};
struct DMesh {
    int32_t bsdf, emitter, flip_normals, has_normals;
    uint32_t prim_offset, nfaces;
    float normalization, sum;      // 1 / area and unnormalised area (DiscreteDistribution, distr_1d.h:233-264)
    uint32_t cdf_offset, pad[3];   // into DScene::area_cdf / area_pmf
};
// One record per scene emitter in the order Scene::Scene collects them (src/render/scene.cpp:22-63): area lights when their shape is added,
// the others where they stand in the scene description.  type: 0 area (src/emitters/area.cpp; `mesh` = its shape), 1 point
// (src/emitters/point.cpp; `radiance` = intensity, `position`), 2 constant environment (src/emitters/constant.cpp), 3 envmap
// (src/emitters/envmap.cpp; pixels and sampling hierarchy live in DScene::env_*)
enum EmitterType : int32_t { EMITTER_AREA = 0, EMITTER_POINT = 1, EMITTER_CONSTANT = 2, EMITTER_ENVMAP = 3 };
struct DEmitter { int32_t mesh; float radiance[3]; int32_t type; float position[3]; };

struct DScene {
    // AM_BVH_WIDTH == 4: 128 B per node = 8 x float4: lo.x[4], hi.x[4], lo.y[4], hi.y[4], lo.z[4], hi.z[4], child refs[4], pad;
    //   child ref 0 = empty slot (the root, node 0, is never a child), > 0 inner node, < 0 leaf as below.
    // AM_BVH_WIDTH == 2: 64 B per node = 4 x float4 (Aila-Laine layout):
    //   n0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)  n1 = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
    //   n2 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)  n3 = (bits c0, bits c1, -, -)
    // child >= 0: inner node index; child < 0: leaf, ~child = (first_tri << 3) | count (count <= 7)
    const float4 *nodes;
    // compressed 4-wide tree over the same leaves (bvh.cpp), 64 B per node = 4 x uint4, used by the wavefront trace stage; nullptr = not built:
    //   q0 = (bits origin.x, bits origin.y, bits origin.z, ex | ey << 8 | ez << 16)      plane = origin + q * 2^(e - 127)
    //   q1 = (qlo.x[4], qlo.y[4], qlo.z[4], qhi.x[4])   q2 = (qhi.y[4], qhi.z[4], ref0, ref1)   q3 = (ref2, ref3, -, -)
    //   one byte per child and plane; ref 0 = empty slot, > 0 inner node, < 0 leaf code as below
    const uint4 *nodes_q4;
    // triangles in BVH leaf order, 48 B each = 3 x float4: (p0, bits prim), (p1, bits mesh), (p2, 0)
    const float4 *tris;
    const float4 *tri_nrm;      // vertex normals in the same order (3 x float4), or nullptr
    const uint32_t *prim_to_tri; // global primitive index -> position in `tris`
    const DMesh *meshes;
    const DBsdf *bsdfs;
    const DEmitter *emitters;
    const DSensor *sensors;
    const float *area_cdf, *area_pmf;
    uint32_t n_tris, n_nodes, n_emitters, n_sensors;
    uint32_t grid_x, grid_y;
    int32_t reverse_x, reverse_y, thinlens, shutter_draw;   // shutter_draw: the sensor's shutter interval is non-empty, every sample draws a time (mvpath_single.h:66-68)
    uint32_t film_w, film_h;
    uint32_t max_depth, rr_depth;
    // film reconstruction filter: 0 box, 1 gaussian (polynomial of src/rfilters/gaussian.cpp:52-103), 2 tent (tent.cpp:48-56),
    // 3 mitchell (coeff[0..6] = a3 a2 a0 b3 b2 b1 b0, mitchell.cpp:63-69), 4 catmullrom, 5 lanczos (radius = lobes)
    int32_t filter_type; float filter_radius, filter_inv_radius; float filter_coeff[10];
    // cross-view sample reuse (mvpath.h:120-129): n_reuse = samples per primary sample after mvpath.cpp:192-218, 0 = off
    const DSensorAux *sensor_aux;
    uint32_t n_reuse; int32_t sa_mis, fast_mis, reuse_debug;
    uint32_t n_adapt; float adapt_w; uint32_t pad5[2];      // fill-in samples per masked lane = min(adaptive, n_reuse - 1), 1 / (n_adapt + 1) (mvpath_multi.h:52-53)
    // emitters other than area lights.  gen_emitters != 0 selects the general emitter code (kernels are instantiated for both cases, so
    // that scenes lit by area lights only keep the short path); env_emitter = index of the environment emitter in `emitters` or -1
    float box_lo[3], box_inv[3], pad4[2];     // bounding box of the geometry and 16 / extent per axis (ray-task sort of the wavefront scheduler)
    int32_t gen_emitters, env_emitter, hide_emitters, env_is_map;
    float bsphere[4];                 // centre, radius of the scene as Emitter::set_scene sees it (constant.cpp:73-88, envmap.cpp:290-304)
    const float *env_data;            // envmap: res_y x res_x x 3 floats, last column = first (envmap.cpp:146-207)
    const float *env_warp;            // Hierarchical2D<Float, 0> levels, concatenated (include/mitsuba/core/distr_2d.h:376-482)
    uint32_t env_res_x, env_res_y, env_n_levels, pad2;
    uint32_t env_level_off[12], env_level_width[12];
    float env_scale, env_patch_size[2], env_inv_patch_size[2]; uint32_t env_max_patch[2], pad3;
    float env_to_world[12], env_from_world[12];      // first three rows of to_world and of its inverse
    // Small scenes (a tree with at most 32 leaves, e.g. the Cornell box: 19): the leaves' boxes as a flat list, (lo.xyz, leaf code), (hi.xyz, -)
    // in tree order.  intersect() tests ALL of them in one loop without divergence, then visits the leaves whose box was hit (dscene.cuh).
    const float4 *flat; uint32_t n_flat;
};

} // namespace amvpt
