/* amvpt.h -- C ABI of the B200-native AMVPT render path (libamvpt_b200.so).
 *
 * The reference (xacond00/mitsuba3-amvpt, Mitsuba 3.6.4 + the `mvpath` integrator) has no C-level operator
 * interface: plugins export only `plugin_name()` / `plugin_descr()` (include/mitsuba/core/class.h:225-230) and
 * everything else is C++ ABI templated on <Float, Spectrum>.  The entry points below are what an FFI for this path
 * binds; each one names the reference interface it replaces.  INTEGRATION.md shows the C++ plugin shim a maintainer
 * would add (an `Integrator` subclass whose render() forwards to amvpt_render()) and the ctypes binding used by the
 * tests.  Conventions: plain pointers and sizes, caller-owned output buffers, int return codes (0 = ok, negative =
 * error; amvpt_last_error() returns the thread-local message, mirroring the reference's Throw(...) strings), one
 * render at a time per scene handle.  All matrices are row-major.  No CPU fallback exists: every compute entry point
 * fails with AMVPT_ERR_CUDA when no sm_100a device is usable.
 */
#ifndef AMVPT_H
#define AMVPT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define AMVPT_OK 0
#define AMVPT_ERR_INVALID (-1)   /* bad argument / scene description error (reference: Throw -> std::runtime_error) */
#define AMVPT_ERR_CUDA (-2)      /* CUDA runtime failure or no device */
#define AMVPT_ERR_IO (-3)        /* file not found / parse error */

typedef struct amvpt_scene amvpt_scene;
typedef struct amvpt_comm amvpt_comm;      /* one NCCL communicator of this process (one process per GPU) */

/* BSDF plugin parameters (src/bsdfs/{diffuse,conductor,dielectric,roughconductor,roughdielectric,roughplastic}.cpp
 * constructors; `twosided` = wrapped in src/bsdfs/twosided.cpp).  type: 0 diffuse, 1 conductor, 2 dielectric,
 * 3 roughconductor, 4 roughdielectric, 5 roughplastic.  distr: 0 beckmann, 1 ggx. */
typedef struct amvpt_bsdf_desc {
    int32_t type, twosided, distr, nonlinear;
    float reflectance[3], eta[3], k[3], specular_reflectance[3], specular_transmittance[3];
    float int_ior, ext_ior, alpha_u, alpha_v;
} amvpt_bsdf_desc;

/* `grid` multi-sensor over a perspective / thinlens sub-camera (src/sensors/grid.cpp:84-236,
 * src/sensors/perspective.cpp:139-181, src/sensors/thinlens.cpp:152-187).  res_x / res_y = quilt size.
 * fov_axis: 0 x, 1 y, 2 diagonal, 3 smaller, 4 larger (src/render/sensor.cpp:177-231). */
typedef struct amvpt_grid_sensor_desc {
    int32_t thinlens, fov_axis; uint32_t grid_x, grid_y, res_x, res_y; int32_t reverse_x, reverse_y;
    int32_t use_cone, cam_center;
    double to_world[16], fov;
    float near_clip, far_clip, focus_distance, aperture_radius, cone_deg, cam_dist;
    float cam_dir[3], cam_off[3];
} amvpt_grid_sensor_desc;

/* `mvpath` integrator parameters (src/integrators/mvpath.h:120-129; MonteCarloIntegrator
 * src/render/integrator.cpp:503-520) + independent sampler seed / sample_count (src/samplers/independent.cpp). */
typedef struct amvpt_integrator_desc {
    int64_t max_depth; uint32_t rr_depth; int32_t hide_emitters; uint32_t spp_pass_lim;
    uint32_t sampler_seed, sample_count;
} amvpt_integrator_desc;
/* Integrator `timeout` (seconds, <= 0 = none; src/render/integrator.cpp:22-28, should_stop() integrator.h:303-306): checked between passes */
int amvpt_scene_set_timeout(struct amvpt_scene *s, float seconds);

/* Integrator::render(scene, sensor, seed, spp, develop, evaluate) arguments (include/mitsuba/render/integrator.h:77-96)
 * plus the build-defined extensions (view sharding, scheduler choice, adaptive allocation). */
typedef struct amvpt_render_desc {
    uint32_t seed, spp;
    int32_t scheduler;             /* 0 warp batches, 1 persistent + path regeneration, 2 stream (unified traversal loop, ballot-compacted refill),
                                    * 3 wavefront (regen / trace / shade-per-BSDF-type stages over a pool of path slots),
                                    * 4 auto (default): wavefront for scenes with >= 3 BSDF plugin types or >= 32768 triangles, else 1 */
    int32_t collect_stats;         /* bit 0: count rays / BVH nodes / triangle tests (slower); bit 1: time the wavefront trace stage */
    uint32_t rank, world;          /* this process renders the view tiles v with v mod world == rank */
    int32_t adaptive; float adaptive_threshold; uint32_t adaptive_min_spp;
    void *stream;                  /* cudaStream_t or NULL */
    void (*allreduce)(float *dev, uint32_t n, void *user); void *allreduce_user;   /* stats all-reduce hook (any collective library); must complete before it returns */
    struct amvpt_comm *comm;       /* or: an amvpt_comm -- the library all-reduces the statistics itself (ncclAllReduce on `stream`) */
} amvpt_render_desc;

typedef struct amvpt_render_stats {
    double seconds;                /* device time of the pass loop (CUDA events) */
    uint64_t paths, closest_rays, shadow_rays, bvh_nodes, tri_tests, path_vertices;
    uint32_t passes, launches, spp_done, scheduler_used;   /* scheduler_used: what `auto` resolved to */
    double trace_seconds;          /* wavefront only, if timed: device time of the trace stage kernels */
    uint32_t trace_launches, pad;
} amvpt_render_stats;

const char *amvpt_last_error(void);
int amvpt_device_count(void);

/* ---- scene construction: Scene / Shape / BSDF / Emitter / Sensor plugin constructors --------------------------- */
int amvpt_scene_create(amvpt_scene **out);
void amvpt_scene_destroy(amvpt_scene *s);
/* xml::load_file (src/core/xml.cpp) for the subset of SURVEY.md sec.8b; defines = "key=value" strings (-D) */
int amvpt_scene_from_xml(const char *path, const char *const *defines, int n_defines, amvpt_scene **out);
int amvpt_scene_from_xml_string(const char *xml, const char *base_dir, const char *const *defines, int n_defines, amvpt_scene **out);
int amvpt_scene_add_bsdf(amvpt_scene *s, const amvpt_bsdf_desc *d);                     /* returns BSDF index */
/* Mesh (src/render/mesh.cpp) with optional vertex normals and optional area emitter radiance (src/emitters/area.cpp) */
int amvpt_scene_add_mesh(amvpt_scene *s, const float *positions, uint32_t n_vertices, const uint32_t *faces, uint32_t n_faces,
                         const float *normals, int bsdf, const float *radiance, int flip_normals);   /* returns mesh index */
/* emitters that are not attached to a shape, appended to the scene's emitter list in call order like Scene::Scene does
 * (src/render/scene.cpp:22-63); each returns the emitter index.  PointLight (src/emitters/point.cpp:65-85), ConstantBackgroundEmitter
 * (src/emitters/constant.cpp:55-68), EnvironmentMapEmitter (src/emitters/envmap.cpp:111-214: rgb = the bitmap as linear RGB float32,
 * h rows of w pixels, top row first; to_world16 may be NULL = identity).  At most one environment emitter (constant or envmap) */
int amvpt_scene_add_emitter_point(amvpt_scene *s, const float *position3, const float *intensity3);
int amvpt_scene_add_emitter_constant(amvpt_scene *s, const float *radiance3);
int amvpt_scene_add_emitter_envmap(amvpt_scene *s, const float *rgb, uint32_t w, uint32_t h, float scale, const float *to_world16);
int amvpt_scene_set_sensor_grid(amvpt_scene *s, const amvpt_grid_sensor_desc *d);
/* `batch` multi-sensor (src/sensors/batch.cpp:94-180): n explicit perspective sub-cameras, film = n*res_x by res_y */
int amvpt_scene_set_sensor_batch(amvpt_scene *s, uint32_t n, uint32_t res_x, uint32_t res_y, const float *sample_to_camera,
                                 const float *to_world, float near_clip, float far_clip, int reverse_x);
/* any MultiSensor as the table of its sub-cameras in sensor-index order (MultiSensor::sensors() / grid_dim() / reverse_axis(), src/sensors/grid.cpp:
 * 435-457): grid_x * grid_y records of sample_to_camera (16 floats), to_world (16 floats) and { near_clip, far_clip, aperture_radius,
 * focus_distance }; the film is grid_x * view_res_x by grid_y * view_res_y.  This is what a plugin shim inside the reference calls
 * (integration/mvpath_b200.cpp) */
int amvpt_scene_set_sensor_table(amvpt_scene *s, uint32_t grid_x, uint32_t grid_y, uint32_t view_res_x, uint32_t view_res_y, int reverse_x, int reverse_y,
                                 int thinlens, const float *sample_to_camera, const float *to_world, const float *scalars4);
int amvpt_scene_set_integrator(amvpt_scene *s, const amvpt_integrator_desc *d);
/* cross-view sample reuse of the mvpath integrator (src/integrators/mvpath.h:120-129: sa_reuse, sa_mis, fast_mis, reuse_count,
 * adaptive; render_multisample / sample_multi, mvpath_multi.h:8-358).  Reuse is active when sa_reuse, more than one sub-sensor
 * and reuse_count != 1; *n_samples receives the number of cameras per group after the divisor rule of mvpath.cpp:192-218
 * (0 = off).  `adaptive` = number of fill-in paths for samples that no other camera could reuse (mvpath_multi.h:52-58, 79-115; single rank);
 * amvpt_scene_set_reuse_debug switches the `debug` output (the adapt mask instead of the image, mvpath_multi.h:54-56) */
int amvpt_scene_set_reuse(amvpt_scene *s, int sa_reuse, int sa_mis, int fast_mis, uint32_t reuse_count, uint32_t adaptive, uint32_t *n_samples);
int amvpt_scene_set_reuse_debug(amvpt_scene *s, int debug);
/* number of fill-in paths the last amvpt_render traced (the reference logs it as "Adaptive samples", mvpath_multi.h:89) */
int amvpt_last_fill_paths(amvpt_scene *s, uint64_t *n);
/* reconstruction filter of the film (src/rfilters/{box,gaussian,tent,mitchell,catmullrom,lanczos}.cpp; ImageBlock::put,
 * src/render/imageblock.cpp:174-558): type 0 box (default of this API), 1 gaussian (param = stddev, Mitsuba's default filter with
 * stddev 0.5), 2 tent (param = radius), 3 mitchell (param = B, param2 = C; both default to 1/3 in the reference), 4 catmullrom,
 * 5 lanczos (param = lobes, a positive integer; reference default 3).  param2 is ignored by every type but 3. */
int amvpt_scene_set_rfilter(amvpt_scene *s, int type, float param, float param2);
/* builds the BVH and uploads everything to `device` (Scene::Scene accel_init, src/render/scene.cpp:70-73) */
int amvpt_scene_commit(amvpt_scene *s, int device);
int amvpt_scene_info(amvpt_scene *s, uint32_t *film_w, uint32_t *film_h, uint32_t *n_views, uint32_t *n_prims,
                     uint32_t *n_bvh_nodes, uint32_t *sample_count);
int amvpt_scene_get_sensor(amvpt_scene *s, uint32_t index, float *sample_to_camera16, float *to_world16);
/* scalars4 = { near_clip, far_clip, aperture_radius, focus_distance } of sub-sensor `index`; layout5 = { grid_x, grid_y, reverse_x, reverse_y, thinlens } */
int amvpt_scene_get_sensor_params(amvpt_scene *s, uint32_t index, float *scalars4, int32_t *layout5);
int amvpt_scene_get_bsdf_tables(amvpt_scene *s, int bsdf, float *ext_transmittance64, float *scalars3);
/* read-back of the host scene description (what the XML loader or the add_* calls produced; works before commit).
 * amvpt_scene_get_mesh: pass NULL buffers to query sizes first.  amvpt_scene_get_bsdf returns int_ior = int_ior / ext_ior, ext_ior = 1. */
int amvpt_scene_counts(amvpt_scene *s, uint32_t *n_meshes, uint32_t *n_bsdfs, uint32_t *n_emitters);
int amvpt_scene_get_mesh(amvpt_scene *s, uint32_t index, uint32_t *n_vertices, uint32_t *n_faces, int32_t *bsdf, int32_t *has_normals,
                         int32_t *flip_normals, int32_t *is_emitter, float *radiance3, float *positions, uint32_t *faces, float *normals);
int amvpt_scene_get_bsdf(amvpt_scene *s, int bsdf, amvpt_bsdf_desc *d);
/* emitter `index` of the scene's emitter list: type 0 area (mesh = its shape), 1 point, 2 constant, 3 envmap */
int amvpt_scene_get_emitter(amvpt_scene *s, uint32_t index, int32_t *type, int32_t *mesh, float *radiance3, float *position3);
/* the environment map as loaded (rgb may be NULL to query the size), its scale and to_world; *w = 0 if the scene has none */
int amvpt_scene_get_envmap(amvpt_scene *s, uint32_t *w, uint32_t *h, float *scale, float *to_world16, float *rgb);
/* sensor flag that changes the sampler stream: 1 if every sample draws a shutter time (shutter_close > shutter_open, mvpath_single.h:66-68) */
int amvpt_scene_get_shutter_draw(amvpt_scene *s, int32_t *draw);
int amvpt_scene_set_shutter(amvpt_scene *s, float shutter_open, float shutter_close);
int amvpt_scene_get_integrator(amvpt_scene *s, amvpt_integrator_desc *d, int32_t *is_mvpath, int32_t *sa_reuse, uint32_t *reuse_count);

/* parity tap of the reuse branch: for lanes [lane_begin, lane_begin + n) the n_samples records of render_multisample, 9 floats
 * each = { pos.x, pos.y, weight, result.rgb, valid, indirect, sensor index } (mirrors oracle orc_trace_lanes_multi) */
int amvpt_trace_lanes_multi(amvpt_scene *s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, float *out9);

/* ---- the hot path ----------------------------------------------------------------------------------------------- */
/* MVPathIntegrator::render (src/integrators/mvpath.cpp:7-278): accumulates into the scene's film (RGBW, float32) */
int amvpt_render(amvpt_scene *s, const amvpt_render_desc *d, amvpt_render_stats *stats_out);
/* rectangles of the quilt that `rank` of `world` renders (view sharding, build-defined): fills up to `cap`
 * records of 5 uint32 {x0, y0, w, h, view} and returns the number of rectangles */
int amvpt_shard_tiles(amvpt_scene *s, uint32_t rank, uint32_t world, uint32_t *out5, uint32_t cap);
/* the rectangle of every view (sub-sensor) in the quilt, ordered by view index = ix + grid_x * iy with the tile position following
 * GridSensor::sample_ray's axis reversal (src/sensors/grid.cpp:246-255): same 5-uint32 records, returns their number.  This is
 * the per-view split amvpt_film_write_pfm / _exr(view) cut by */
int amvpt_view_tiles(amvpt_scene *s, uint32_t *out5, uint32_t cap);
/* ---- multi-GPU hand-off (build-defined; the reference is single-device).  NCCL is bound at run time.
 * amvpt_comm_unique_id: ncclGetUniqueId on one rank (128 bytes, hand them to the other ranks by any means: MPI, a file, torch.distributed);
 * amvpt_comm_init: ncclCommInitRank on `device`; amvpt_film_gather: the root's film receives every other rank's view tiles (each rank
 * sends only what it rendered, 1 / world of the quilt, plus the border pixels where a jittered sample can round up into the next pixel; with sample reuse or a wide reconstruction filter the whole quilt is sum-reduced);
 * enqueued on `stream` (NULL = default stream), returns without synchronising */
int amvpt_comm_unique_id(uint8_t *id128);
int amvpt_comm_init(const uint8_t *id128, uint32_t rank, uint32_t world, int device, amvpt_comm **out);
void amvpt_comm_destroy(amvpt_comm *c);
int amvpt_film_gather(amvpt_scene *s, amvpt_comm *c, uint32_t root, void *stream);
int amvpt_comm_allreduce_sum(amvpt_comm *c, float *dev, uint32_t n, void *stream);
/* the rectangles amvpt_film_gather moves for `rank` of `world` (host logic, no device needed): the rank's view tiles and 1-pixel-wide runs of
 * the foreign pixels just right of / below them (every pixel at most once), records of 5 uint32 {x0, y0, w, h, pixel offset in the staging
 * buffer}; returns their number */
int amvpt_gather_tiles(amvpt_scene *s, uint32_t rank, uint32_t world, uint32_t *out5, uint32_t cap);
/* the adaptive allocator's decisions of the last amvpt_render(adaptive = 1): out[pass * n_views + view] = samples per pixel (lane
 * indices k < that number of the pass's lane map were traced); pass out = NULL to query the sizes.  Identical on every rank */
int amvpt_adaptive_alloc(amvpt_scene *s, uint32_t *out, uint32_t cap, uint32_t *n_passes, uint32_t *n_views);
int amvpt_cancel(amvpt_scene *s);                                        /* Integrator::cancel (integrator.h:303-306) */
int amvpt_film_clear(amvpt_scene *s);                                    /* Film::prepare / clear (hdrfilm.cpp:245-273) */
int amvpt_film_raw(amvpt_scene *s, float *rgbw_host);                    /* film storage, H*W*4 floats to host */
int amvpt_film_develop(amvpt_scene *s, float *rgb_host);                 /* HDRFilm::develop (hdrfilm.cpp:321-408), H*W*3 to host */
int amvpt_film_develop_device(amvpt_scene *s, float *rgb_dev, void *stream);
/* develop into a page-locked host buffer owned by the library (one per process, grown on demand, reused by every call and every scene):
 * *rgb_host stays valid until the next call; the copy runs at PCIe rate instead of through a pageable bounce buffer */
int amvpt_film_develop_pinned(amvpt_scene *s, float **rgb_host);
int amvpt_film_device_ptr(amvpt_scene *s, void **rgbw_dev, void **moments_dev);   /* for the NCCL film gather */
int amvpt_film_add_host(amvpt_scene *s, const float *rgbw_host);         /* ImageBlock::put_block (imageblock.cpp:135-172) */

/* ---- output path and display step --------------------------------------------------- */
/* HDRFilm::write with file_format = pfm (hdrfilm.cpp:420-560 -> Bitmap::write_pfm, src/core/bitmap.cpp:2270-2299): the developed
 * quilt (view < 0) or one view's tile of it (the per-view split of the reference's to_gif.py), RGB float32, bottom row first */
int amvpt_film_write_pfm(amvpt_scene *s, const char *path, int32_t view);
/* HDRFilm::write with file_format = openexr (hdrfilm.cpp:420-560 -> Bitmap::write_exr, src/core/bitmap.cpp:1332-1440): same
 * image and view selection as above, channels R G B, half_float != 0 stores float16 (hdrfilm's default component_format,
 * hdrfilm.cpp:146-147, 191-192; round to nearest even), 0 stores float32.  The OpenEXR library is not in this tree, so the file is
 * written directly as a single-part scan-line OpenEXR 2 file with NO_COMPRESSION (the reference compresses with PIZ,
 * bitmap.cpp:1350): any EXR reader decodes both to the same pixel values */
int amvpt_film_write_exr(amvpt_scene *s, const char *path, int32_t view, int32_t half_float);
/* the component conversion amvpt_film_write_exr applies (host code, no device needed): float32 -> float16 bits, round to nearest even */
int amvpt_f32_to_f16(const float *src, uint16_t *dst, uint64_t n);
/* what the viewer does to a rendered quilt when it loads it: Bitmap::convert(PixelFormat::RGB, Struct::Type::UInt8, srgb_gamma)
 * (src/mitsuba/program.h:55; per component src/core/struct.cpp:1645-1662 + dr::linear_to_srgb).  The 8-bit quilt stays on the
 * device for amvpt_interlace; rgb8_host (H*W*3) may be NULL */
int amvpt_film_to_srgb8(amvpt_scene *s, uint8_t *rgb8_host);
/* lenticular display calibration (src/mitsuba/preset.h:12-17) + the viewer toggles "Display quilt" / "Nearest scaling"
 * (src/mitsuba/program.cpp:130-132) */
typedef struct amvpt_preset { float center, focus, pitch, tilt, subp, view[2]; int32_t grid[2]; int32_t flip_rgb, quilt, nearest; } amvpt_preset;
/* Program::display_image (src/mitsuba/program.cpp:199-272): quilt -> interlaced image for a slanted-lens display, dst_ch = 3
 * (the screenshot path; pass flip_rgb = 1 as the reference does) or 4 (the window surface, alpha = 255).  src_host = NULL uses the
 * scene's own 8-bit quilt; otherwise any src_h * src_w * src_ch image (src_ch >= 3).  The window itself is unavailable on this
 * box: dst_host (dst_h * dst_w * dst_ch bytes) is the headless surface */
int amvpt_interlace(amvpt_scene *s, const uint8_t *src_host, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const amvpt_preset *p,
                    uint32_t dst_w, uint32_t dst_h, uint32_t dst_ch, uint8_t *dst_host);
int amvpt_display_last_ms(amvpt_scene *s, float *srgb8_ms, float *interlace_ms);   /* device time of the last two kernels above */

/* ---- parity taps (used by tests/; same meaning as the oracle's orc_* entry points) ------------------------------ */
int amvpt_trace_lanes(amvpt_scene *s, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, int trace,
                      float *sample_pos, float *ray_o, float *ray_d, float *ray_maxt, uint32_t *view, float *radiance,
                      uint32_t *pixel, uint32_t max_hits, uint32_t *hit_prim, float *hit_t, uint64_t *stats2);
/* any_hit: bit 0 = Scene::ray_test semantics (first accepted triangle), bit 1 = walk the compressed 4-wide tree of the wavefront trace stage,
 * bit 2 = no traversal counters (stats2 stays 0): scenes with at most 32 leaves then take the flat leaf list the render kernels use */
int amvpt_intersect(amvpt_scene *s, const float *o, const float *d, const float *maxt, uint32_t n, int any_hit,
                    uint32_t *prim, float *tuv, uint64_t *stats2);
/* amvpt_intersect without any allocation (csrc/noalloc.cu).  No counterpart in the reference; added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same results
 * bit for bit.  Rays, results and counters are staged in ONE file-scope device array of AMVPT_STATIC_MAX_RAYS x 44 B + 64 B = 2.75 MiB that
 * every GPU loading the library holds until exit; n > AMVPT_STATIC_MAX_RAYS returns AMVPT_ERR_INVALID.  All calls share that array: calls
 * that overlap on the GPU (two host threads, two streams) race -- serialise them. */
#define AMVPT_STATIC_MAX_RAYS 65536u
int amvpt_intersect_static(amvpt_scene *s, const float *o, const float *d, const float *maxt, uint32_t n, int any_hit,
                           uint32_t *prim, float *tuv, uint64_t *stats2);
int amvpt_bsdf_eval_pdf_sample(amvpt_scene *s, int bsdf, const float *wi, const float *wo, const float *s1, const float *s2, uint32_t n,
                               float *f, float *pdf, float *bs_wo, float *bs_pdf, float *bs_eta, uint32_t *bs_type, float *weight);
int amvpt_math1(int device, int op, const float *in, uint32_t n, float *out);
/* Hierarchical2D<Float, 0>::sample / eval of the scene's environment map on the device (include/mitsuba/core/distr_2d.h:490-575, 668-698):
 * u = n x 2 uniform numbers -> uv (n x 2), pdf (n), and eval(uv) (n) */
int amvpt_env_warp(amvpt_scene *s, const float *u, uint32_t n, float *uv, float *pdf, float *eval_back);
/* host-side readers of the scene loader, exposed for the tests: Bitmap(path) -> linear RGB float32 (OpenEXR scan-line none / PIZ, PFM;
 * rgb may be NULL to query the size), and <spectrum value="wavelength:value ..."> -> linear sRGB (src/core/spectrum.cpp:85-130 behind
 * src/core/xml.cpp:1170-1268; n = 1 is the uniform case) */
int amvpt_image_read(const char *path, uint32_t *w, uint32_t *h, float *rgb);
int amvpt_spectrum_to_srgb(const double *wavelengths, const double *values, uint32_t n, int within_emitter, int unbounded_name, double *rgb3);
int amvpt_sampler_stream(int device, uint32_t seed_value, uint32_t lane, uint32_t n, float *out);
/* memory-system micro-benchmark used by bench.py for the L2 line of the roofline ("L2 bandwidth to be measured by
 * a builder micro-benchmark"): kind 0 = every thread streams 16-byte ld.global.cg loads over a zero-filled buffer of `bytes`,
 * `iters` times after a warm-up sweep; returns GB/s.  bytes well below the 126 MB L2 -> L2 read bandwidth, several L2 sizes -> HBM */
int amvpt_membench(int device, int kind, uint64_t bytes, uint32_t iters, double *gb_per_s);
/* host-side checks of the BVH the kernels traverse: every triangle inside its leaf box, every child box inside its parent */
int amvpt_bvh_validate(amvpt_scene *s, uint32_t *max_depth, uint32_t *n_leaves);
/* the same for the compressed 4-wide tree of the wavefront trace stage: every triangle inside the decoded box of every slot above it;
 * *n_nodes = 0 when the scene has no such tree (a single leaf) */
/* FNV-1a over the flattened tree as it is uploaded (nodes | leaf-ordered triangles | compressed nodes): the build is deterministic and does not
 * depend on AMVPT_BVH_THREADS (tests/test_cpu_host.py) */
int amvpt_bvh_checksum(amvpt_scene *s, uint64_t *checksum);
int amvpt_bvh_q4_validate(amvpt_scene *s, uint32_t *n_nodes, uint32_t *max_depth, uint32_t *n_leaves);

#ifdef __cplusplus
}
#endif
#endif
