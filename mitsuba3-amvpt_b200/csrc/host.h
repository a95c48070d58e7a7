// host.h -- host-side scene of the AMVPT path: the objects the reference's plugins would hold (meshes, BSDF
// parameters, the grid / batch multi-sensor, HDR film, independent sampler, mvpath integrator parameters), their
// device mirrors, and the render entry point that stands in for MVPathIntegrator::render().
#pragma once
#include "kernels.cuh"
#include "htransform.h"
#include <stdexcept>
#include <string>
#include <vector>

namespace amvpt {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };
#define AM_THROW(...) throw ::amvpt::Error(::amvpt::strfmt(__VA_ARGS__))
std::string strfmt(const char *fmt, ...);
void cuda_check(cudaError_t e, const char *what);
#define AM_CUDA(x) ::amvpt::cuda_check((x), #x)

// Mesh (src/render/mesh.cpp): positions / normals / faces + area distribution (mesh.cpp:444-486)
struct HMesh {
    std::string id;
    std::vector<float> pos, nrm;          // 3 floats per vertex; nrm empty = no vertex normals
    std::vector<uint32_t> faces;          // 3 indices per face
    int bsdf = 0, emitter = -1, flip_normals = 0;
    bool had_uv = false;                  // the file carried texture coordinates (dropped: no textures on this path; see the anisotropy guard in xml.cpp)
    uint32_t prim_offset = 0;
    std::vector<float> pmf, cdf; float sum = 0.f, normalization = 0.f;
    uint32_t nfaces() const { return (uint32_t) (faces.size() / 3); }
    void build_area_distr();
};

// src/integrators/mvpath.h:120-129 + MonteCarloIntegrator (src/render/integrator.cpp:503-520)
struct IntegratorParams {
    int64_t max_depth = -1; uint32_t rr_depth = 5; bool hide_emitters = false;
    uint32_t spp_pass_lim = 16; bool sa_reuse = false, sa_mis = false, fast_mis = false, force_eval = true, debug = false;
    uint32_t reuse_count = 1, adaptive = 0, block_size = 0; float timeout = -1.f;
    bool is_mvpath = true;
};
// GridSensor parameters (src/sensors/grid.cpp:84-180) over a perspective / thinlens sub-camera
struct GridSensorParams {
    bool thinlens = false; int fov_axis = 0;   // 0 x, 1 y, 2 diagonal, 3 smaller, 4 larger
    uint32_t grid_x = 1, grid_y = 1, res_x = 0, res_y = 0; bool reverse_x = false, reverse_y = true;
    bool use_cone = false, cam_center = false;
    bool single = false;   // a plain perspective / thinlens sensor (one view, to_world used as given)
    M4d to_world = m4d_identity();
    double fov = 45.0; bool has_focal_length = false; std::string focal_length = "50mm";
    float near_clip = 1e-2f, far_clip = 1e4f, focus_distance = 1e4f, aperture_radius = 0.f, cone_deg = 0.f, cam_dist = 0.1f;
    float cam_dir[3] = { 1.f, 0.f, 0.f }, cam_off[3] = { 0.f, 0.f, 0.f };
};
// EnvironmentMapEmitter (src/emitters/envmap.cpp:111-214): the bitmap as loaded, the radiance array with its periodic extra column and
// the luminance * sin(theta) hierarchy that importance-samples it (Hierarchical2D<Float, 0>, include/mitsuba/core/distr_2d.h:376-482)
struct EnvMap {
    bool present = false; uint32_t src_w = 0, src_h = 0; std::vector<float> src_rgb;
    uint32_t res_x = 0, res_y = 0; std::vector<float> data;
    float scale = 1.f; M4 to_world = m4_identity(), from_world = m4_identity();
    std::vector<float> warp; std::vector<uint32_t> level_off, level_width;
    float patch_size[2] = { 0, 0 }, inv_patch_size[2] = { 0, 0 }; uint32_t max_patch[2] = { 0, 0 };
    void build(const float *rgb, uint32_t w, uint32_t h, float scale, const M4d &to_world);
};
// xml.cpp create_texture_from_spectrum (src/core/xml.cpp:1170-1268) in an RGB variant: wavelength:value pairs -> linear sRGB
// (spectrum_list_to_srgb, src/core/spectrum.cpp:85-130), a uniform value inside an emitter -> value * xyz_to_srgb(1, 1, 1)
void spectrum_to_srgb(std::vector<double> wavelengths, std::vector<double> values, bool within_emitter, bool unbounded_name, double *rgb);
void uniform_emitter_spectrum_to_srgb(double value, double *rgb);
void spectrum_list_to_srgb(const std::vector<double> &wavelengths, const std::vector<double> &values, bool bounded, bool d65, double *rgb);
// imageio.cpp: Bitmap(path) -> convert(RGB, Float32, linear), top row first
void load_image_rgb(const std::string &path, uint32_t *w, uint32_t *h, std::vector<float> &rgb);

enum SchedulerMode { SCHED_WARP_BATCH = 0, SCHED_REGENERATE = 1, SCHED_STREAM = 2, SCHED_WAVEFRONT = 3, SCHED_AUTO = 4, SCHED_REUSE = 5 /* reported only */, SCHED_SPLIT = 6 };

struct BvhOut { std::vector<float4> nodes, tris, tri_nrm; std::vector<uint32_t> prim_to_tri; uint32_t depth = 0;
                std::vector<uint4> nodes_q4; uint32_t depth_q4 = 0;
                std::vector<float4> flat; };     // flat: the leaves' boxes of a tree with at most 32 leaves, two float4 each (DScene::flat), else empty     // nodes_q4: compressed 4-wide tree over the same leaves (empty = not built)
struct RenderStats {
    double seconds = 0.0;            // device time of the pass loop (CUDA events)
    unsigned long long paths = 0, closest_rays = 0, shadow_rays = 0, nodes = 0, tris = 0, vertices = 0;
    uint32_t passes = 0, launches = 0, spp_done = 0;
    double trace_seconds = 0.0; uint32_t trace_launches = 0, scheduler_used = 0;   // wavefront: device time of the trace stage (if timed)
};
struct Shard { uint32_t rank = 0, world = 1; };   // view tiles owned: {v : v mod world == rank}

struct Scene {
    // ---- host description
    std::vector<HMesh> meshes; std::vector<DBsdf> bsdfs; std::vector<DEmitter> emitters; std::vector<DSensor> sensors;
    std::vector<std::string> bsdf_ids;
    uint32_t grid_x = 1, grid_y = 1, res_x = 0, res_y = 0; bool reverse_x = false, reverse_y = true, thinlens = false;
    bool multi_sensor = false, shutter_draw = false;   // shutter_draw: shutter_close > shutter_open (Sensor ctor, src/render/sensor.cpp:18-22)
    uint32_t film_w = 0, film_h = 0;
    uint32_t sample_count = 4, sampler_seed = 0;
    // film reconstruction filter: 0 box, 1 gaussian (stddev), 2 tent (radius), 3 mitchell (B, C), 4 catmullrom, 5 lanczos (lobes)
    int rfilter_type = 0; float rfilter_param = 0.5f, rfilter_param2 = 0.f;
    void set_rfilter(int type, float param, float param2 = 0.f);
    IntegratorParams integrator;
    // samples per primary sample of the reuse branch (MVPT::render, mvpath.cpp:17-22 + 192-218), 0 = reuse off
    uint32_t reuse_samples() const;
    void set_reuse(bool sa_reuse, bool sa_mis, bool fast_mis, uint32_t reuse_count, uint32_t adaptive);
    // ---- device mirror
    int device = -1, sm_count = 148;
    bool committed = false;
    DScene d{};
    std::vector<void *> allocs;
    float4 *film = nullptr; float2 *moments = nullptr; float *rgb = nullptr;
    unsigned long long *counters = nullptr; TileDesc *tiles_dev = nullptr; float *view_stats_dev = nullptr;
    uint32_t tiles_cap = 0;
    uint8_t *srgb8 = nullptr;                              // display step: the developed quilt as 8-bit sRGB (allocated on first use)
    float display_ms[2] = { 0.f, 0.f };                    // device time of the last srgb8 / interlace kernel
    float4 *split_buf = nullptr; uint32_t split_cap = 0;   // "split" scheduler: first-vertex records of one chunk
    uint8_t *adapt_mask = nullptr; uint32_t *adapt_scratch = nullptr, *adapt_compact = nullptr; unsigned long long adapt_cap = 0; size_t adapt_compact_cap = 0;   // reuse `adaptive` fill-in
    unsigned long long last_fill_paths = 0;
    std::vector<uint32_t> alloc_log;      // adaptive allocator: samples per pixel given to view v in pass p of the last render, [pass][view]
    float4 *reuse_buf = nullptr; uint32_t reuse_cap = 0;   // staged sample-reuse schedule: per-chunk hit + projection records
    WfPool wf{}; uint32_t wf_type_mask = 0;     // wavefront scheduler pool (allocated on first use)
    uint32_t n_nodes = 0, n_tris = 0;
    char *bvh_dev = nullptr; size_t bvh_bytes = 0, bvh_hot_bytes = 0, l2_persist_max = 0, l2_window_max = 0;   // contiguous nodes | triangles | compressed nodes; hot = the first two
    volatile bool stop = false;
    RenderStats last;
    // host copy of the tree the kernels traverse (built by prepare_host, checked by amvpt_bvh_validate)
    BvhOut bvh; bool host_ready = false;

    ~Scene();
    int add_bsdf(const DBsdf &b, const std::string &id = "");
    int add_mesh(HMesh &&m, const float *radiance);
    // scene-level emitters (src/emitters/{point,constant,envmap}.cpp); each returns the emitter index
    EnvMap env; int env_emitter = -1;
    int add_emitter_point(const float *position, const float *intensity);
    int add_emitter_constant(const float *radiance);
    int add_emitter_envmap(const float *rgb, uint32_t w, uint32_t h, float scale, const M4d &to_world);
    bool general_emitters() const { for (auto &e : emitters) if (e.type != EMITTER_AREA) return true; return false; }
    void bounding_sphere(float *center3, float *radius) const;
    void set_sensor_grid(const GridSensorParams &g);
    void set_sensor_explicit(uint32_t n, uint32_t res_x, uint32_t res_y, const float *s2c, const float *to_world,
                             float near_clip, float far_clip, bool reverse_x);
    void set_sensor_table(uint32_t grid_x, uint32_t grid_y, uint32_t res_x, uint32_t res_y, bool reverse_x, bool reverse_y, bool thinlens,
                          const float *s2c, const float *to_world, const float *scalars4);
    void prepare_host();
    void commit(int device);
    void film_clear(cudaStream_t st);
    void *dalloc(size_t bytes); void release_device_blocks(); std::vector<size_t> alloc_bytes;      // sizes of `allocs` (device block cache, scene.cpp)
    template <class T> T *upload(const std::vector<T> &v) {
        T *p = (T *) dalloc(std::max<size_t>(v.size(), 1) * sizeof(T));
        if (!v.empty()) AM_CUDA(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        return p;
    }
    uint32_t nprims() const { uint32_t n = 0; for (auto &m : meshes) n += m.nfaces(); return n; }
    std::vector<TileDesc> view_tiles(const Shard &sh) const;
};

// roughplastic.cpp:207-252 parameters_changed(): fills the derived members of a DBsdf record
void bsdf_prepare(DBsdf &b);

// bvh.cpp
void build_bvh(const Scene &s, float scene_scale, BvhOut &out);

// render.cpp -- MVPathIntegrator::render() (src/integrators/mvpath.cpp:7-278), wavefront branch
struct RenderOptions {
    uint32_t seed = 0, spp = 0; int scheduler = SCHED_AUTO; bool stats = false, time_stages = false; uint32_t regen_min = 0, trace_min = 0;
    Shard shard; cudaStream_t stream = nullptr;
    // adaptive allocator (build-defined): 0 = fixed spp; otherwise spp is the mean budget per pixel
    bool adaptive = false; float adaptive_threshold = 0.f; uint32_t adaptive_min_spp = 16;
    // optional collective hook: sums `n` floats across ranks in place (NCCL all-reduce in the multi-GPU driver)
    void (*allreduce)(float *dev, uint32_t n, void *user) = nullptr; void *allreduce_user = nullptr;
    void *comm = nullptr;            // amvpt_comm: the statistics all-reduce goes through NCCL on the render stream (takes precedence over the hook)
};
void render(Scene &s, const RenderOptions &opt);
// comm.cpp: NCCL behind the C ABI
void comm_allreduce_sum(void *comm, float *dev, uint32_t n, cudaStream_t st);
void film_gather(Scene &s, void *comm, uint32_t root, cudaStream_t st);
std::vector<TileDesc> rank_tiles(const Scene &s, const Shard &sh, bool per_view);

// xml.cpp
Scene *load_xml_file(const std::string &path, const std::vector<std::pair<std::string, std::string>> &defines);
Scene *load_xml_string(const std::string &xml, const std::string &base_dir, const std::vector<std::pair<std::string, std::string>> &defines);
// meshio.cpp
void load_obj(const std::string &path, const M4d &to_world, bool face_normals, bool flip_tex, HMesh &out);
void load_ply(const std::string &path, const M4d &to_world, bool face_normals, HMesh &out);
void load_serialized(const std::string &path, int shape_index, const M4d &to_world, bool face_normals, HMesh &out);
void make_cube(const M4d &to_world, HMesh &out);

} // namespace amvpt
