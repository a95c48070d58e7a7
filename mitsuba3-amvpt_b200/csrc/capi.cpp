// capi.cpp -- extern "C" boundary of libamvpt_b200.so (declared in include/amvpt.h).
#include "../../include/amvpt.h"
#include "host.h"
#include <cstring>
#include <functional>
#include <cfloat>
#include <cmath>

using namespace amvpt;

struct amvpt_scene { Scene *s; };
static thread_local std::string g_err;

static int guard(const std::function<int()> &fn) {
    try { return fn(); }
    catch (const Error &e) {
        g_err = e.what();
        return g_err.rfind("CUDA error", 0) == 0 ? AMVPT_ERR_CUDA : (g_err.rfind("io:", 0) == 0 ? AMVPT_ERR_IO : AMVPT_ERR_INVALID);
    }
    catch (const std::exception &e) { g_err = e.what(); return AMVPT_ERR_INVALID; }
}
#define NEED(x) do { if (!(x)) AM_THROW("%s: null or invalid argument (%s)", __func__, #x); } while (0)

// scratch device buffer that frees itself
struct DBuf {
    void *p = nullptr;
    explicit DBuf(size_t bytes) { AM_CUDA(cudaMalloc(&p, bytes ? bytes : 16)); }
    ~DBuf() { cudaFree(p); }
    template <class T> T *as() { return (T *) p; }
    void up(const void *h, size_t bytes) { if (bytes) AM_CUDA(cudaMemcpy(p, h, bytes, cudaMemcpyHostToDevice)); }
    void down(void *h, size_t bytes) { if (bytes) AM_CUDA(cudaMemcpy(h, p, bytes, cudaMemcpyDeviceToHost)); }
};
static void need_device(int dev) {
    int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) AM_THROW("CUDA error: no CUDA device available (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (dev < 0 || dev >= n) AM_THROW("CUDA error: device %d out of range (%d devices)", dev, n);
    AM_CUDA(cudaSetDevice(dev));
}
static Scene &committed(amvpt_scene *h) {
    if (!h || !h->s) AM_THROW("null scene handle");
    if (!h->s->committed) AM_THROW("scene not committed (call amvpt_scene_commit first)");
    AM_CUDA(cudaSetDevice(h->s->device));
    return *h->s;
}
static std::vector<std::pair<std::string, std::string>> parse_defines(const char *const *defines, int n) {
    std::vector<std::pair<std::string, std::string>> out;
    for (int i = 0; i < n; ++i) {
        std::string d = defines[i]; size_t eq = d.find('=');
        if (eq == std::string::npos) AM_THROW("define '%s' is not of the form key=value", d.c_str());
        out.emplace_back(d.substr(0, eq), d.substr(eq + 1));
    }
    return out;
}

extern "C" int amvpt_guard_call(const std::function<int()> &fn) { return guard(fn); }      // comm.cpp shares the exception -> error code mapping

extern "C" {

const char *amvpt_last_error(void) { return g_err.c_str(); }
int amvpt_device_count(void) { int n = 0; return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0; }

int amvpt_scene_create(amvpt_scene **out) {
    return guard([&] { NEED(out); *out = new amvpt_scene{ new Scene() }; return AMVPT_OK; });
}
void amvpt_scene_destroy(amvpt_scene *s) { if (s) { delete s->s; delete s; } }
int amvpt_scene_from_xml(const char *path, const char *const *defines, int n_defines, amvpt_scene **out) {
    return guard([&] { NEED(path); NEED(out); *out = new amvpt_scene{ load_xml_file(path, parse_defines(defines, n_defines)) }; return AMVPT_OK; });
}
int amvpt_scene_from_xml_string(const char *xml, const char *base_dir, const char *const *defines, int n_defines, amvpt_scene **out) {
    return guard([&] { NEED(xml); NEED(out);
        *out = new amvpt_scene{ load_xml_string(xml, base_dir ? base_dir : ".", parse_defines(defines, n_defines)) }; return AMVPT_OK; });
}
int amvpt_scene_add_bsdf(amvpt_scene *h, const amvpt_bsdf_desc *d) {
    return guard([&] { NEED(h); NEED(d);
        if (d->type < 0 || d->type > 5) AM_THROW("unknown BSDF type %d", d->type);
        DBsdf b{}; b.type = d->type; b.twosided = d->twosided; b.distr = d->distr; b.nonlinear = d->nonlinear;
        for (int i = 0; i < 3; ++i) { b.reflectance[i] = d->reflectance[i]; b.eta[i] = d->eta[i]; b.k[i] = d->k[i];
            b.spec_refl[i] = d->specular_reflectance[i]; b.spec_trans[i] = d->specular_transmittance[i]; }
        b.ior_eta = d->int_ior / d->ext_ior; b.alpha_u = d->alpha_u; b.alpha_v = d->alpha_v;
        return h->s->add_bsdf(b); });
}
int amvpt_scene_add_mesh(amvpt_scene *h, const float *pos, uint32_t nv, const uint32_t *faces, uint32_t nf, const float *normals,
                         int bsdf, const float *radiance, int flip_normals) {
    return guard([&] { NEED(h); NEED(pos || !nv); NEED(faces || !nf);
        HMesh m; m.pos.assign(pos, pos + 3 * (size_t) nv); m.faces.assign(faces, faces + 3 * (size_t) nf);
        if (normals) m.nrm.assign(normals, normals + 3 * (size_t) nv);
        m.bsdf = bsdf; m.flip_normals = flip_normals;
        return h->s->add_mesh(std::move(m), radiance); });
}
int amvpt_scene_add_emitter_point(amvpt_scene *h, const float *position, const float *intensity) {
    return guard([&] { NEED(h); NEED(position); NEED(intensity); return h->s->add_emitter_point(position, intensity); });
}
int amvpt_scene_add_emitter_constant(amvpt_scene *h, const float *radiance) {
    return guard([&] { NEED(h); NEED(radiance); return h->s->add_emitter_constant(radiance); });
}
int amvpt_scene_add_emitter_envmap(amvpt_scene *h, const float *rgb, uint32_t w, uint32_t hh, float scale, const float *to_world16) {
    return guard([&] { NEED(h); NEED(rgb); M4d tw = m4d_identity();
        if (to_world16) for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) tw.m[i][j] = to_world16[4 * i + j];
        return h->s->add_emitter_envmap(rgb, w, hh, scale, tw); });
}
int amvpt_scene_get_emitter(amvpt_scene *h, uint32_t index, int32_t *type, int32_t *mesh, float *radiance3, float *position3) {
    return guard([&] { NEED(h); Scene &s = *h->s; if (index >= s.emitters.size()) AM_THROW("emitter index %u out of range", index);
        const DEmitter &e = s.emitters[index]; if (type) *type = e.type; if (mesh) *mesh = e.mesh;
        for (int k = 0; k < 3; ++k) { if (radiance3) radiance3[k] = e.radiance[k]; if (position3) position3[k] = e.position[k]; }
        return AMVPT_OK; });
}
int amvpt_scene_get_envmap(amvpt_scene *h, uint32_t *w, uint32_t *hh, float *scale, float *to_world16, float *rgb) {
    return guard([&] { NEED(h); NEED(w); NEED(hh); Scene &s = *h->s;
        if (!s.env.present) { *w = *hh = 0; return AMVPT_OK; }
        *w = s.env.src_w; *hh = s.env.src_h; if (scale) *scale = s.env.scale;
        if (to_world16) std::memcpy(to_world16, s.env.to_world.m, 64);
        if (rgb) std::memcpy(rgb, s.env.src_rgb.data(), s.env.src_rgb.size() * sizeof(float));
        return AMVPT_OK; });
}
int amvpt_scene_get_sensor_params(amvpt_scene *h, uint32_t index, float *scalars4, int32_t *layout5) {
    return guard([&] { NEED(h); Scene &s = *h->s; if (index >= s.sensors.size()) AM_THROW("sensor index %u out of range", index);
        const DSensor &c = s.sensors[index];
        if (scalars4) { scalars4[0] = c.near_clip; scalars4[1] = c.far_clip; scalars4[2] = c.aperture_radius; scalars4[3] = c.focus_distance; }
        if (layout5) { layout5[0] = (int32_t) s.grid_x; layout5[1] = (int32_t) s.grid_y; layout5[2] = s.reverse_x; layout5[3] = s.reverse_y; layout5[4] = s.thinlens; }
        return AMVPT_OK; });
}
int amvpt_scene_get_shutter_draw(amvpt_scene *h, int32_t *draw) { return guard([&] { NEED(h); NEED(draw); *draw = h->s->shutter_draw ? 1 : 0; return AMVPT_OK; }); }
int amvpt_scene_set_shutter(amvpt_scene *h, float shutter_open, float shutter_close) {
    return guard([&] { NEED(h); const float t = shutter_close - shutter_open;
        if (t < 0.f) AM_THROW("Shutter opening time must be less than or equal to the shutter closing time!");
        h->s->shutter_draw = t > 0.f; if (h->s->committed) h->s->d.shutter_draw = t > 0.f; return AMVPT_OK; });
}
int amvpt_image_read(const char *path, uint32_t *w, uint32_t *hh, float *rgb) {
    return guard([&] { NEED(path); NEED(w); NEED(hh); std::vector<float> px; load_image_rgb(path, w, hh, px);
        if (rgb) std::memcpy(rgb, px.data(), px.size() * sizeof(float)); return AMVPT_OK; });
}
int amvpt_spectrum_to_srgb(const double *wavelengths, const double *values, uint32_t n, int within_emitter, int unbounded_name, double *rgb3) {
    return guard([&] { NEED(values); NEED(rgb3); NEED(n >= 1);
        if (n == 1) { if (within_emitter && !unbounded_name) uniform_emitter_spectrum_to_srgb(values[0], rgb3); else rgb3[0] = rgb3[1] = rgb3[2] = values[0]; return AMVPT_OK; }
        NEED(wavelengths);
        spectrum_to_srgb(std::vector<double>(wavelengths, wavelengths + n), std::vector<double>(values, values + n), within_emitter != 0, unbounded_name != 0, rgb3);
        return AMVPT_OK; });
}
int amvpt_env_warp(amvpt_scene *h, const float *u, uint32_t n, float *uv, float *pdf, float *eval_back) {
    return guard([&] { Scene &s = committed(h); NEED(u); NEED(uv); NEED(pdf); NEED(eval_back);
        if (!s.d.env_is_map) AM_THROW("amvpt_env_warp: the scene has no environment map");
        DBuf du(sizeof(float) * 2 * n), dout(sizeof(float) * 4 * n); du.up(u, sizeof(float) * 2 * n);
        launch_env_warp_tap(s.d, du.as<float>(), n, dout.as<float>(), nullptr); AM_CUDA(cudaDeviceSynchronize());
        std::vector<float> out(4 * (size_t) n); dout.down(out.data(), sizeof(float) * 4 * n);
        for (uint32_t i = 0; i < n; ++i) { uv[2 * i] = out[4 * i]; uv[2 * i + 1] = out[4 * i + 1]; pdf[i] = out[4 * i + 2]; eval_back[i] = out[4 * i + 3]; }
        return AMVPT_OK; });
}
int amvpt_scene_set_sensor_grid(amvpt_scene *h, const amvpt_grid_sensor_desc *d) {
    return guard([&] { NEED(h); NEED(d);
        GridSensorParams g; g.thinlens = d->thinlens; g.fov_axis = d->fov_axis; g.grid_x = d->grid_x; g.grid_y = d->grid_y;
        g.res_x = d->res_x; g.res_y = d->res_y; g.reverse_x = d->reverse_x; g.reverse_y = d->reverse_y; g.use_cone = d->use_cone; g.cam_center = d->cam_center;
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) g.to_world.m[i][j] = d->to_world[4 * i + j];
        g.fov = d->fov; g.near_clip = d->near_clip; g.far_clip = d->far_clip; g.focus_distance = d->focus_distance;
        g.aperture_radius = d->aperture_radius; g.cone_deg = d->cone_deg; g.cam_dist = d->cam_dist;
        for (int i = 0; i < 3; ++i) { g.cam_dir[i] = d->cam_dir[i]; g.cam_off[i] = d->cam_off[i]; }
        h->s->set_sensor_grid(g); return AMVPT_OK; });
}
int amvpt_scene_set_sensor_batch(amvpt_scene *h, uint32_t n, uint32_t res_x, uint32_t res_y, const float *s2c, const float *to_world,
                                 float near_clip, float far_clip, int reverse_x) {
    return guard([&] { NEED(h); NEED(s2c); NEED(to_world);
        h->s->set_sensor_explicit(n, res_x, res_y, s2c, to_world, near_clip, far_clip, reverse_x != 0); return AMVPT_OK; });
}
int amvpt_scene_set_sensor_table(amvpt_scene *h, uint32_t grid_x, uint32_t grid_y, uint32_t view_res_x, uint32_t view_res_y, int reverse_x, int reverse_y, int thinlens,
                                 const float *s2c, const float *to_world, const float *scalars4) {
    return guard([&] { NEED(h); NEED(s2c); NEED(to_world); NEED(scalars4);
        h->s->set_sensor_table(grid_x, grid_y, view_res_x, view_res_y, reverse_x != 0, reverse_y != 0, thinlens != 0, s2c, to_world, scalars4); return AMVPT_OK; });
}
int amvpt_scene_set_integrator(amvpt_scene *h, const amvpt_integrator_desc *d) {
    return guard([&] { NEED(h); NEED(d);
        IntegratorParams &ip = h->s->integrator;
        if (d->max_depth < 0 && d->max_depth != -1) AM_THROW("\"max_depth\" must be set to -1 (infinite) or a value >= 0");   // integrator.cpp:510-512
        ip.max_depth = d->max_depth; ip.rr_depth = d->rr_depth; ip.hide_emitters = d->hide_emitters != 0; ip.spp_pass_lim = d->spp_pass_lim;
        if (ip.rr_depth == 0) AM_THROW("\"rr_depth\" must be set to a value greater than zero!");                              // integrator.cpp:517-518
        h->s->sampler_seed = d->sampler_seed; if (d->sample_count) h->s->sample_count = d->sample_count;
        if (h->s->committed) { h->s->d.max_depth = ip.max_depth < 0 ? 0xffffffffu : (uint32_t) ip.max_depth; h->s->d.rr_depth = ip.rr_depth; h->s->d.hide_emitters = ip.hide_emitters ? 1 : 0; }
        return AMVPT_OK; });
}
int amvpt_scene_set_reuse(amvpt_scene *h, int sa_reuse, int sa_mis, int fast_mis, uint32_t reuse_count, uint32_t adaptive, uint32_t *n_samples) {
    return guard([&] { NEED(h); h->s->set_reuse(sa_reuse != 0, sa_mis != 0, fast_mis != 0, reuse_count, adaptive);
        if (n_samples) *n_samples = h->s->reuse_samples(); return AMVPT_OK; });
}
int amvpt_scene_set_timeout(amvpt_scene *h, float seconds) { return guard([&] { NEED(h); h->s->integrator.timeout = seconds; return AMVPT_OK; }); }
int amvpt_scene_set_reuse_debug(amvpt_scene *h, int debug) {
    return guard([&] { NEED(h); h->s->integrator.debug = debug != 0; if (h->s->committed) h->s->d.reuse_debug = debug != 0; return AMVPT_OK; });
}
int amvpt_adaptive_alloc(amvpt_scene *h, uint32_t *out, uint32_t cap, uint32_t *n_passes, uint32_t *n_views) {
    return guard([&] { NEED(h); Scene &s = *h->s; const uint32_t nv = s.grid_x * s.grid_y, np = nv ? (uint32_t) (s.alloc_log.size() / nv) : 0;
        if (n_passes) *n_passes = np; if (n_views) *n_views = nv;
        if (out) { if (cap < s.alloc_log.size()) AM_THROW("amvpt_adaptive_alloc: buffer too small (%zu entries)", s.alloc_log.size()); std::memcpy(out, s.alloc_log.data(), s.alloc_log.size() * 4); }
        return AMVPT_OK; });
}
int amvpt_last_fill_paths(amvpt_scene *h, uint64_t *n) { return guard([&] { NEED(h); NEED(n); *n = h->s->last_fill_paths; return AMVPT_OK; }); }
int amvpt_scene_set_rfilter(amvpt_scene *h, int type, float param, float param2) {
    return guard([&] { NEED(h); h->s->set_rfilter(type, param, param2); return AMVPT_OK; });
}
int amvpt_scene_commit(amvpt_scene *h, int device) {
    return guard([&] { NEED(h); need_device(device); h->s->commit(device); return AMVPT_OK; });
}
int amvpt_scene_info(amvpt_scene *h, uint32_t *film_w, uint32_t *film_h, uint32_t *n_views, uint32_t *n_prims, uint32_t *n_nodes, uint32_t *sample_count) {
    return guard([&] { NEED(h); Scene &s = *h->s;
        if (film_w) *film_w = s.film_w; if (film_h) *film_h = s.film_h; if (n_views) *n_views = (uint32_t) s.sensors.size();
        if (n_prims) *n_prims = s.nprims(); if (n_nodes) *n_nodes = s.n_nodes; if (sample_count) *sample_count = s.sample_count;
        return AMVPT_OK; });
}
int amvpt_scene_get_sensor(amvpt_scene *h, uint32_t index, float *s2c, float *to_world) {
    return guard([&] { NEED(h); NEED(index < h->s->sensors.size()); const DSensor &c = h->s->sensors[index];
        if (s2c) memcpy(s2c, c.s2c, 64);
        if (to_world) { memcpy(to_world, c.to_world, 48); to_world[12] = to_world[13] = to_world[14] = 0.f; to_world[15] = 1.f; }
        return AMVPT_OK; });
}
int amvpt_scene_get_bsdf_tables(amvpt_scene *h, int bsdf, float *ext_t, float *scalars) {
    return guard([&] { NEED(h); NEED(bsdf >= 0 && bsdf < (int) h->s->bsdfs.size()); const DBsdf &b = h->s->bsdfs[bsdf];
        if (ext_t) memcpy(ext_t, b.ext_transmittance, sizeof(b.ext_transmittance));
        if (scalars) { scalars[0] = b.inv_eta_2; scalars[1] = b.spec_sampling_weight; scalars[2] = b.internal_reflectance; }
        return AMVPT_OK; });
}
int amvpt_scene_counts(amvpt_scene *h, uint32_t *n_meshes, uint32_t *n_bsdfs, uint32_t *n_emitters) {
    return guard([&] { NEED(h); Scene &s = *h->s;
        if (n_meshes) *n_meshes = (uint32_t) s.meshes.size(); if (n_bsdfs) *n_bsdfs = (uint32_t) s.bsdfs.size(); if (n_emitters) *n_emitters = (uint32_t) s.emitters.size();
        return AMVPT_OK; });
}
int amvpt_scene_get_mesh(amvpt_scene *h, uint32_t index, uint32_t *n_vertices, uint32_t *n_faces, int32_t *bsdf, int32_t *has_normals, int32_t *flip_normals,
                         int32_t *is_emitter, float *radiance3, float *positions, uint32_t *faces, float *normals) {
    return guard([&] { NEED(h); NEED(index < h->s->meshes.size()); const HMesh &m = h->s->meshes[index];
        if (n_vertices) *n_vertices = (uint32_t) (m.pos.size() / 3); if (n_faces) *n_faces = m.nfaces(); if (bsdf) *bsdf = m.bsdf;
        if (has_normals) *has_normals = m.nrm.empty() ? 0 : 1; if (flip_normals) *flip_normals = m.flip_normals; if (is_emitter) *is_emitter = m.emitter >= 0;
        if (radiance3 && m.emitter >= 0) memcpy(radiance3, h->s->emitters[m.emitter].radiance, 12);
        if (positions) memcpy(positions, m.pos.data(), m.pos.size() * 4); if (faces) memcpy(faces, m.faces.data(), m.faces.size() * 4);
        if (normals && !m.nrm.empty()) memcpy(normals, m.nrm.data(), m.nrm.size() * 4);
        return AMVPT_OK; });
}
int amvpt_scene_get_bsdf(amvpt_scene *h, int bsdf, amvpt_bsdf_desc *d) {
    return guard([&] { NEED(h); NEED(d); NEED(bsdf >= 0 && bsdf < (int) h->s->bsdfs.size()); const DBsdf &b = h->s->bsdfs[bsdf];
        d->type = b.type; d->twosided = b.twosided; d->distr = b.distr; d->nonlinear = b.nonlinear;
        for (int i = 0; i < 3; ++i) { d->reflectance[i] = b.reflectance[i]; d->eta[i] = b.eta[i]; d->k[i] = b.k[i];
            d->specular_reflectance[i] = b.spec_refl[i]; d->specular_transmittance[i] = b.spec_trans[i]; }
        d->int_ior = b.ior_eta; d->ext_ior = 1.f; d->alpha_u = b.alpha_u; d->alpha_v = b.alpha_v;
        return AMVPT_OK; });
}
int amvpt_scene_get_integrator(amvpt_scene *h, amvpt_integrator_desc *d, int32_t *is_mvpath, int32_t *sa_reuse, uint32_t *reuse_count) {
    return guard([&] { NEED(h); NEED(d); const IntegratorParams &ip = h->s->integrator;
        d->max_depth = ip.max_depth; d->rr_depth = ip.rr_depth; d->hide_emitters = ip.hide_emitters; d->spp_pass_lim = ip.spp_pass_lim;
        d->sampler_seed = h->s->sampler_seed; d->sample_count = h->s->sample_count;
        if (is_mvpath) *is_mvpath = ip.is_mvpath; if (sa_reuse) *sa_reuse = ip.sa_reuse; if (reuse_count) *reuse_count = ip.reuse_count;
        return AMVPT_OK; });
}

// ---- hot path
int amvpt_render(amvpt_scene *h, const amvpt_render_desc *d, amvpt_render_stats *out) {
    return guard([&] { NEED(d); Scene &s = committed(h);
        RenderOptions o; o.seed = d->seed; o.spp = d->spp; o.scheduler = d->scheduler; o.stats = (d->collect_stats & 1) != 0; o.time_stages = (d->collect_stats & 2) != 0;
        o.shard.rank = d->rank; o.shard.world = d->world ? d->world : 1; o.stream = (cudaStream_t) d->stream;
        o.adaptive = d->adaptive != 0; o.adaptive_threshold = d->adaptive_threshold; o.adaptive_min_spp = d->adaptive_min_spp ? d->adaptive_min_spp : 16;
        o.allreduce = d->allreduce; o.allreduce_user = d->allreduce_user; o.comm = d->comm;
        render(s, o);
        if (out) { const RenderStats &r = s.last; out->seconds = r.seconds; out->paths = r.paths; out->closest_rays = r.closest_rays; out->shadow_rays = r.shadow_rays;
            out->bvh_nodes = r.nodes; out->tri_tests = r.tris; out->path_vertices = r.vertices; out->passes = r.passes; out->launches = r.launches; out->spp_done = r.spp_done; out->scheduler_used = r.scheduler_used;
            out->trace_seconds = r.trace_seconds; out->trace_launches = r.trace_launches; out->pad = 0; }
        return AMVPT_OK; });
}
int amvpt_shard_tiles(amvpt_scene *h, uint32_t rank, uint32_t world, uint32_t *out5, uint32_t cap) {
    return guard([&] { NEED(h); NEED(world > 0 && rank < world); if (h->s->sensors.empty()) AM_THROW("no sensor");
        Shard sh; sh.rank = rank; sh.world = world; std::vector<TileDesc> t = rank_tiles(*h->s, sh, false);
        for (size_t i = 0; i < t.size() && i < cap && out5; ++i) { uint32_t *o = out5 + 5 * i; o[0] = t[i].x0; o[1] = t[i].y0; o[2] = t[i].w; o[3] = t[i].h; o[4] = t[i].pad; }
        return (int) t.size(); });
}
int amvpt_view_tiles(amvpt_scene *h, uint32_t *out5, uint32_t cap) {
    return guard([&] { NEED(h); if (h->s->sensors.empty()) AM_THROW("no sensor");
        Shard all; std::vector<TileDesc> t = h->s->view_tiles(all);
        for (size_t i = 0; i < t.size() && i < cap && out5; ++i) { uint32_t *o = out5 + 5 * i; o[0] = t[i].x0; o[1] = t[i].y0; o[2] = t[i].w; o[3] = t[i].h; o[4] = t[i].pad; }
        return (int) t.size(); });
}
int amvpt_cancel(amvpt_scene *h) { return guard([&] { NEED(h); h->s->stop = true; return AMVPT_OK; }); }
int amvpt_film_clear(amvpt_scene *h) { return guard([&] { Scene &s = committed(h); s.film_clear(nullptr); AM_CUDA(cudaDeviceSynchronize()); return AMVPT_OK; }); }
int amvpt_film_raw(amvpt_scene *h, float *rgbw) {
    return guard([&] { NEED(rgbw); Scene &s = committed(h);
        AM_CUDA(cudaMemcpy(rgbw, s.film, sizeof(float4) * (size_t) s.film_w * s.film_h, cudaMemcpyDeviceToHost)); return AMVPT_OK; });
}
int amvpt_film_develop_device(amvpt_scene *h, float *rgb_dev, void *stream) {
    return guard([&] { NEED(rgb_dev); Scene &s = committed(h);
        launch_develop(s.film, rgb_dev, (size_t) s.film_w * s.film_h, (cudaStream_t) stream); AM_CUDA(cudaGetLastError()); return AMVPT_OK; });
}
int amvpt_film_develop(amvpt_scene *h, float *rgb) {
    return guard([&] { NEED(rgb); Scene &s = committed(h); size_t npix = (size_t) s.film_w * s.film_h;
        if (!s.rgb) s.rgb = (float *) s.dalloc(npix * 3 * sizeof(float));
        launch_develop(s.film, s.rgb, npix, nullptr); AM_CUDA(cudaGetLastError());
        AM_CUDA(cudaMemcpy(rgb, s.rgb, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost)); return AMVPT_OK; });
}
int amvpt_film_develop_pinned(amvpt_scene *h, float **rgb_host) {
    return guard([&] { NEED(rgb_host); Scene &s = committed(h); const size_t npix = (size_t) s.film_w * s.film_h, bytes = npix * 3 * sizeof(float);
        static float *pinned = nullptr; static size_t pinned_bytes = 0;        // process-wide: page-locking 200 MB costs tens of milliseconds, so it happens once
        if (pinned_bytes < bytes) { if (pinned) cudaFreeHost(pinned); pinned = nullptr; pinned_bytes = 0; AM_CUDA(cudaHostAlloc((void **) &pinned, bytes, cudaHostAllocDefault)); pinned_bytes = bytes; }
        if (!s.rgb) s.rgb = (float *) s.dalloc(bytes);
        launch_develop(s.film, s.rgb, npix, nullptr); AM_CUDA(cudaGetLastError());
        AM_CUDA(cudaMemcpyAsync(pinned, s.rgb, bytes, cudaMemcpyDeviceToHost, nullptr)); AM_CUDA(cudaStreamSynchronize(nullptr));
        *rgb_host = pinned; return AMVPT_OK; });
}
int amvpt_film_device_ptr(amvpt_scene *h, void **rgbw, void **moments) {
    return guard([&] { Scene &s = committed(h); if (rgbw) *rgbw = s.film; if (moments) *moments = s.moments; return AMVPT_OK; });
}
int amvpt_film_add_host(amvpt_scene *h, const float *rgbw) {
    return guard([&] { NEED(rgbw); Scene &s = committed(h); size_t n = (size_t) s.film_w * s.film_h * 4;
        std::vector<float> cur(n); AM_CUDA(cudaMemcpy(cur.data(), s.film, n * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) cur[i] += rgbw[i];
        AM_CUDA(cudaMemcpy(s.film, cur.data(), n * 4, cudaMemcpyHostToDevice)); return AMVPT_OK; });
}

// ---- output path and display step
namespace {
struct EventTimer {     // device time of the kernels launched between the constructor and stop() on the null stream
    cudaEvent_t a, b; EventTimer() { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, nullptr); }
    float stop() { float ms = 0.f; cudaEventRecord(b, nullptr); cudaEventSynchronize(b); cudaEventElapsedTime(&ms, a, b); return ms; }
    ~EventTimer() { cudaEventDestroy(a); cudaEventDestroy(b); }
};
}
int amvpt_film_to_srgb8(amvpt_scene *h, uint8_t *rgb8) {
    return guard([&] { Scene &s = committed(h); size_t npix = (size_t) s.film_w * s.film_h;
        if (!s.srgb8) s.srgb8 = (uint8_t *) s.dalloc(npix * 3 + 16);
        { EventTimer t; launch_srgb8(s.film, s.srgb8, npix, s.sm_count, nullptr); AM_CUDA(cudaGetLastError()); s.display_ms[0] = t.stop(); }
        if (rgb8) AM_CUDA(cudaMemcpy(rgb8, s.srgb8, npix * 3, cudaMemcpyDeviceToHost));
        return AMVPT_OK; });
}
int amvpt_interlace(amvpt_scene *h, const uint8_t *src, uint32_t src_w, uint32_t src_h, uint32_t src_ch, const amvpt_preset *p,
                    uint32_t dst_w, uint32_t dst_h, uint32_t dst_ch, uint8_t *dst) {
    return guard([&] { Scene &s = committed(h); NEED(p); NEED(dst); NEED(dst_ch == 3 || dst_ch == 4); NEED(dst_w >= 2 && dst_h >= 2);
        NEED(p->grid[0] > 0 && p->grid[1] > 0);
        static_assert(sizeof(DisplayPreset) == sizeof(amvpt_preset), "preset layouts differ");
        DisplayPreset dp; memcpy(&dp, p, sizeof(dp));
        const uint8_t *dsrc; DBuf up(0);
        if (src) { NEED(src_w >= 1 && src_h >= 1 && src_ch >= 3); size_t n = (size_t) src_w * src_h * src_ch; DBuf b(n); b.up(src, n); std::swap(up.p, b.p); dsrc = up.as<uint8_t>(); }
        else { if (!s.srgb8) AM_THROW("interlace: no source image and amvpt_film_to_srgb8 has not been called"); dsrc = s.srgb8; src_w = s.film_w; src_h = s.film_h; src_ch = 3; }
        size_t nd = (size_t) dst_w * dst_h * dst_ch; DBuf out(nd);
        { EventTimer t; launch_interlace(dsrc, src_w, src_h, src_ch, dp, out.as<uint8_t>(), dst_w, dst_h, dst_ch, nullptr); AM_CUDA(cudaGetLastError()); s.display_ms[1] = t.stop(); }
        AM_CUDA(cudaMemcpy(dst, out.p, nd, cudaMemcpyDeviceToHost)); return AMVPT_OK; });
}
int amvpt_display_last_ms(amvpt_scene *h, float *srgb8_ms, float *interlace_ms) {
    return guard([&] { Scene &s = committed(h); if (srgb8_ms) *srgb8_ms = s.display_ms[0]; if (interlace_ms) *interlace_ms = s.display_ms[1]; return AMVPT_OK; });
}
// developed quilt on the host + the rectangle `view` selects (view < 0: the whole quilt)
struct Developed { std::vector<float> rgb; uint32_t x0 = 0, y0 = 0, w = 0, h = 0, stride = 0; };
static Developed develop_view(Scene &s, int32_t view, const char *who) {
    size_t npix = (size_t) s.film_w * s.film_h; Developed d;
    if (!s.rgb) s.rgb = (float *) s.dalloc(npix * 3 * sizeof(float));
    launch_develop(s.film, s.rgb, npix, nullptr); AM_CUDA(cudaGetLastError());
    d.rgb.resize(npix * 3); AM_CUDA(cudaMemcpy(d.rgb.data(), s.rgb, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    d.w = d.stride = s.film_w; d.h = s.film_h;
    if (view >= 0) {
        Shard all; std::vector<TileDesc> t = s.view_tiles(all); bool found = false;      // one rectangle per sub-sensor, pad = view index
        for (auto &r : t) if ((int32_t) r.pad == view) { d.x0 = r.x0; d.y0 = r.y0; d.w = r.w; d.h = r.h; found = true; break; }
        if (!found) AM_THROW("%s: view %d out of range (%zu views)", who, view, t.size());
    }
    return d;
}
int amvpt_film_write_pfm(amvpt_scene *h, const char *path, int32_t view) {
    return guard([&] { NEED(path); Scene &s = committed(h); Developed d = develop_view(s, view, "write_pfm");
        FILE *f = fopen(path, "wb"); if (!f) AM_THROW("write_pfm: cannot open %s", path);
        fprintf(f, "PF\n%u %u\n-1\n", d.w, d.h);                      // Bitmap::write_pfm (src/core/bitmap.cpp:2270-2299): little endian, bottom row first
        bool ok = true;
        for (uint32_t y = 0; y < d.h && ok; ++y) ok = fwrite(d.rgb.data() + ((size_t) (d.y0 + d.h - 1 - y) * d.stride + d.x0) * 3, sizeof(float), (size_t) d.w * 3, f) == (size_t) d.w * 3;
        ok = (fclose(f) == 0) && ok; if (!ok) AM_THROW("write_pfm: short write to %s", path);
        return AMVPT_OK; });
}
// float32 -> float16, round to nearest even (what dr::half / Struct conversion to Float16 does; overflow -> inf, NaN kept quiet)
static uint16_t f32_to_f16(float v) {
    uint32_t x; memcpy(&x, &v, 4); uint32_t sign = (x >> 16) & 0x8000u; x &= 0x7fffffffu;
    if (x >= 0x7f800000u) return (uint16_t) (sign | 0x7c00u | (x > 0x7f800000u ? 0x200u | ((x >> 13) & 0x3ffu) : 0u));
    if (x >= 0x477ff000u) return (uint16_t) (sign | 0x7c00u);          // rounds to >= 65520 -> inf
    if (x < 0x33000001u) return (uint16_t) sign;                        // <= 2^-25 -> +-0 (ties to even)
    uint32_t e = x >> 23, m = (x & 0x7fffffu) | 0x800000u, shift, hbits;
    if (e < 113) { shift = 126 - e; hbits = 0; } else { shift = 13; hbits = (e - 112) << 10; m &= 0x7fffffu; }
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    q += (rem > half) || (rem == half && (q & 1u));
    return (uint16_t) (sign | (hbits + q));                             // a mantissa carry moves into the exponent, as it should
}
int amvpt_f32_to_f16(const float *src, uint16_t *dst, uint64_t n) {
    return guard([&] { if (n) { NEED(src); NEED(dst); } for (uint64_t i = 0; i < n; ++i) dst[i] = f32_to_f16(src[i]); return AMVPT_OK; });
}
int amvpt_film_write_exr(amvpt_scene *h, const char *path, int32_t view, int32_t half_float) {
    return guard([&] { NEED(path); Scene &s = committed(h); Developed d = develop_view(s, view, "write_exr");
        // OpenEXR 2 single-part scan-line file, NO_COMPRESSION, channels B G R (alphabetical, as the format requires), top row first
        std::vector<uint8_t> hd; auto put = [&](const void *p, size_t n) { hd.insert(hd.end(), (const uint8_t *) p, (const uint8_t *) p + n); };
        auto str = [&](const char *z) { put(z, strlen(z) + 1); }; auto i32 = [&](int32_t v) { put(&v, 4); }; auto f32 = [&](float v) { put(&v, 4); };
        auto attr = [&](const char *name, const char *type, int32_t size) { str(name); str(type); i32(size); };
        i32(20000630); i32(2);
        attr("channels", "chlist", 3 * 18 + 1);
        for (const char *c : { "B", "G", "R" }) { str(c); i32(half_float ? 1 : 2); uint8_t z4[4] = { 0, 0, 0, 0 }; put(z4, 4); i32(1); i32(1); }
        hd.push_back(0);
        attr("compression", "compression", 1); hd.push_back(0);
        attr("dataWindow", "box2i", 16); i32(0); i32(0); i32((int32_t) d.w - 1); i32((int32_t) d.h - 1);
        attr("displayWindow", "box2i", 16); i32(0); i32(0); i32((int32_t) d.w - 1); i32((int32_t) d.h - 1);
        attr("lineOrder", "lineOrder", 1); hd.push_back(0);
        attr("pixelAspectRatio", "float", 4); f32(1.f);
        attr("screenWindowCenter", "v2f", 8); f32(0.f); f32(0.f);
        attr("screenWindowWidth", "float", 4); f32(1.f);
        hd.push_back(0);
        size_t bpc = half_float ? 2 : 4, row = (size_t) d.w * 3 * bpc; uint64_t off = hd.size() + (uint64_t) d.h * 8;
        for (uint32_t y = 0; y < d.h; ++y) { uint64_t o = off + (uint64_t) y * (8 + row); put(&o, 8); }
        FILE *f = fopen(path, "wb"); if (!f) AM_THROW("write_exr: cannot open %s", path);
        bool ok = fwrite(hd.data(), 1, hd.size(), f) == hd.size(); std::vector<uint8_t> line(8 + row);
        for (uint32_t y = 0; y < d.h && ok; ++y) {
            int32_t yy = (int32_t) y, n = (int32_t) row; memcpy(line.data(), &yy, 4); memcpy(line.data() + 4, &n, 4);
            const float *src = d.rgb.data() + ((size_t) (d.y0 + y) * d.stride + d.x0) * 3;
            for (int c = 0; c < 3; ++c) {                               // file channel c = B, G, R <- source component 2 - c
                uint8_t *dst = line.data() + 8 + (size_t) c * d.w * bpc;
                for (uint32_t x = 0; x < d.w; ++x) { float v = src[(size_t) x * 3 + (2 - c)];
                    if (half_float) { uint16_t hv = f32_to_f16(v); memcpy(dst + (size_t) x * 2, &hv, 2); } else memcpy(dst + (size_t) x * 4, &v, 4); }
            }
            ok = fwrite(line.data(), 1, line.size(), f) == line.size();
        }
        ok = (fclose(f) == 0) && ok; if (!ok) AM_THROW("write_exr: short write to %s", path);
        return AMVPT_OK; });
}

// ---- parity taps
int amvpt_trace_lanes_multi(amvpt_scene *h, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, float *out9) {
    return guard([&] { Scene &s = committed(h); NEED(spp_per_pass > 0); NEED(out9);
        if (!s.d.n_reuse) AM_THROW("trace_lanes_multi: sample reuse is off");
        size_t bytes = (size_t) n * s.d.n_reuse * 9 * 4; DBuf b(bytes);
        launch_reuse_tap(s.d, seed_value, spp_per_pass, lane_begin, n, b.as<float>(), nullptr);
        AM_CUDA(cudaGetLastError()); AM_CUDA(cudaDeviceSynchronize());
        b.down(out9, bytes); return AMVPT_OK; });
}
int amvpt_trace_lanes(amvpt_scene *h, uint32_t seed_value, uint32_t spp_per_pass, uint32_t lane_begin, uint32_t n, int trace,
                      float *sample_pos, float *ray_o, float *ray_d, float *ray_maxt, uint32_t *view, float *radiance,
                      uint32_t *pixel, uint32_t max_hits, uint32_t *hit_prim, float *hit_t, uint64_t *stats2) {
    return guard([&] { Scene &s = committed(h); NEED(spp_per_pass > 0);
        size_t N = n;
        DBuf b_sp(N * 8), b_o(N * 12), b_d(N * 12), b_mt(N * 4), b_v(N * 4), b_r(N * 16), b_px(N * 4), b_hp(N * max_hits * 4), b_ht(N * max_hits * 4), b_c(8 * CTR_COUNT);
        AM_CUDA(cudaMemset(b_c.p, 0, 8 * CTR_COUNT));
        LaneTaps t; t.sample_pos = b_sp.as<float>(); t.ray_o = b_o.as<float>(); t.ray_d = b_d.as<float>(); t.ray_maxt = b_mt.as<float>();
        t.view = b_v.as<uint32_t>(); t.radiance = b_r.as<float>(); t.pixel = b_px.as<uint32_t>(); t.max_hits = max_hits;
        t.hit_prim = b_hp.as<uint32_t>(); t.hit_t = b_ht.as<float>();
        launch_trace_lanes(s.d, seed_value, spp_per_pass, lane_begin, n, trace, t, b_c.as<unsigned long long>(), nullptr);
        AM_CUDA(cudaGetLastError()); AM_CUDA(cudaDeviceSynchronize());
        if (sample_pos) b_sp.down(sample_pos, N * 8);
        if (ray_o) b_o.down(ray_o, N * 12);
        if (ray_d) b_d.down(ray_d, N * 12);
        if (ray_maxt) b_mt.down(ray_maxt, N * 4);
        if (view) b_v.down(view, N * 4);
        if (radiance) b_r.down(radiance, N * 16);
        if (pixel) b_px.down(pixel, N * 4);
        if (hit_prim && max_hits) b_hp.down(hit_prim, N * max_hits * 4);
        if (hit_t && max_hits) b_ht.down(hit_t, N * max_hits * 4);
        if (stats2) { unsigned long long c[CTR_COUNT]; b_c.down(c, sizeof(c)); stats2[0] = c[CTR_NODES]; stats2[1] = c[CTR_TRIS]; }
        return AMVPT_OK; });
}
int amvpt_intersect(amvpt_scene *h, const float *o, const float *d, const float *maxt, uint32_t n, int any_hit, uint32_t *prim, float *tuv, uint64_t *stats2) {
    return guard([&] { Scene &s = committed(h); NEED(o && d && maxt && prim); size_t N = n;
        DBuf b_o(N * 12), b_d(N * 12), b_m(N * 4), b_p(N * 4), b_t(N * 12), b_c(8 * CTR_COUNT);
        b_o.up(o, N * 12); b_d.up(d, N * 12); b_m.up(maxt, N * 4); AM_CUDA(cudaMemset(b_c.p, 0, 8 * CTR_COUNT));
        launch_intersect(s.d, b_o.as<float>(), b_d.as<float>(), b_m.as<float>(), n, any_hit, b_p.as<uint32_t>(), b_t.as<float>(), b_c.as<unsigned long long>(), nullptr);
        AM_CUDA(cudaGetLastError()); AM_CUDA(cudaDeviceSynchronize());
        b_p.down(prim, N * 4); if (tuv) b_t.down(tuv, N * 12);
        if (stats2) { unsigned long long c[CTR_COUNT]; b_c.down(c, sizeof(c)); stats2[0] = c[CTR_NODES]; stats2[1] = c[CTR_TRIS]; }
        return AMVPT_OK; });
}
int amvpt_bsdf_eval_pdf_sample(amvpt_scene *h, int bsdf, const float *wi, const float *wo, const float *s1, const float *s2, uint32_t n,
                               float *f, float *pdf, float *bs_wo, float *bs_pdf, float *bs_eta, uint32_t *bs_type, float *weight) {
    return guard([&] { Scene &s = committed(h); NEED(bsdf >= 0 && bsdf < (int) s.bsdfs.size()); NEED(wi && wo && s1 && s2); size_t N = n;
        DBuf b_wi(N * 12), b_wo(N * 12), b_s1(N * 4), b_s2(N * 8), b_out(N * 64);
        b_wi.up(wi, N * 12); b_wo.up(wo, N * 12); b_s1.up(s1, N * 4); b_s2.up(s2, N * 8);
        launch_bsdf_tap(s.d, bsdf, b_wi.as<float>(), b_wo.as<float>(), b_s1.as<float>(), b_s2.as<float>(), n, b_out.as<float>(), nullptr);
        AM_CUDA(cudaGetLastError()); AM_CUDA(cudaDeviceSynchronize());
        std::vector<float> q(N * 16); b_out.down(q.data(), N * 64);
        for (size_t i = 0; i < N; ++i) { const float *r = &q[16 * i];
            if (f) { f[3 * i] = r[0]; f[3 * i + 1] = r[1]; f[3 * i + 2] = r[2]; } if (pdf) pdf[i] = r[3];
            if (bs_wo) { bs_wo[3 * i] = r[4]; bs_wo[3 * i + 1] = r[5]; bs_wo[3 * i + 2] = r[6]; } if (bs_pdf) bs_pdf[i] = r[7]; if (bs_eta) bs_eta[i] = r[8];
            if (bs_type) memcpy(&bs_type[i], &r[9], 4); if (weight) { weight[3 * i] = r[10]; weight[3 * i + 1] = r[11]; weight[3 * i + 2] = r[12]; } }
        return AMVPT_OK; });
}
int amvpt_math1(int device, int op, const float *in, uint32_t n, float *out) {
    return guard([&] { NEED(in && out); need_device(device); size_t N = n; DBuf a(N * 4), b(N * 4); a.up(in, N * 4);
        launch_math_tap(op, a.as<float>(), n, b.as<float>(), nullptr); AM_CUDA(cudaGetLastError()); AM_CUDA(cudaDeviceSynchronize()); b.down(out, N * 4); return AMVPT_OK; });
}
int amvpt_sampler_stream(int device, uint32_t seed_value, uint32_t lane, uint32_t n, float *out) {
    return guard([&] { NEED(out); need_device(device); DBuf b((size_t) n * 4);
        launch_sampler_tap(seed_value, lane, n, b.as<float>(), nullptr); AM_CUDA(cudaGetLastError()); AM_CUDA(cudaDeviceSynchronize()); b.down(out, (size_t) n * 4); return AMVPT_OK; });
}
// Host-side structural check of the tree that was uploaded (no GPU work): containment of triangles and child boxes.
int amvpt_membench(int device, int kind, uint64_t bytes, uint32_t iters, double *gb_per_s) {
    return guard([&] { NEED(gb_per_s); need_device(device);
        if (kind != 0) AM_THROW("amvpt_membench: unknown kind %d (0 = 16-byte reads)", kind);
        int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const double v = run_membench_read((size_t) bytes, iters, sms, nullptr);
        if (v < 0.0) AM_THROW("CUDA error: membench failed (%s)", cudaGetErrorString(cudaGetLastError()));
        *gb_per_s = v; return AMVPT_OK; });
}
int amvpt_bvh_checksum(amvpt_scene *h, uint64_t *checksum) {
    return guard([&] { NEED(h); NEED(checksum); Scene &s = *h->s; s.prepare_host();
        uint64_t x = 1469598103934665603ull;
        auto eat = [&](const void *p, size_t n) { const unsigned char *b = (const unsigned char *) p; for (size_t i = 0; i < n; ++i) { x ^= b[i]; x *= 1099511628211ull; } };
        eat(s.bvh.nodes.data(), s.bvh.nodes.size() * sizeof(float4)); eat(s.bvh.tris.data(), s.bvh.tris.size() * sizeof(float4));
        eat(s.bvh.nodes_q4.data(), s.bvh.nodes_q4.size() * sizeof(uint4)); eat(s.bvh.prim_to_tri.data(), s.bvh.prim_to_tri.size() * 4);
        *checksum = x; return AMVPT_OK; });
}
int amvpt_bvh_q4_validate(amvpt_scene *h, uint32_t *n_nodes, uint32_t *max_depth, uint32_t *n_leaves) {
    return guard([&] { NEED(h); Scene &s = *h->s; s.prepare_host();
        const std::vector<uint4> &q = s.bvh.nodes_q4;
        if (n_nodes) *n_nodes = (uint32_t) (q.size() / 4);
        uint32_t leaves = 0, maxd = 0; std::vector<uint8_t> seen(s.bvh.tris.size() / 3, 0);
        if (q.empty()) { if (max_depth) *max_depth = 0; if (n_leaves) *n_leaves = 0; return AMVPT_OK; }
        // every triangle under a child slot must lie inside that slot's decoded box (origin + q * 2^(e - 127), evaluated as the kernel does)
        struct B { float lo[3], hi[3]; };
        std::function<B(int32_t, uint32_t)> rec = [&](int32_t ref, uint32_t depth) -> B {
            maxd = std::max(maxd, depth); B tb; for (int a = 0; a < 3; ++a) { tb.lo[a] = FLT_MAX; tb.hi[a] = -FLT_MAX; }
            if (ref < 0) {
                uint32_t code = ~(uint32_t) ref, first = code >> 3, count = code & 7u; leaves++;
                for (uint32_t i = first; i < first + count; ++i) {
                    if (seen[i]++) AM_THROW("bvh q4: triangle %u referenced twice", i);
                    for (int k = 0; k < 3; ++k) { const float4 &v = s.bvh.tris[3 * (size_t) i + k]; const float p[3] = { v.x, v.y, v.z };
                        for (int a = 0; a < 3; ++a) { tb.lo[a] = std::min(tb.lo[a], p[a]); tb.hi[a] = std::max(tb.hi[a], p[a]); } }
                }
                return tb;
            }
            const uint4 *n = &q[4 * (size_t) ref]; float org[3]; memcpy(org, &n[0].x, 12);
            const uint32_t ql[3] = { n[1].x, n[1].y, n[1].z }, qh[3] = { n[1].w, n[2].x, n[2].y }; const int32_t c[4] = { (int32_t) n[2].z, (int32_t) n[2].w, (int32_t) n[3].x, (int32_t) n[3].y };
            for (int i = 0; i < 4; ++i) {
                if (c[i] == 0) continue;
                if (c[i] > 0 && c[i] <= ref) AM_THROW("bvh q4: child %d of node %d is not numbered after its parent", c[i], ref);
                const B kb = rec(c[i], depth + 1);
                for (int a = 0; a < 3; ++a) {
                    uint32_t eb = ((n[0].w >> (8 * a)) & 0xffu) << 23; float scale; memcpy(&scale, &eb, 4);
                    const float lo = std::fma((float) ((ql[a] >> (8 * i)) & 0xffu), scale, org[a]), hi = std::fma((float) ((qh[a] >> (8 * i)) & 0xffu), scale, org[a]);
                    if (kb.lo[a] < lo || kb.hi[a] > hi) AM_THROW("bvh q4: geometry under slot %d of node %d leaves its decoded box on axis %d", i, ref, a);
                    tb.lo[a] = std::min(tb.lo[a], kb.lo[a]); tb.hi[a] = std::max(tb.hi[a], kb.hi[a]);
                }
            }
            return tb;
        };
        rec(0, 0);
        for (size_t i = 0; i < seen.size(); ++i) if (!seen[i]) AM_THROW("bvh q4: triangle %zu is not referenced by any leaf", i);
        if (max_depth) *max_depth = maxd; if (n_leaves) *n_leaves = leaves; return AMVPT_OK; });
}
int amvpt_bvh_validate(amvpt_scene *h, uint32_t *max_depth, uint32_t *n_leaves) {
    return guard([&] { NEED(h); Scene &s = *h->s; s.prepare_host();
        struct B { float lo[3], hi[3]; };
        uint32_t leaves = 0, maxd = 0; std::vector<uint8_t> seen(s.bvh.tris.size() / 3, 0);
        std::function<void(int32_t, const B *, uint32_t)> rec = [&](int32_t ref, const B *box, uint32_t depth) {
            maxd = std::max(maxd, depth);
            if (ref < 0) {
                uint32_t code = ~(uint32_t) ref, first = code >> 3, count = code & 7u; leaves++;
                for (uint32_t i = first; i < first + count; ++i) {
                    if (seen[i]++) AM_THROW("bvh: triangle %u referenced twice", i);
                    for (int k = 0; k < 3; ++k) { const float4 &v = s.bvh.tris[3 * (size_t) i + k]; float p[3] = { v.x, v.y, v.z };
                        for (int a = 0; a < 3; ++a) if (box && (p[a] < box->lo[a] || p[a] > box->hi[a])) AM_THROW("bvh: triangle %u outside its leaf box", i); }
                }
                return;
            }
#if AM_BVH_WIDTH == 4
            const float4 *n = &s.bvh.nodes[8 * (size_t) ref];
            const float *lo[3] = { &n[0].x, &n[2].x, &n[4].x }, *hi[3] = { &n[1].x, &n[3].x, &n[5].x };
            int32_t c[4]; memcpy(c, &n[6].x, 16);
            for (int i = 0; i < 4; ++i) {
                if (c[i] == 0) continue;                 // empty slot
                if (c[i] > 0 && c[i] <= ref) AM_THROW("bvh: child %d of node %d is not numbered after its parent", c[i], ref);
                B b = { { lo[0][i], lo[1][i], lo[2][i] }, { hi[0][i], hi[1][i], hi[2][i] } };
                if (box) for (int k = 0; k < 3; ++k) if (b.lo[k] < box->lo[k] || b.hi[k] > box->hi[k]) AM_THROW("bvh: child box outside parent at node %d", ref);
                rec(c[i], &b, depth + 1);
            }
#else
            const float4 *n = &s.bvh.nodes[4 * (size_t) ref];
#if AM_FFMA2
            B a = { { n[0].x, n[0].y, n[2].x }, { n[0].z, n[0].w, n[2].y } }, b = { { n[1].x, n[1].y, n[2].z }, { n[1].z, n[1].w, n[2].w } };
#else
            B a = { { n[0].x, n[0].z, n[2].x }, { n[0].y, n[0].w, n[2].y } }, b = { { n[1].x, n[1].z, n[2].z }, { n[1].y, n[1].w, n[2].w } };
#endif
            int32_t ca, cb; memcpy(&ca, &n[3].x, 4); memcpy(&cb, &n[3].y, 4);
            bool a_empty = ca == -1, b_empty = cb == -1;
            if (box) for (int k = 0; k < 3; ++k) {
                if (!a_empty && (a.lo[k] < box->lo[k] || a.hi[k] > box->hi[k])) AM_THROW("bvh: child box outside parent at node %d", ref);
                if (!b_empty && (b.lo[k] < box->lo[k] || b.hi[k] > box->hi[k])) AM_THROW("bvh: child box outside parent at node %d", ref);
            }
            rec(ca, &a, depth + 1); rec(cb, &b, depth + 1);
#endif
        };
        rec(0, nullptr, 0);
        for (size_t i = 0; i < seen.size(); ++i) if (!seen[i]) AM_THROW("bvh: triangle %zu not referenced", i);
        if (max_depth) *max_depth = maxd; if (n_leaves) *n_leaves = leaves; return AMVPT_OK; });
}

} // extern "C"
