// integration/mvpath_b200.cpp -- the plugin a maintainer of xacond00/mitsuba3-amvpt adds to the reference tree (as
// src/integrators/mvpath_b200.cpp, registered in src/integrators/CMakeLists.txt like mvpath, linked with -lamvpt_b200) so that
// <integrator type="mvpath_b200"> renders through this repository's C ABI (include/amvpt.h) while XML loading, Scene, Film, CLI and Python stay
// the reference's.  It is written against the reference's headers only (no header of this repository other than amvpt.h); tests/test_cpu_host.py
// runs `g++ -std=c++17 -fsyntax-only` on it against those headers where /root/reference is mounted.
//
// Data flow: Scene::shapes() -> Mesh buffers -> amvpt_scene_add_mesh; BSDF / emitter parameters are read through Object::traverse (the
// mechanism mi.traverse uses); the MultiSensor's sub-cameras (MultiSensor::sensors(), grid_dim(), reverse_axis(), src/sensors/grid.cpp:435-457)
// become one table of sample_to_camera / to_world matrices -> amvpt_scene_set_sensor_table, so the quilt layout is the sensor's own; the RGBW
// film comes back through amvpt_film_raw into the reference's ImageBlock (Film::put_block), after a size check.
#include <mitsuba/core/properties.h>
#include <mitsuba/core/rfilter.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/mesh.h>
#include <mitsuba/render/sampler.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/texture.h>
#include <amvpt.h>

#include <cstring>
#include <map>
#include <string>
#include <typeinfo>
#include <vector>

NAMESPACE_BEGIN(mitsuba)

// Collects what Object::traverse() reports: scalar / colour parameters by name and referenced objects (textures).
template <typename Float, typename Spectrum> struct ParamCollector : TraversalCallback {
    MI_IMPORT_TYPES(Texture)
    using ScalarColor3f_ = Color<dr::scalar_t<Float>, 3>;
    std::map<std::string, float> scalars; std::map<std::string, std::array<float, 3>> colors; std::map<std::string, Object *> objects;
    void put_object(const std::string &name, Object *obj, uint32_t) override { objects[name] = obj; }
    void put_parameter_impl(const std::string &name, void *ptr, uint32_t, const std::type_info &type) override {
        if (type == typeid(float)) scalars[name] = *(float *) ptr;
        else if (type == typeid(double)) scalars[name] = (float) *(double *) ptr;
        else if (type == typeid(Float)) scalars[name] = (float) dr::slice(*(Float *) ptr, 0);
        else if (type == typeid(bool)) scalars[name] = *(bool *) ptr ? 1.f : 0.f;
        else if (type == typeid(ScalarColor3f_)) { auto &c = *(ScalarColor3f_ *) ptr; colors[name] = { (float) c[0], (float) c[1], (float) c[2] }; }
        else if (type == typeid(Color<Float, 3>)) { auto &c = *(Color<Float, 3> *) ptr; colors[name] = { (float) dr::slice(c[0], 0), (float) dr::slice(c[1], 0), (float) dr::slice(c[2], 0) }; }
    }
    // a constant texture (`uniform`, `srgb`) referenced under `name` -> RGB; textured parameters are outside this path
    bool color(const std::string &name, float *rgb) {
        auto it = objects.find(name);
        if (it == objects.end()) { auto c = colors.find(name); if (c == colors.end()) return false; for (int k = 0; k < 3; ++k) rgb[k] = c->second[k]; return true; }
        ParamCollector sub; it->second->traverse(&sub);
        if (auto c = sub.colors.find("color"); c != sub.colors.end()) { for (int k = 0; k < 3; ++k) rgb[k] = c->second[k]; return true; }      // srgb.cpp
        if (auto v = sub.scalars.find("value"); v != sub.scalars.end()) { rgb[0] = rgb[1] = rgb[2] = v->second; return true; }                  // uniform.cpp
        Throw("mvpath_b200: parameter \"%s\" is not a constant colour (textures are not on this path)", name);
    }
    float scalar(const std::string &name, float def) { auto it = scalars.find(name); if (it != scalars.end()) return it->second;
        float c[3]; if (color(name, c)) return c[0]; return def; }
};

template <typename Float, typename Spectrum>
class MVPathB200 final : public MonteCarloIntegrator<Float, Spectrum> {
public:
    MI_IMPORT_BASE(MonteCarloIntegrator, m_max_depth, m_rr_depth, m_hide_emitters, m_timeout)
    MI_IMPORT_TYPES(Scene, Sensor, Film, ImageBlock, Mesh, BSDF, Emitter, Sampler, Shape, ReconstructionFilter)
    using Collector = ParamCollector<Float, Spectrum>;

    MVPathB200(const Properties &props) : Base(props) {
        m_spp_pass_lim = props.get<uint32_t>("spp_pass_lim", 16);                      // src/integrators/mvpath.h:120-129
        m_sa_reuse = props.get<bool>("sa_reuse", false); m_sa_mis = props.get<bool>("sa_mis", false);
        m_fast_mis = props.get<bool>("fast_mis", false); m_reuse_count = props.get<uint32_t>("reuse_count", 1);
        m_adaptive = props.get<uint32_t>("adaptive", 0); m_debug = props.get<bool>("debug", false);
        m_device = (int) props.get<int64_t>("device", 0);
    }

    TensorXf render(Scene *scene, Sensor *sensor, UInt32 seed, uint32_t spp, bool develop, bool /*evaluate*/) override {
        auto *ms = dynamic_cast<MultiSensor<Float, Spectrum> *>(sensor);
        if (!ms) Throw("This integrator can only be used with MultiSensor !");           // src/integrators/mvpath.cpp:15-17
        amvpt_scene *h = nullptr;
        check(amvpt_scene_create(&h));
        struct Guard { amvpt_scene *h; ~Guard() { amvpt_scene_destroy(h); } } guard{ h };
        std::map<const BSDF *, int> bsdf_index;
        for (auto &shape : scene->shapes()) {
            auto *mesh = dynamic_cast<Mesh *>(shape.get());
            if (!mesh) Throw("mvpath_b200 handles triangle meshes only (shape \"%s\")", shape->id());
            const BSDF *bsdf = mesh->bsdf(); auto it = bsdf_index.find(bsdf);
            if (it == bsdf_index.end()) { amvpt_bsdf_desc d = describe(bsdf); it = bsdf_index.emplace(bsdf, check(amvpt_scene_add_bsdf(h, &d))).first; }
            float radiance[3]; const bool emits = mesh->is_emitter();
            if (emits) { Collector c; const_cast<Emitter *>(mesh->emitter())->traverse(&c); if (!c.color("radiance", radiance)) Throw("mvpath_b200: area emitter without a constant radiance"); }
            check(amvpt_scene_add_mesh(h, mesh->vertex_positions_buffer().data(), (uint32_t) mesh->vertex_count(),
                                       (const uint32_t *) mesh->faces_buffer().data(), (uint32_t) mesh->face_count(),
                                       mesh->has_vertex_normals() ? mesh->vertex_normals_buffer().data() : nullptr,
                                       it->second, emits ? radiance : nullptr, 0 /* Mesh applies flip_normals to its normal buffers at load time */));
        }
        for (auto &e : scene->emitters()) {                                                // emitters that are not attached to a shape
            if (has_flag(e->flags(), EmitterFlags::Surface)) continue;
            Collector c; const_cast<Emitter *>(e.get())->traverse(&c); float rgb[3];
            const std::string cls = e->class_()->name();
            if (cls == "PointLight") { float pos[3] = { 0, 0, 0 }; if (auto p = c.colors.find("position"); p != c.colors.end()) for (int k = 0; k < 3; ++k) pos[k] = p->second[k];
                c.color("intensity", rgb); check(amvpt_scene_add_emitter_point(h, pos, rgb)); }
            else if (cls == "ConstantBackgroundEmitter") { c.color("radiance", rgb); check(amvpt_scene_add_emitter_constant(h, rgb)); }
            else Throw("mvpath_b200: emitter \"%s\" is handed over through the XML loader of the library (amvpt_scene_from_xml), not through this shim", cls);
        }
        // ---- the MultiSensor's sub-cameras as a table, in sensor-index order; the quilt layout (grid, axis reversal) is the sensor's own
        const auto subs = ((const MultiSensor<Float, Spectrum> *) ms)->sensors();
        const ScalarVector2u grid = ms->grid_dim(); const auto [rev_x, rev_y] = ms->reverse_axis();
        const ScalarVector2u film_size = sensor->film()->size(), view = film_size / grid;
        if (subs.size() != (size_t) grid.x() * grid.y()) Throw("mvpath_b200: %u sub-sensors do not fill a %u x %u grid", (uint32_t) subs.size(), grid.x(), grid.y());
        std::vector<float> s2c(16 * subs.size()), tw(16 * subs.size()), sc(4 * subs.size()); bool thinlens = false;
        for (size_t i = 0; i < subs.size(); ++i) {
            Collector c; const_cast<Sensor *>(subs[i])->traverse(&c);
            const float x_fov = c.scalar("x_fov", 45.f), near_ = c.scalar("near_clip", 1e-2f), far_ = c.scalar("far_clip", 1e4f);
            const float aperture = c.scalar("aperture_radius", 0.f), focus = c.scalar("focus_distance", far_);
            thinlens |= aperture > 0.f;
            // ProjectiveCamera::update_camera_transforms (src/sensors/perspective.cpp:173-182): sample_to_camera = perspective_projection(...)^-1
            const auto c2s = perspective_projection(ScalarVector2i(view), ScalarVector2i(view), ScalarVector2i(0), x_fov, near_, far_);
            const auto m = c2s.inverse().matrix; const auto w = subs[i]->world_transform().matrix;
            for (int r = 0; r < 4; ++r) for (int k = 0; k < 4; ++k) { s2c[16 * i + 4 * r + k] = (float) m.entry(r, k); tw[16 * i + 4 * r + k] = (float) dr::slice(w.entry(r, k), 0); }
            sc[4 * i] = near_; sc[4 * i + 1] = far_; sc[4 * i + 2] = aperture; sc[4 * i + 3] = focus;
        }
        check(amvpt_scene_set_sensor_table(h, grid.x(), grid.y(), view.x(), view.y(), rev_x, rev_y, thinlens, s2c.data(), tw.data(), sc.data()));
        amvpt_integrator_desc id{};
        id.max_depth = m_max_depth == (uint32_t) -1 ? -1 : (int64_t) m_max_depth; id.rr_depth = m_rr_depth; id.hide_emitters = m_hide_emitters;
        id.spp_pass_lim = m_spp_pass_lim; id.sample_count = sensor->sampler()->sample_count();
        { Collector c; sensor->sampler()->traverse(&c); id.sampler_seed = (uint32_t) c.scalar("seed", 0.f); }      // Sampler keeps `seed` as m_base_seed (no accessor): 0 unless the plugin exposes it
        check(amvpt_scene_set_integrator(h, &id));
        check(amvpt_scene_set_timeout(h, (float) m_timeout));
        check(amvpt_scene_set_reuse(h, m_sa_reuse, m_sa_mis, m_fast_mis, m_reuse_count, m_adaptive, nullptr));
        check(amvpt_scene_set_reuse_debug(h, m_debug));
        int ftype; float fp, fp2; rfilter_of(sensor->film()->rfilter(), &ftype, &fp, &fp2);
        check(amvpt_scene_set_rfilter(h, ftype, fp, fp2));
        check(amvpt_scene_commit(h, m_device));
        // ---- the hot path
        amvpt_render_desc rd{}; rd.seed = (uint32_t) dr::slice(seed, 0); rd.spp = spp; rd.scheduler = 4 /* auto */; rd.world = 1;
        check(amvpt_render(h, &rd, nullptr));
        // ---- RGBW film back into the reference's Film (Film::prepare, ImageBlock, Film::put_block: src/films/hdrfilm.cpp:245-297)
        uint32_t fw = 0, fh = 0; check(amvpt_scene_info(h, &fw, &fh, nullptr, nullptr, nullptr, nullptr));
        Film *film = sensor->film(); film->prepare({});
        ref<ImageBlock> block = film->create_block();
        if (fw != film_size.x() || fh != film_size.y() || block->channel_count() != 4 || block->border_size() != 0)
            Throw("mvpath_b200: film layout mismatch (library %u x %u x 4, film %u x %u x %u, border %u)", fw, fh, film_size.x(), film_size.y(), (uint32_t) block->channel_count(), block->border_size());
        std::vector<float> rgbw((size_t) fw * fh * 4); check(amvpt_film_raw(h, rgbw.data()));
        block->tensor() = TensorXf(rgbw.data(), 3, block->tensor().shape().data());
        film->put_block(block);
        return develop ? film->develop() : TensorXf();
    }

    MI_DECLARE_CLASS()
private:
    static int check(int rc) { if (rc < 0) Throw("%s", amvpt_last_error()); return rc; }

    // BSDF plugin -> amvpt_bsdf_desc: the class name selects the type, traverse() gives the parameters (names as in src/bsdfs/*.cpp)
    static amvpt_bsdf_desc describe(const BSDF *bsdf_) {
        BSDF *bsdf = const_cast<BSDF *>(bsdf_); amvpt_bsdf_desc d{}; std::string cls = bsdf->class_()->name();
        Collector outer; bsdf->traverse(&outer);
        if (cls == "TwoSidedBRDF") {                              // src/bsdfs/twosided.cpp: one nested BRDF on both sides
            d.twosided = 1; auto it = outer.objects.find("brdf_0"); if (it == outer.objects.end()) Throw("mvpath_b200: twosided without a nested BSDF");
            bsdf = dynamic_cast<BSDF *>(it->second); cls = bsdf->class_()->name(); outer = Collector(); bsdf->traverse(&outer);
        }
        for (int k = 0; k < 3; ++k) { d.reflectance[k] = .5f; d.k[k] = 1.f; d.specular_reflectance[k] = 1.f; d.specular_transmittance[k] = 1.f; }
        d.int_ior = 1.5046f; d.ext_ior = 1.000277f; d.alpha_u = d.alpha_v = .1f;
        const bool ggx = outer.scalars.count("distribution") ? outer.scalars["distribution"] != 0.f : false; d.distr = ggx ? 1 : 0;
        if (cls == "SmoothDiffuse") { d.type = 0; outer.color("reflectance", d.reflectance); }
        else if (cls == "SmoothConductor" || cls == "RoughConductor") { d.type = cls == "SmoothConductor" ? 1 : 3; outer.color("eta", d.eta); outer.color("k", d.k); outer.color("specular_reflectance", d.specular_reflectance); }
        else if (cls == "SmoothDielectric" || cls == "RoughDielectric") { d.type = cls == "SmoothDielectric" ? 2 : 4; d.int_ior = outer.scalar("eta", 1.5046f / 1.000277f); d.ext_ior = 1.f;
            outer.color("specular_reflectance", d.specular_reflectance); outer.color("specular_transmittance", d.specular_transmittance); }
        else if (cls == "RoughPlastic") { d.type = 5; d.int_ior = outer.scalar("eta", 1.49f / 1.000277f); d.ext_ior = 1.f; d.nonlinear = outer.scalar("nonlinear", 0.f) != 0.f;
            outer.color("diffuse_reflectance", d.reflectance); outer.color("specular_reflectance", d.specular_reflectance); }
        else Throw("mvpath_b200: BSDF \"%s\" is not part of this path", cls);
        if (d.type >= 3) { const float a = outer.scalar("alpha", .1f); d.alpha_u = outer.scalar("alpha_u", a); d.alpha_v = outer.scalar("alpha_v", a); }
        return d;
    }
    // ReconstructionFilter plugin -> (type, param, param2) of amvpt_scene_set_rfilter
    static void rfilter_of(const ReconstructionFilter *rf, int *type, float *p, float *p2) {
        const std::string cls = rf->class_()->name(); *p2 = 0.f;
        if (cls == "BoxFilter") { *type = 0; *p = .5f; }
        else if (cls == "GaussianFilter") { *type = 1; *p = (float) rf->radius() / 4.f; }              // radius = 4 stddev (src/rfilters/gaussian.cpp:44)
        else if (cls == "TentFilter") { *type = 2; *p = (float) rf->radius(); }
        else if (cls == "MitchellNetravaliFilter") { *type = 3; *p = *p2 = 1.f / 3.f; }                // B, C are not exposed: the plugin's defaults
        else if (cls == "CatmullRomFilter") { *type = 4; *p = 0.f; }
        else if (cls == "LanczosSincFilter") { *type = 5; *p = (float) rf->radius(); }
        else Throw("mvpath_b200: reconstruction filter \"%s\" is not part of this path", cls);
    }

    uint32_t m_spp_pass_lim, m_reuse_count, m_adaptive; bool m_sa_reuse, m_sa_mis, m_fast_mis, m_debug; int m_device;
};

MI_IMPLEMENT_CLASS_VARIANT(MVPathB200, MonteCarloIntegrator)
MI_EXPORT_PLUGIN(MVPathB200, "B200 multi-view path tracer (libamvpt_b200)")
NAMESPACE_END(mitsuba)
