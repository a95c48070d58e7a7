"""mitsuba3-amvpt_b200 -- Python binding of the B200-native AMVPT render path (ctypes over include/amvpt.h).

The product is the C++/CUDA library ``libamvpt_b200.so``; this module only mirrors the part of the reference's
Python surface that drives the path (``load_file`` / ``Scene`` / ``integrator.render(scene, seed=, spp=, develop=)``,
reference: src/render/python/integrator_v.cpp, src/python/python/util.py) so that tests and bench read like the
reference's own.  There is no CPU fallback: without the built extension the import fails, and every compute call
fails without a CUDA device.
"""
import ctypes as C
import importlib
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("AMVPT_LIB", os.path.join(_HERE, "libamvpt_b200.so"))


class AmvptError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the CUDA extension is the product; there is no fallback path)")
    return C.CDLL(LIB_PATH)


lib = _load()
f32p, u32p, u64p, u8p = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint8)


class Preset(C.Structure):      # amvpt_preset (include/amvpt.h); src/mitsuba/preset.h:12-17
    _fields_ = [("center", C.c_float), ("focus", C.c_float), ("pitch", C.c_float), ("tilt", C.c_float), ("subp", C.c_float), ("view", C.c_float * 2),
                ("grid", C.c_int32 * 2), ("flip_rgb", C.c_int32), ("quilt", C.c_int32), ("nearest", C.c_int32)]


class BsdfDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("twosided", C.c_int32), ("distr", C.c_int32), ("nonlinear", C.c_int32),
                ("reflectance", C.c_float * 3), ("eta", C.c_float * 3), ("k", C.c_float * 3),
                ("specular_reflectance", C.c_float * 3), ("specular_transmittance", C.c_float * 3),
                ("int_ior", C.c_float), ("ext_ior", C.c_float), ("alpha_u", C.c_float), ("alpha_v", C.c_float)]


class GridSensorDesc(C.Structure):
    _fields_ = [("thinlens", C.c_int32), ("fov_axis", C.c_int32), ("grid_x", C.c_uint32), ("grid_y", C.c_uint32),
                ("res_x", C.c_uint32), ("res_y", C.c_uint32), ("reverse_x", C.c_int32), ("reverse_y", C.c_int32),
                ("use_cone", C.c_int32), ("cam_center", C.c_int32), ("to_world", C.c_double * 16), ("fov", C.c_double),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("focus_distance", C.c_float), ("aperture_radius", C.c_float),
                ("cone_deg", C.c_float), ("cam_dist", C.c_float), ("cam_dir", C.c_float * 3), ("cam_off", C.c_float * 3)]


class IntegratorDesc(C.Structure):
    _fields_ = [("max_depth", C.c_int64), ("rr_depth", C.c_uint32), ("hide_emitters", C.c_int32), ("spp_pass_lim", C.c_uint32),
                ("sampler_seed", C.c_uint32), ("sample_count", C.c_uint32)]


ALLREDUCE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_uint32, C.c_void_p)


class RenderDesc(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("spp", C.c_uint32), ("scheduler", C.c_int32), ("collect_stats", C.c_int32),
                ("rank", C.c_uint32), ("world", C.c_uint32), ("adaptive", C.c_int32), ("adaptive_threshold", C.c_float),
                ("adaptive_min_spp", C.c_uint32), ("stream", C.c_void_p), ("allreduce", ALLREDUCE_FN), ("allreduce_user", C.c_void_p), ("comm", C.c_void_p)]


class RenderStats(C.Structure):
    _fields_ = [("seconds", C.c_double), ("paths", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("bvh_nodes", C.c_uint64), ("tri_tests", C.c_uint64), ("path_vertices", C.c_uint64),
                ("passes", C.c_uint32), ("launches", C.c_uint32), ("spp_done", C.c_uint32), ("scheduler_used", C.c_uint32),
                ("trace_seconds", C.c_double), ("trace_launches", C.c_uint32), ("pad", C.c_uint32)]


lib.amvpt_last_error.restype = C.c_char_p
EXPORTS = ["amvpt_last_error", "amvpt_device_count", "amvpt_scene_create", "amvpt_scene_destroy", "amvpt_scene_from_xml",
           "amvpt_scene_from_xml_string", "amvpt_scene_add_bsdf", "amvpt_scene_add_mesh", "amvpt_scene_set_sensor_grid",
           "amvpt_scene_set_sensor_batch", "amvpt_scene_set_sensor_table", "amvpt_scene_add_emitter_point", "amvpt_scene_add_emitter_constant", "amvpt_scene_add_emitter_envmap", "amvpt_scene_get_emitter", "amvpt_scene_get_sensor_params", "amvpt_scene_get_envmap", "amvpt_scene_get_shutter_draw", "amvpt_scene_set_shutter", "amvpt_env_warp", "amvpt_image_read", "amvpt_spectrum_to_srgb", "amvpt_scene_set_integrator", "amvpt_scene_set_rfilter", "amvpt_scene_set_reuse", "amvpt_scene_set_reuse_debug", "amvpt_scene_set_timeout", "amvpt_last_fill_paths", "amvpt_scene_commit", "amvpt_scene_info",
           "amvpt_scene_get_sensor", "amvpt_scene_get_bsdf_tables", "amvpt_scene_counts", "amvpt_scene_get_mesh",
           "amvpt_scene_get_bsdf", "amvpt_scene_get_integrator", "amvpt_render", "amvpt_shard_tiles", "amvpt_view_tiles", "amvpt_adaptive_alloc", "amvpt_comm_unique_id", "amvpt_comm_init", "amvpt_comm_destroy", "amvpt_film_gather", "amvpt_comm_allreduce_sum", "amvpt_gather_tiles", "amvpt_film_develop_pinned", "amvpt_cancel", "amvpt_film_clear",
           "amvpt_film_raw", "amvpt_film_develop", "amvpt_film_develop_device", "amvpt_film_device_ptr", "amvpt_film_add_host",
           "amvpt_trace_lanes", "amvpt_trace_lanes_multi", "amvpt_intersect", "amvpt_intersect_static", "amvpt_bsdf_eval_pdf_sample", "amvpt_math1", "amvpt_sampler_stream",
           "amvpt_bvh_validate", "amvpt_bvh_checksum", "amvpt_bvh_q4_validate", "amvpt_membench", "amvpt_film_write_pfm", "amvpt_film_write_exr", "amvpt_f32_to_f16", "amvpt_film_to_srgb8", "amvpt_interlace", "amvpt_display_last_ms"]
lib.amvpt_scene_destroy.restype = None
lib.amvpt_scene_destroy.argtypes = [C.c_void_p]
lib.amvpt_comm_destroy.restype = None
lib.amvpt_comm_destroy.argtypes = [C.c_void_p]


def _chk(rc):
    if rc < 0:
        raise AmvptError(lib.amvpt_last_error().decode())
    return rc


def _p(a, t=f32p):
    return None if a is None else a.ctypes.data_as(t)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if shape is None else a.reshape(shape)


BSDF_TYPES = {"diffuse": 0, "conductor": 1, "dielectric": 2, "roughconductor": 3, "roughdielectric": 4, "roughplastic": 5}
FOV_AXES = {"x": 0, "y": 1, "diagonal": 2, "smaller": 3, "larger": 4}
RFILTERS = {"box": 0, "gaussian": 1, "tent": 2, "mitchell": 3, "catmullrom": 4, "lanczos": 5}
SCHEDULERS = {"warp_batch": 0, "regenerate": 1, "stream": 2, "wavefront": 3, "auto": 4, "split": 6}
SCHEDULER_NAMES = {0: "warp_batch", 1: "regenerate", 2: "stream", 3: "wavefront", 5: "reuse", 6: "split"}


def bsdf_desc(type, twosided=False, distribution="beckmann", nonlinear=False, reflectance=(.5, .5, .5), eta=(0., 0., 0.),
              k=(1., 1., 1.), specular_reflectance=(1., 1., 1.), specular_transmittance=(1., 1., 1.), int_ior=1.5046,
              ext_ior=1.000277, alpha=0.1, alpha_u=None, alpha_v=None):
    """Parameter names and defaults follow the reference's BSDF plugin constructors (src/bsdfs/*.cpp)."""
    d = BsdfDesc()
    d.type = BSDF_TYPES[type]; d.twosided = int(twosided); d.distr = {"beckmann": 0, "ggx": 1}[distribution]; d.nonlinear = int(nonlinear)
    for name, val in (("reflectance", reflectance), ("eta", eta), ("k", k), ("specular_reflectance", specular_reflectance),
                      ("specular_transmittance", specular_transmittance)):
        v = np.broadcast_to(np.asarray(val, dtype=np.float32), (3,))
        setattr(d, name, (C.c_float * 3)(*v))
    d.int_ior = int_ior; d.ext_ior = ext_ior
    d.alpha_u = alpha if alpha_u is None else alpha_u; d.alpha_v = alpha if alpha_v is None else alpha_v
    return d


class Scene:
    """Scene + MultiSensor + Film + Sampler + Integrator of one render job (the objects Scene::Scene collects in
    src/render/scene.cpp:22-96).  Built either from XML (load_file / load_string) or programmatically."""

    def __init__(self, handle=None):
        if handle is None:
            handle = C.c_void_p()
            _chk(lib.amvpt_scene_create(C.byref(handle)))
        self._h = handle
        self._keep = []

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:          # `lib` is already gone when the interpreter shuts down
            lib.amvpt_scene_destroy(h)

    # ---- construction
    def add_bsdf(self, desc=None, **kw):
        desc = desc if desc is not None else bsdf_desc(**kw)
        return _chk(lib.amvpt_scene_add_bsdf(self._h, C.byref(desc)))

    def add_mesh(self, positions, faces, bsdf, normals=None, radiance=None, flip_normals=False):
        pos = _f32(positions, (-1, 3)); fc = np.ascontiguousarray(faces, dtype=np.uint32).reshape(-1, 3)
        nrm = None if normals is None else _f32(normals, (-1, 3))
        rad = None if radiance is None else _f32(np.broadcast_to(np.asarray(radiance, dtype=np.float32), (3,)))
        return _chk(lib.amvpt_scene_add_mesh(self._h, _p(pos), C.c_uint32(len(pos)), _p(fc, u32p), C.c_uint32(len(fc)), _p(nrm),
                                             C.c_int(bsdf), _p(rad), C.c_int(int(flip_normals))))

    def add_emitter_point(self, position, intensity):
        """PointLight (src/emitters/point.cpp): `position`, `intensity` (RGB)."""
        return _chk(lib.amvpt_scene_add_emitter_point(self._h, _p(_f32(position, (3,))), _p(_f32(np.broadcast_to(np.asarray(intensity, np.float32), (3,))))))

    def add_emitter_constant(self, radiance=1.0):
        """ConstantBackgroundEmitter (src/emitters/constant.cpp)."""
        return _chk(lib.amvpt_scene_add_emitter_constant(self._h, _p(_f32(np.broadcast_to(np.asarray(radiance, np.float32), (3,))))))

    def add_emitter_envmap(self, rgb, scale=1.0, to_world=None):
        """EnvironmentMapEmitter (src/emitters/envmap.cpp): rgb = (H, W, 3) linear radiance, latitude-longitude."""
        img = _f32(rgb); h, w = img.shape[:2]; tw = None if to_world is None else _f32(to_world, (16,))
        return _chk(lib.amvpt_scene_add_emitter_envmap(self._h, _p(img), C.c_uint32(w), C.c_uint32(h), C.c_float(scale), _p(tw)))

    def set_shutter(self, shutter_open=0.0, shutter_close=0.0):
        _chk(lib.amvpt_scene_set_shutter(self._h, C.c_float(shutter_open), C.c_float(shutter_close)))

    def emitters(self):
        """The scene's emitter list: dicts with type ('area' | 'point' | 'constant' | 'envmap'), mesh, radiance, position."""
        out = []
        for i in range(self.counts()["n_emitters"]):
            t, m = C.c_int32(), C.c_int32(); rad = np.zeros(3, np.float32); pos = np.zeros(3, np.float32)
            _chk(lib.amvpt_scene_get_emitter(self._h, C.c_uint32(i), C.byref(t), C.byref(m), _p(rad), _p(pos)))
            out.append(dict(type=("area", "point", "constant", "envmap")[t.value], mesh=m.value, radiance=rad, position=pos))
        return out

    def envmap(self):
        w, h, sc = C.c_uint32(), C.c_uint32(), C.c_float(); tw = np.zeros(16, np.float32)
        _chk(lib.amvpt_scene_get_envmap(self._h, C.byref(w), C.byref(h), C.byref(sc), _p(tw), None))
        if w.value == 0:
            return None
        rgb = np.zeros((h.value, w.value, 3), np.float32)
        _chk(lib.amvpt_scene_get_envmap(self._h, C.byref(w), C.byref(h), C.byref(sc), _p(tw), _p(rgb)))
        return dict(rgb=rgb, scale=sc.value, to_world=tw.reshape(4, 4))

    def shutter_draw(self):
        d = C.c_int32(); _chk(lib.amvpt_scene_get_shutter_draw(self._h, C.byref(d))); return bool(d.value)

    def env_warp(self, u):
        u = _f32(u, (-1, 2)); n = len(u); uv = np.zeros((n, 2), np.float32); pdf = np.zeros(n, np.float32); ev = np.zeros(n, np.float32)
        _chk(lib.amvpt_env_warp(self._h, _p(u), C.c_uint32(n), _p(uv), _p(pdf), _p(ev))); return uv, pdf, ev

    def set_sensor_grid(self, grid_x, grid_y, res_x, res_y, to_world, fov, fov_axis="x", near_clip=1e-2, far_clip=1e4,
                        focus_distance=1.0, thinlens=False, aperture_radius=0.0, reverse_x=False, reverse_y=True, cone_deg=None,
                        cam_dir=(1., 0., 0.), cam_dist=None, cam_center=True, cam_off=(0., 0., 0.)):
        """`grid` sensor (src/sensors/grid.cpp:84-180); res_x / res_y are the quilt size."""
        d = GridSensorDesc()
        d.thinlens = int(thinlens); d.fov_axis = FOV_AXES[fov_axis]; d.grid_x = grid_x; d.grid_y = grid_y; d.res_x = res_x; d.res_y = res_y
        d.reverse_x = int(reverse_x); d.reverse_y = int(reverse_y); d.use_cone = int(cone_deg is not None); d.cam_center = int(cam_center)
        d.to_world = (C.c_double * 16)(*np.asarray(to_world, dtype=np.float64).reshape(16))
        d.fov = fov; d.near_clip = near_clip; d.far_clip = far_clip; d.focus_distance = focus_distance; d.aperture_radius = aperture_radius
        d.cone_deg = 0.0 if cone_deg is None else cone_deg
        cd = np.asarray(cam_dir, dtype=np.float32)
        d.cam_dist = float(np.float32(np.sqrt(np.float32(cd[0] * cd[0] + cd[1] * cd[1] + cd[2] * cd[2])))) if cam_dist is None else cam_dist
        d.cam_dir = (C.c_float * 3)(*cd); d.cam_off = (C.c_float * 3)(*np.asarray(cam_off, dtype=np.float32))
        _chk(lib.amvpt_scene_set_sensor_grid(self._h, C.byref(d)))

    def set_sensor_batch(self, res_x, res_y, sample_to_camera, to_world, near_clip, far_clip, reverse_x=False):
        s2c = _f32(sample_to_camera, (-1, 16)); tw = _f32(to_world, (-1, 16))
        _chk(lib.amvpt_scene_set_sensor_batch(self._h, C.c_uint32(len(s2c)), C.c_uint32(res_x), C.c_uint32(res_y), _p(s2c), _p(tw),
                                              C.c_float(near_clip), C.c_float(far_clip), C.c_int(int(reverse_x))))

    def set_integrator(self, max_depth=-1, rr_depth=5, hide_emitters=False, spp_pass_lim=16, sampler_seed=0, sample_count=0):
        d = IntegratorDesc(max_depth, rr_depth, int(hide_emitters), spp_pass_lim, sampler_seed, sample_count)
        _chk(lib.amvpt_scene_set_integrator(self._h, C.byref(d)))

    def set_reuse(self, sa_reuse=True, sa_mis=False, fast_mis=False, reuse_count=0, adaptive=0):
        """Cross-view sample reuse of the mvpath integrator (sa_reuse / sa_mis / fast_mis / reuse_count, mvpath.h:120-129).
        Returns the number of cameras per reuse group (0 = reuse off)."""
        n = C.c_uint32(0)
        _chk(lib.amvpt_scene_set_reuse(self._h, C.c_int(int(sa_reuse)), C.c_int(int(sa_mis)), C.c_int(int(fast_mis)), C.c_uint32(reuse_count),
                                       C.c_uint32(adaptive), C.byref(n)))
        self.n_reuse = n.value
        return n.value

    def set_timeout(self, seconds):
        """Integrator `timeout`: the render stops after the pass during which this many seconds have elapsed (should_stop(), integrator.h:303-306)."""
        _chk(lib.amvpt_scene_set_timeout(self._h, C.c_float(seconds)))

    def set_reuse_debug(self, debug=True):
        _chk(lib.amvpt_scene_set_reuse_debug(self._h, C.c_int(int(debug))))

    def last_fill_paths(self):
        n = C.c_uint64(); _chk(lib.amvpt_last_fill_paths(self._h, C.byref(n))); return n.value

    def trace_lanes_multi(self, seed_value, spp_per_pass, lane_begin, n):
        """Reuse-branch tap: (n, n_reuse, 9) = pos.xy, weight, result.rgb, valid, indirect, sensor index."""
        out = np.zeros((n, self.n_reuse, 9), np.float32)
        _chk(lib.amvpt_trace_lanes_multi(self._h, C.c_uint32(seed_value), C.c_uint32(spp_per_pass), C.c_uint32(lane_begin), C.c_uint32(n), _p(out)))
        return out

    def set_rfilter(self, type="box", param=0.5, param2=0.0):
        """Film reconstruction filter: 'box', 'gaussian' (param = stddev), 'tent' (param = radius), 'mitchell' (param = B,
        param2 = C), 'catmullrom' or 'lanczos' (param = lobes)."""
        _chk(lib.amvpt_scene_set_rfilter(self._h, C.c_int(RFILTERS[type]), C.c_float(param), C.c_float(param2)))

    def commit(self, device=0):
        _chk(lib.amvpt_scene_commit(self._h, C.c_int(device)))
        return self

    # ---- queries
    def info(self):
        v = [C.c_uint32() for _ in range(6)]
        _chk(lib.amvpt_scene_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("film_w", "film_h", "n_views", "n_prims", "n_bvh_nodes", "sample_count"), [x.value for x in v]))

    def sensor(self, index):
        s2c = np.zeros(16, np.float32); tw = np.zeros(16, np.float32)
        _chk(lib.amvpt_scene_get_sensor(self._h, C.c_uint32(index), _p(s2c), _p(tw)))
        return s2c.reshape(4, 4), tw.reshape(4, 4)

    def sensor_params(self, index=0):
        sc = np.zeros(4, np.float32); lay = np.zeros(5, np.int32)
        _chk(lib.amvpt_scene_get_sensor_params(self._h, C.c_uint32(index), _p(sc), lay.ctypes.data_as(C.POINTER(C.c_int32))))
        return dict(near_clip=float(sc[0]), far_clip=float(sc[1]), aperture_radius=float(sc[2]), focus_distance=float(sc[3]),
                    grid_x=int(lay[0]), grid_y=int(lay[1]), reverse_x=bool(lay[2]), reverse_y=bool(lay[3]), thinlens=bool(lay[4]))

    def bsdf_tables(self, bsdf):
        t = np.zeros(64, np.float32); s = np.zeros(3, np.float32)
        _chk(lib.amvpt_scene_get_bsdf_tables(self._h, C.c_int(bsdf), _p(t), _p(s)))
        return t, s

    # ---- read-back of the host description (what XML loading produced)
    def counts(self):
        a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _chk(lib.amvpt_scene_counts(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(n_meshes=a.value, n_bsdfs=b.value, n_emitters=c.value)

    def mesh(self, index):
        nv, nf, bsdf, hn, fl, em = C.c_uint32(), C.c_uint32(), C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        rad = np.zeros(3, np.float32); null = C.c_void_p()
        _chk(lib.amvpt_scene_get_mesh(self._h, C.c_uint32(index), C.byref(nv), C.byref(nf), C.byref(bsdf), C.byref(hn), C.byref(fl), C.byref(em),
                                      _p(rad), null, null, null))
        pos = np.zeros((nv.value, 3), np.float32); faces = np.zeros((nf.value, 3), np.uint32)
        nrm = np.zeros((nv.value, 3), np.float32) if hn.value else None
        _chk(lib.amvpt_scene_get_mesh(self._h, C.c_uint32(index), None, None, None, None, None, None, None, _p(pos), _p(faces, u32p),
                                      _p(nrm) if nrm is not None else null))
        return dict(positions=pos, faces=faces, normals=nrm, bsdf=bsdf.value, flip_normals=bool(fl.value),
                    radiance=tuple(float(x) for x in rad) if em.value else None)

    def bsdf(self, index):
        d = BsdfDesc(); _chk(lib.amvpt_scene_get_bsdf(self._h, C.c_int(index), C.byref(d)))
        inv = {v: k for k, v in BSDF_TYPES.items()}
        return dict(type=inv[d.type], twosided=bool(d.twosided), distribution=("beckmann", "ggx")[d.distr], nonlinear=bool(d.nonlinear),
                    reflectance=tuple(d.reflectance), eta=tuple(d.eta), k=tuple(d.k), specular_reflectance=tuple(d.specular_reflectance),
                    specular_transmittance=tuple(d.specular_transmittance), int_ior=d.int_ior, ext_ior=d.ext_ior, alpha_u=d.alpha_u, alpha_v=d.alpha_v)

    def integrator(self):
        d = IntegratorDesc(); mv, ru, rc = C.c_int32(), C.c_int32(), C.c_uint32()
        _chk(lib.amvpt_scene_get_integrator(self._h, C.byref(d), C.byref(mv), C.byref(ru), C.byref(rc)))
        return dict(max_depth=d.max_depth, rr_depth=d.rr_depth, hide_emitters=bool(d.hide_emitters), spp_pass_lim=d.spp_pass_lim,
                    sampler_seed=d.sampler_seed, sample_count=d.sample_count, is_mvpath=bool(mv.value), sa_reuse=bool(ru.value), reuse_count=rc.value)

    def bvh_validate(self):
        d, n = C.c_uint32(), C.c_uint32()
        _chk(lib.amvpt_bvh_validate(self._h, C.byref(d), C.byref(n)))
        return d.value, n.value

    def bvh_checksum(self):
        x = C.c_uint64(); _chk(lib.amvpt_bvh_checksum(self._h, C.byref(x))); return x.value

    def bvh_q4_validate(self):
        n, d, l = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _chk(lib.amvpt_bvh_q4_validate(self._h, C.byref(n), C.byref(d), C.byref(l)))
        return n.value, d.value, l.value

    # ---- the hot path: Integrator.render(scene, sensor, seed, spp, develop, evaluate)
    def render(self, seed=0, spp=0, develop=True, scheduler="auto", stats=False, rank=0, world=1, adaptive=False,
               adaptive_threshold=0.0, adaptive_min_spp=16, stream=None, allreduce=None, clear=True, time_stages=False, comm=None):
        """Mirrors MVPathIntegrator::render (src/integrators/mvpath.cpp:7-278): accumulates `spp` samples per pixel into the
        film and returns the developed (H, W, 3) image if `develop` (else None).  `clear=False` keeps accumulating."""
        if clear:
            self.film_clear()
        d = RenderDesc()
        d.seed = seed; d.spp = spp; d.scheduler = SCHEDULERS[scheduler] if isinstance(scheduler, str) else scheduler
        d.collect_stats = int(stats) | (2 if time_stages else 0); d.rank = rank; d.world = world; d.adaptive = int(adaptive)
        d.adaptive_threshold = adaptive_threshold; d.adaptive_min_spp = adaptive_min_spp; d.stream = stream
        if comm is not None:
            d.comm = comm._h
        if allreduce is not None:
            cb = ALLREDUCE_FN(lambda dev, n, user: allreduce(dev, n)); self._keep = [cb]; d.allreduce = cb
        st = RenderStats()
        _chk(lib.amvpt_render(self._h, C.byref(d), C.byref(st)))
        self.last_stats = {k: getattr(st, k) for k, _ in RenderStats._fields_ if k != "pad"}
        return self.develop() if develop else None

    def shard_tiles(self, rank, world):
        """Rectangles (x0, y0, w, h, view) of the quilt that `rank` of `world` renders."""
        buf = np.zeros((4096, 5), np.uint32)
        n = _chk(lib.amvpt_shard_tiles(self._h, C.c_uint32(rank), C.c_uint32(world), _p(buf, u32p), C.c_uint32(len(buf))))
        return buf[:n].copy()

    def gather_tiles(self, rank, world):
        """Rectangles (x0, y0, w, h, staging offset) amvpt_film_gather moves for `rank`: its view tiles grown by one pixel."""
        buf = np.zeros((4096, 5), np.uint32)
        n = _chk(lib.amvpt_gather_tiles(self._h, C.c_uint32(rank), C.c_uint32(world), _p(buf, u32p), C.c_uint32(len(buf))))
        return buf[:n].copy()

    def view_tiles(self):
        """Rectangle (x0, y0, w, h, view) of every view of the quilt, by view index (the per-view split of write_pfm / write_exr)."""
        buf = np.zeros((65536, 5), np.uint32)
        n = _chk(lib.amvpt_view_tiles(self._h, _p(buf, u32p), C.c_uint32(len(buf))))
        return buf[:n].copy()

    def film_clear(self):
        _chk(lib.amvpt_film_clear(self._h))

    def film_raw(self):
        i = self.info(); out = np.zeros((i["film_h"], i["film_w"], 4), np.float32)
        _chk(lib.amvpt_film_raw(self._h, _p(out))); return out

    def develop(self):
        i = self.info(); out = np.zeros((i["film_h"], i["film_w"], 3), np.float32)
        _chk(lib.amvpt_film_develop(self._h, _p(out))); return out

    def film_device_ptr(self):
        a, b = C.c_void_p(), C.c_void_p()
        _chk(lib.amvpt_film_device_ptr(self._h, C.byref(a), C.byref(b))); return a.value, b.value

    def adaptive_alloc(self):
        """(passes, views) samples per pixel the adaptive allocator handed out in the last render."""
        a, b = C.c_uint32(), C.c_uint32(); _chk(lib.amvpt_adaptive_alloc(self._h, None, C.c_uint32(0), C.byref(a), C.byref(b)))
        out = np.zeros((a.value, b.value), np.uint32)
        if out.size:
            _chk(lib.amvpt_adaptive_alloc(self._h, _p(out, u32p), C.c_uint32(out.size), C.byref(a), C.byref(b)))
        return out

    def film_gather(self, comm, root=0, stream=None):
        """Rank-owned view tiles -> the root's film over NCCL (amvpt_film_gather); asynchronous on `stream`."""
        _chk(lib.amvpt_film_gather(self._h, comm._h, C.c_uint32(root), stream))

    def develop_pinned(self):
        """HDRFilm::develop into the library's page-locked host buffer (reused from call to call): (H, W, 3) view, valid until the next call."""
        p = C.POINTER(C.c_float)(); i = self.info()
        _chk(lib.amvpt_film_develop_pinned(self._h, C.byref(p)))
        return np.ctypeslib.as_array(p, shape=(i["film_h"], i["film_w"], 3))

    def film_add(self, rgbw):
        _chk(lib.amvpt_film_add_host(self._h, _p(_f32(rgbw))))

    # ---- output path and display step
    def write_pfm(self, path, view=-1):
        """HDRFilm::write, file_format = pfm (Bitmap::write_pfm, src/core/bitmap.cpp:2270-2299); view >= 0 writes that view's tile only."""
        _chk(lib.amvpt_film_write_pfm(self._h, str(path).encode(), C.c_int32(view)))

    def write_exr(self, path, view=-1, component_format="float16"):
        """HDRFilm::write, file_format = openexr (hdrfilm.cpp:420-560; component_format as hdrfilm.cpp:146-147, 191-196): RGB half or
        float, uncompressed scan lines; view >= 0 writes that view's tile only."""
        if component_format not in ("float16", "float32"):
            raise ValueError(f"component_format must be float16 or float32, not {component_format!r}")
        _chk(lib.amvpt_film_write_exr(self._h, str(path).encode(), C.c_int32(view), C.c_int32(component_format == "float16")))

    def to_srgb8(self, fetch=True):
        """The developed quilt as the viewer loads it: Bitmap::convert(RGB, UInt8, srgb) (src/mitsuba/program.h:55). Stays on the device for interlace()."""
        i = self.info(); out = np.zeros((i["film_h"], i["film_w"], 3), np.uint8) if fetch else None
        _chk(lib.amvpt_film_to_srgb8(self._h, _p(out, u8p))); return out

    def interlace(self, preset, dst_w, dst_h, dst_ch=3, src=None):
        """Program::display_image (src/mitsuba/program.cpp:199-272). preset: dict with center, focus, pitch, tilt, subp, view, grid and
        optionally flip_rgb, quilt, nearest (src/mitsuba/preset.h:12-17); src: optional (H, W, C >= 3) uint8 image instead of the film."""
        pr = Preset(center=preset["center"], focus=preset["focus"], pitch=preset["pitch"], tilt=preset["tilt"], subp=preset["subp"],
                    view=(C.c_float * 2)(*preset["view"]), grid=(C.c_int32 * 2)(*preset["grid"]), flip_rgb=int(preset.get("flip_rgb", dst_ch == 3)),
                    quilt=int(preset.get("quilt", False)), nearest=int(preset.get("nearest", False)))
        out = np.zeros((dst_h, dst_w, dst_ch), np.uint8)
        if src is not None:
            src = np.ascontiguousarray(src, np.uint8); sh, sw, sc = src.shape
        else:
            sh = sw = sc = 0
        _chk(lib.amvpt_interlace(self._h, _p(src, u8p), C.c_uint32(sw), C.c_uint32(sh), C.c_uint32(sc), C.byref(pr), C.c_uint32(dst_w), C.c_uint32(dst_h), C.c_uint32(dst_ch), _p(out, u8p)))
        return out

    def display_last_ms(self):
        a, b = C.c_float(), C.c_float(); _chk(lib.amvpt_display_last_ms(self._h, C.byref(a), C.byref(b))); return a.value, b.value

    # ---- parity taps
    def trace_lanes(self, seed_value, spp_per_pass, lane_begin, n, trace=True, max_hits=0):
        o = dict(sample_pos=np.zeros((n, 2), np.float32), ray_o=np.zeros((n, 3), np.float32), ray_d=np.zeros((n, 3), np.float32),
                 ray_maxt=np.zeros(n, np.float32), view=np.zeros(n, np.uint32), radiance=np.zeros((n, 4), np.float32),
                 pixel=np.zeros(n, np.uint32), hit_prim=np.zeros((n, max(max_hits, 1)), np.uint32), hit_t=np.zeros((n, max(max_hits, 1)), np.float32))
        stats = np.zeros(2, np.uint64)
        _chk(lib.amvpt_trace_lanes(self._h, C.c_uint32(seed_value), C.c_uint32(spp_per_pass), C.c_uint32(lane_begin), C.c_uint32(n), C.c_int(int(trace)),
                                   _p(o["sample_pos"]), _p(o["ray_o"]), _p(o["ray_d"]), _p(o["ray_maxt"]), _p(o["view"], u32p), _p(o["radiance"]),
                                   _p(o["pixel"], u32p), C.c_uint32(max_hits), _p(o["hit_prim"], u32p), _p(o["hit_t"]), _p(stats, u64p)))
        o["stats"] = stats
        return o

    def intersect(self, o, d, maxt, any_hit=False, q4=False, flat=False):
        o = _f32(o, (-1, 3)); d = _f32(d, (-1, 3)); maxt = _f32(maxt, (-1,)); n = len(o)
        prim = np.zeros(n, np.uint32); tuv = np.zeros((n, 3), np.float32); stats = np.zeros(2, np.uint64)
        _chk(lib.amvpt_intersect(self._h, _p(o), _p(d), _p(maxt), C.c_uint32(n), C.c_int(int(any_hit) | (2 if q4 else 0) | (4 if flat else 0)), _p(prim, u32p), _p(tuv), _p(stats, u64p)))
        return prim, tuv, stats

    STATIC_MAX_RAYS = 65536      # AMVPT_STATIC_MAX_RAYS (include/amvpt.h)

    def intersect_static(self, o, d, maxt, any_hit=False, q4=False):
        """amvpt_intersect_static: the same tap with its scratch in a static device array (no allocation; at most STATIC_MAX_RAYS rays)."""
        o = _f32(o, (-1, 3)); d = _f32(d, (-1, 3)); maxt = _f32(maxt, (-1,)); n = len(o)
        prim = np.zeros(n, np.uint32); tuv = np.zeros((n, 3), np.float32); stats = np.zeros(2, np.uint64)
        _chk(lib.amvpt_intersect_static(self._h, _p(o), _p(d), _p(maxt), C.c_uint32(n), C.c_int(int(any_hit) | (2 if q4 else 0)), _p(prim, u32p), _p(tuv), _p(stats, u64p)))
        return prim, tuv, stats

    def bsdf_eval_pdf_sample(self, bsdf, wi, wo, s1, s2):
        wi = _f32(wi, (-1, 3)); wo = _f32(wo, (-1, 3)); s1 = _f32(s1, (-1,)); s2 = _f32(s2, (-1, 2)); n = len(wi)
        r = dict(f=np.zeros((n, 3), np.float32), pdf=np.zeros(n, np.float32), bs_wo=np.zeros((n, 3), np.float32), bs_pdf=np.zeros(n, np.float32),
                 bs_eta=np.zeros(n, np.float32), bs_type=np.zeros(n, np.uint32), weight=np.zeros((n, 3), np.float32))
        _chk(lib.amvpt_bsdf_eval_pdf_sample(self._h, C.c_int(bsdf), _p(wi), _p(wo), _p(s1), _p(s2), C.c_uint32(n), _p(r["f"]), _p(r["pdf"]),
                                            _p(r["bs_wo"]), _p(r["bs_pdf"]), _p(r["bs_eta"]), _p(r["bs_type"], u32p), _p(r["weight"])))
        return r


def _defines(kw):
    items = [f"{k}={v}".encode() for k, v in kw.items()]
    arr = (C.c_char_p * max(len(items), 1))(*items)
    return arr, len(items)


def load_file(path, **defines):
    """mi.load_file(path, **kwargs) (reference: src/core/python/xml.cpp): kwargs substitute $name parameters."""
    h = C.c_void_p(); arr, n = _defines(defines)
    _chk(lib.amvpt_scene_from_xml(str(path).encode(), arr, n, C.byref(h)))
    return Scene(h)


def load_string(xml, base_dir=".", **defines):
    h = C.c_void_p(); arr, n = _defines(defines)
    _chk(lib.amvpt_scene_from_xml_string(xml.encode(), str(base_dir).encode(), arr, n, C.byref(h)))
    return Scene(h)


def f32_to_f16(x):
    """Float32 -> float16 bits as write_exr stores them (Struct conversion to Float16 for hdrfilm's component_format = float16)."""
    x = np.ascontiguousarray(x, np.float32); out = np.empty(x.shape, np.uint16)
    _chk(lib.amvpt_f32_to_f16(_p(x), out.ctypes.data_as(C.c_void_p), C.c_uint64(x.size))); return out


def membench(bytes, iters=20, device=0, kind=0):
    """GB/s of 16-byte loads over a buffer of `bytes` (L2 when it fits, else HBM); bench.py's L2 roofline denominator."""
    out = C.c_double(); _chk(lib.amvpt_membench(C.c_int(device), C.c_int(kind), C.c_uint64(bytes), C.c_uint32(iters), C.byref(out))); return out.value


def image_read(path):
    """Bitmap(path) -> linear RGB float32 (H, W, 3) through the scene loader's readers (OpenEXR scan-line none / PIZ, PFM)."""
    w, h = C.c_uint32(), C.c_uint32(); _chk(lib.amvpt_image_read(str(path).encode(), C.byref(w), C.byref(h), None))
    out = np.zeros((h.value, w.value, 3), np.float32); _chk(lib.amvpt_image_read(str(path).encode(), C.byref(w), C.byref(h), _p(out))); return out


def spectrum_to_srgb(wavelengths, values, within_emitter=False, unbounded_name=False):
    """<spectrum value="wavelength:value ..."> (or a single uniform value) -> linear sRGB as the XML loader converts it."""
    wl = np.ascontiguousarray(wavelengths, np.float64); v = np.ascontiguousarray(values, np.float64); out = np.zeros(3, np.float64)
    dp = C.POINTER(C.c_double)
    _chk(lib.amvpt_spectrum_to_srgb(wl.ctypes.data_as(dp), v.ctypes.data_as(dp), C.c_uint32(len(v)), C.c_int(int(within_emitter)), C.c_int(int(unbounded_name)), out.ctypes.data_as(dp)))
    return out


class Comm:
    """One NCCL communicator behind the C ABI (amvpt_comm).  `unique_id()` on one rank, hand the 128 bytes to the others, then Comm(id, rank, world, device)."""

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)(); _chk(lib.amvpt_comm_unique_id(buf)); return bytes(buf)

    def __init__(self, unique_id, rank, world, device):
        self._h = C.c_void_p(); buf = (C.c_uint8 * 128)(*unique_id)
        _chk(lib.amvpt_comm_init(buf, C.c_uint32(rank), C.c_uint32(world), C.c_int(device), C.byref(self._h)))
        self.rank, self.world = rank, world

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:
            lib.amvpt_comm_destroy(h)


def device_count():
    return lib.amvpt_device_count()


def math1(op, x, device=0):
    x = _f32(x, (-1,)); out = np.zeros_like(x)
    _chk(lib.amvpt_math1(C.c_int(device), C.c_int(op), _p(x), C.c_uint32(len(x)), _p(out))); return out


def sampler_stream(seed_value, lane, n, device=0):
    out = np.zeros(n, np.float32)
    _chk(lib.amvpt_sampler_stream(C.c_int(device), C.c_uint32(seed_value), C.c_uint32(lane), C.c_uint32(n), _p(out))); return out
