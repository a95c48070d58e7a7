"""ctypes binding of the parity oracle (oracle/libamvpt_oracle.so) and of the reference-header shim
(oracle/_ref/libamvpt_refshim.so).  TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess
import numpy as np

RFILTERS = {"box": 0, "gaussian": 1, "tent": 2, "mitchell": 3, "catmullrom": 4, "lanczos": 5}

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libamvpt_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libamvpt_refshim.so")
f32p, u32p, u64p, i32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int)


def build_oracle():
    src = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".h"))]
    if not os.path.exists(ORACLE_SO) or any(os.path.getmtime(s) > os.path.getmtime(ORACLE_SO) for s in src):
        subprocess.run(["make", "-C", ORACLE_DIR, "oracle"], check=True, capture_output=True)
    return ORACLE_SO


def build_ref():
    """Only where the reference tree is mounted (this container); elsewhere the prebuilt .so is used if present."""
    if os.path.isdir("/root/reference/include"):
        subprocess.run(["make", "-C", ORACLE_DIR, "_ref"], check=True, capture_output=True)
    return REF_SO if os.path.exists(REF_SO) else None


_orc = None


def orc():
    global _orc
    if _orc is None:
        _orc = C.CDLL(build_oracle())
        _orc.orc_scene_new.restype = C.c_void_p
        _orc.orc_render.restype = C.c_double
        _orc.orc_render_rows.restype = C.c_double
        _orc.orc_num_prims.restype = C.c_uint32
    return _orc


def ref():
    p = REF_SO if os.path.exists(REF_SO) else build_ref()
    if not p:
        return None
    lib = C.CDLL(p)
    lib.ref_const.restype = C.c_float
    lib.ref_parse_fov_axis.restype = C.c_double
    return lib


class OrcBsdfDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("twosided", C.c_int32), ("distr", C.c_int32), ("nonlinear", C.c_int32),
                ("reflectance", C.c_float * 3), ("eta", C.c_float * 3), ("k", C.c_float * 3), ("spec_refl", C.c_float * 3),
                ("spec_trans", C.c_float * 3), ("int_ior", C.c_float), ("ext_ior", C.c_float), ("alpha_u", C.c_float), ("alpha_v", C.c_float)]


class OrcSensorDesc(C.Structure):
    _fields_ = [("thinlens", C.c_int32), ("fov_axis", C.c_int32), ("grid_x", C.c_uint32), ("grid_y", C.c_uint32), ("res_x", C.c_uint32),
                ("res_y", C.c_uint32), ("reverse_x", C.c_int32), ("reverse_y", C.c_int32), ("use_cone", C.c_int32), ("cam_center", C.c_int32),
                ("to_world", C.c_double * 16), ("to_world_inv_t", C.c_double * 16), ("fov", C.c_double),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("focus_distance", C.c_float), ("aperture_radius", C.c_float),
                ("cone_deg", C.c_float), ("cam_dist", C.c_float), ("cam_dir", C.c_float * 3), ("cam_off", C.c_float * 3)]


BSDF_TYPES = {"diffuse": 0, "conductor": 1, "dielectric": 2, "roughconductor": 3, "roughdielectric": 4, "roughplastic": 5}
FOV_AXES = {"x": 0, "y": 1, "diagonal": 2, "smaller": 3, "larger": 4}


def _p(a, t=f32p):
    return None if a is None else a.ctypes.data_as(t)


def _v3(v):
    return (C.c_float * 3)(*np.broadcast_to(np.asarray(v, np.float32), (3,)))


class OracleScene:
    """Same construction calls as the product's Scene, feeding oracle/amvpt_oracle.cpp."""

    def __init__(self, desc=None):
        self.o = orc(); self.h = C.c_void_p(self.o.orc_scene_new())
        if desc is not None:
            for b in desc["bsdfs"]:
                self.add_bsdf(**b)
            for m in desc["meshes"]:
                self.add_mesh(m["positions"], m["faces"], m["bsdf"], m["normals"], m["radiance"], m["flip_normals"])
            self.set_sensor_grid(**desc["sensor"]); self.set_integrator(**desc["integrator"])
            if desc.get("rfilter"):
                self.set_rfilter(*desc["rfilter"])
            if desc.get("reuse"):
                self.set_reuse(**desc["reuse"])
            self.prepare()

    @classmethod
    def from_product(cls, scene):
        """The same scene the product holds (e.g. loaded from XML by amvpt_scene_from_xml): meshes, BSDF parameters, emitters in scene order,
        the raw environment-map pixels, the sub-sensor table and the integrator parameters go to the oracle, which rebuilds everything
        derived (area CDFs, roughplastic tables, environment hierarchy, bounding sphere, its own BVH) by itself."""
        o = cls(); c = scene.counts()
        for i in range(c["n_bsdfs"]):
            b = scene.bsdf(i); b["alpha_u"], b["alpha_v"] = float(b["alpha_u"]), float(b["alpha_v"]); o.add_bsdf(**b)
        for i in range(c["n_meshes"]):
            m = scene.mesh(i); o.add_mesh(m["positions"], m["faces"], m["bsdf"], m["normals"], m["radiance"], m["flip_normals"])
        ems = scene.emitters(); tmap = {"area": 0, "point": 1, "constant": 2, "envmap": 3}
        for e in ems:
            if e["type"] == "point":
                o.o.orc_add_emitter_point(o.h, _p(np.ascontiguousarray(e["position"], np.float32)), _p(np.ascontiguousarray(e["radiance"], np.float32)))
            elif e["type"] == "constant":
                o.o.orc_add_emitter_constant(o.h, _p(np.ascontiguousarray(e["radiance"], np.float32)))
            elif e["type"] == "envmap":
                env = scene.envmap(); rgb = np.ascontiguousarray(env["rgb"], np.float32); tw = np.ascontiguousarray(env["to_world"], np.float32)
                o.o.orc_add_emitter_envmap(o.h, _p(rgb), C.c_uint32(rgb.shape[1]), C.c_uint32(rgb.shape[0]), C.c_float(env["scale"]), _p(tw))
        ty = np.array([tmap[e["type"]] for e in ems], np.int32); me = np.array([e["mesh"] for e in ems], np.int32)
        o.o.orc_order_emitters(o.h, ty.ctypes.data_as(C.POINTER(C.c_int32)), me.ctypes.data_as(C.POINTER(C.c_int32)), C.c_uint32(len(ems)))
        info = scene.info(); sp0 = scene.sensor_params(0); n = info["n_views"]
        s2c = np.zeros((n, 16), np.float32); tw = np.zeros((n, 16), np.float32); sc = np.zeros((n, 4), np.float32)
        for i in range(n):
            a, b = scene.sensor(i); s2c[i] = a.reshape(16); tw[i] = b.reshape(16); q = scene.sensor_params(i)
            sc[i] = [q["near_clip"], q["far_clip"], q["aperture_radius"], q["focus_distance"]]
        o.o.orc_set_sensor_raw(o.h, C.c_uint32(sp0["grid_x"]), C.c_uint32(sp0["grid_y"]), C.c_uint32(info["film_w"] // sp0["grid_x"]), C.c_uint32(info["film_h"] // sp0["grid_y"]),
                               C.c_int(int(sp0["reverse_x"])), C.c_int(int(sp0["reverse_y"])), C.c_int(int(sp0["thinlens"])), _p(s2c), _p(tw), _p(sc))
        ig = scene.integrator()
        o.set_integrator(max_depth=ig["max_depth"], rr_depth=ig["rr_depth"], hide_emitters=ig["hide_emitters"], spp_pass_lim=ig["spp_pass_lim"], sampler_seed=ig["sampler_seed"])
        o.o.orc_set_shutter_draw(o.h, C.c_int(int(scene.shutter_draw())))
        o.prepare()
        return o

    def __del__(self):
        if getattr(self, "h", None):
            self.o.orc_scene_free(self.h); self.h = None

    def add_bsdf(self, type, twosided=False, distribution="beckmann", nonlinear=False, reflectance=(.5, .5, .5), eta=(0., 0., 0.), k=(1., 1., 1.),
                 specular_reflectance=(1., 1., 1.), specular_transmittance=(1., 1., 1.), int_ior=1.5046, ext_ior=1.000277, alpha=0.1,
                 alpha_u=None, alpha_v=None):
        d = OrcBsdfDesc(BSDF_TYPES[type], int(twosided), {"beckmann": 0, "ggx": 1}[distribution], int(nonlinear), _v3(reflectance), _v3(eta), _v3(k),
                        _v3(specular_reflectance), _v3(specular_transmittance), int_ior, ext_ior,
                        alpha if alpha_u is None else alpha_u, alpha if alpha_v is None else alpha_v)
        return self.o.orc_add_bsdf(self.h, C.byref(d))

    def add_mesh(self, positions, faces, bsdf, normals=None, radiance=None, flip_normals=False):
        pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3); fc = np.ascontiguousarray(faces, np.uint32).reshape(-1, 3)
        nrm = None if normals is None else np.ascontiguousarray(normals, np.float32)
        rad = None if radiance is None else np.ascontiguousarray(np.broadcast_to(np.asarray(radiance, np.float32), (3,)))
        return self.o.orc_add_mesh(self.h, _p(pos), C.c_uint32(len(pos)), _p(fc, u32p), C.c_uint32(len(fc)), _p(nrm), C.c_int(bsdf), _p(rad), C.c_int(int(flip_normals)))

    def set_sensor_grid(self, grid_x, grid_y, res_x, res_y, to_world, fov, fov_axis="x", near_clip=1e-2, far_clip=1e4, focus_distance=1.0,
                        thinlens=False, aperture_radius=0.0, reverse_x=False, reverse_y=True, cone_deg=None, cam_dir=(1., 0., 0.), cam_dist=None,
                        cam_center=True, cam_off=(0., 0., 0.)):
        tw = np.asarray(to_world, np.float64).reshape(4, 4); it = np.linalg.inv(tw).T
        cd = np.asarray(cam_dir, np.float32)
        dist = float(np.float32(np.sqrt(np.float32(cd[0] * cd[0] + cd[1] * cd[1] + cd[2] * cd[2])))) if cam_dist is None else cam_dist
        d = OrcSensorDesc(int(thinlens), FOV_AXES[fov_axis], grid_x, grid_y, res_x // grid_x, res_y // grid_y, int(reverse_x), int(reverse_y),
                          int(cone_deg is not None), int(cam_center), (C.c_double * 16)(*tw.reshape(16)), (C.c_double * 16)(*it.reshape(16)), fov,
                          near_clip, far_clip, focus_distance, aperture_radius, 0.0 if cone_deg is None else cone_deg, dist, _v3(cd), _v3(cam_off))
        self.o.orc_set_sensor_grid(self.h, C.byref(d))

    def set_sensor_batch(self, res_x, res_y, sample_to_camera, to_world, near_clip, far_clip):
        s2c = np.ascontiguousarray(sample_to_camera, np.float32).reshape(-1, 16); tw = np.ascontiguousarray(to_world, np.float32).reshape(-1, 16)
        self.o.orc_set_sensor_explicit(self.h, C.c_uint32(len(s2c)), C.c_uint32(res_x), C.c_uint32(res_y), _p(s2c), _p(tw), C.c_float(near_clip), C.c_float(far_clip))

    def set_integrator(self, max_depth=-1, rr_depth=5, hide_emitters=False, spp_pass_lim=16, sampler_seed=0, sample_count=0):
        self.spp_pass_lim = spp_pass_lim
        self.o.orc_set_integrator(self.h, C.c_int64(max_depth), C.c_uint32(rr_depth), C.c_int(int(hide_emitters)), C.c_uint32(sampler_seed))

    def set_reuse(self, sa_reuse=True, sa_mis=False, fast_mis=False, reuse_count=0, adaptive=0, debug=False):
        """MVPathIntegrator sa_reuse / sa_mis / fast_mis / reuse_count; returns the samples per primary sample (0 = off)."""
        self.o.orc_set_reuse.restype = C.c_uint32
        self.n_reuse = self.o.orc_set_reuse(self.h, C.c_int(int(sa_reuse)), C.c_int(int(sa_mis)), C.c_int(int(fast_mis)), C.c_uint32(reuse_count))
        self.o.orc_set_reuse_adaptive(self.h, C.c_uint32(adaptive), C.c_int(int(debug)))
        return self.n_reuse

    def trace_lanes_multi(self, seed_value, spp_per_pass, lane_begin, n, brute=True):
        """Reuse-branch tap: (n, n_reuse, 9) = pos.xy, weight, result.rgb, valid, indirect, sensor index."""
        out = np.zeros((n, self.n_reuse, 9), np.float32)
        self.o.orc_trace_lanes_multi(self.h, C.c_uint32(seed_value), C.c_uint32(spp_per_pass), C.c_uint32(lane_begin), C.c_uint32(n), C.c_int(int(brute)), _p(out))
        return out

    def set_rfilter(self, type="box", param=0.5, param2=0.0):
        self.o.orc_set_rfilter(self.h, C.c_int(RFILTERS[type]), C.c_float(param), C.c_float(param2))

    def prepare(self):
        self.o.orc_prepare(self.h)

    def film_size(self):
        w, h = C.c_uint32(), C.c_uint32(); self.o.orc_film_size(self.h, C.byref(w), C.byref(h)); return w.value, h.value

    def sensor(self, i):
        a = np.zeros(16, np.float32); b = np.zeros(16, np.float32); self.o.orc_get_sensor(self.h, C.c_uint32(i), _p(a), _p(b))
        return a.reshape(4, 4), b.reshape(4, 4)

    def bsdf_tables(self, bsdf):
        t = np.zeros(64, np.float32); s = np.zeros(3, np.float32); self.o.orc_bsdf_tables(self.h, C.c_int(bsdf), _p(t), _p(s)); return t, s

    def trace_lanes(self, seed_value, spp_per_pass, lane_begin, n, trace=True, max_hits=0, brute=False):
        o = dict(sample_pos=np.zeros((n, 2), np.float32), ray_o=np.zeros((n, 3), np.float32), ray_d=np.zeros((n, 3), np.float32),
                 ray_maxt=np.zeros(n, np.float32), view=np.zeros(n, np.uint32), radiance=np.zeros((n, 4), np.float32), pixel=np.zeros(n, np.uint32),
                 hit_prim=np.zeros((n, max(max_hits, 1)), np.uint32), hit_t=np.zeros((n, max(max_hits, 1)), np.float32))
        stats = np.zeros(2, np.uint64)
        self.o.orc_trace_lanes(self.h, C.c_uint32(seed_value), C.c_uint32(spp_per_pass), C.c_uint32(lane_begin), C.c_uint32(n), C.c_int(int(brute)),
                               C.c_int(int(trace)), _p(o["sample_pos"]), _p(o["ray_o"]), _p(o["ray_d"]), _p(o["ray_maxt"]), _p(o["view"], u32p),
                               _p(o["radiance"]), _p(o["pixel"], u32p), C.c_uint32(max_hits), _p(o["hit_prim"], u32p), _p(o["hit_t"]), _p(stats, u64p))
        o["stats"] = stats
        return o

    def intersect(self, o, d, maxt, brute=False):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3); d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        maxt = np.ascontiguousarray(maxt, np.float32); n = len(o)
        prim = np.zeros(n, np.uint32); tuv = np.zeros((n, 3), np.float32); stats = np.zeros(2, np.uint64)
        self.o.orc_intersect(self.h, _p(o), _p(d), _p(maxt), C.c_int(n), C.c_int(int(brute)), _p(prim, u32p), _p(tuv), _p(stats, u64p))
        return prim, tuv, stats

    def ray_test(self, o, d, maxt, brute=False):
        o = np.ascontiguousarray(o, np.float32).reshape(-1, 3); d = np.ascontiguousarray(d, np.float32).reshape(-1, 3)
        maxt = np.ascontiguousarray(maxt, np.float32); n = len(o); occ = np.zeros(n, np.int32)
        self.o.orc_ray_test(self.h, _p(o), _p(d), _p(maxt), C.c_int(n), C.c_int(int(brute)), _p(occ, i32p)); return occ

    def bsdf_eval_pdf_sample(self, bsdf, wi, wo, s1, s2):
        wi = np.ascontiguousarray(wi, np.float32).reshape(-1, 3); wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        s1 = np.ascontiguousarray(s1, np.float32); s2 = np.ascontiguousarray(s2, np.float32).reshape(-1, 2); n = len(wi)
        r = dict(f=np.zeros((n, 3), np.float32), pdf=np.zeros(n, np.float32), bs_wo=np.zeros((n, 3), np.float32), bs_pdf=np.zeros(n, np.float32),
                 bs_eta=np.zeros(n, np.float32), bs_type=np.zeros(n, np.uint32), weight=np.zeros((n, 3), np.float32))
        self.o.orc_bsdf_eval_pdf_sample(self.h, C.c_int(bsdf), _p(wi), _p(wo), _p(s1), _p(s2), C.c_int(n), _p(r["f"]), _p(r["pdf"]), _p(r["bs_wo"]),
                                        _p(r["bs_pdf"]), _p(r["bs_eta"]), _p(r["bs_type"], u32p), _p(r["weight"]))
        return r

    def render_alloc(self, alloc, seed=0, spp_per_pass=16, brute=False, nthreads=1):
        """Film of the samples an adaptive render claims: alloc[pass][view] samples per pixel (single thread by default: exact row ownership)."""
        alloc = np.ascontiguousarray(alloc, np.uint32); w, h = self.film_size(); film = np.zeros((h, w, 4), np.float32)
        self.o.orc_render_alloc(self.h, C.c_uint32(seed), C.c_uint32(spp_per_pass), C.c_uint32(alloc.shape[0]), _p(alloc, u32p), C.c_int(int(brute)), C.c_int(nthreads), _p(film))
        return film

    def render(self, seed=0, spp=16, brute=False, nthreads=None, rows=None, row_stride=1):
        """Returns (RGBW film [H, W, 4], seconds in the pass loop, (nodes, tris))."""
        w, h = self.film_size(); film = np.zeros((h, w, 4), np.float32); stats = np.zeros(2, np.uint64)
        r0, r1 = rows if rows is not None else (0, h)
        nthreads = nthreads or os.cpu_count() or 1
        sec = self.o.orc_render_rows(self.h, C.c_uint32(seed), C.c_uint32(spp), C.c_uint32(self.spp_pass_lim), C.c_int(int(brute)), C.c_int(nthreads),
                                     _p(film), C.c_uint32(r0), C.c_uint32(r1), C.c_uint32(row_stride), _p(stats, u64p))
        return film, sec, stats


def develop(rgbw):
    w = rgbw[..., 3:4]
    return rgbw[..., :3] / np.where(w == 0, 1, w)


# ---- display step (oracle/o_display.h)
u8p = C.POINTER(C.c_uint8)


def to_srgb8(rgb):
    x = np.ascontiguousarray(rgb, np.float32); out = np.zeros(x.shape, np.uint8); orc().orc_to_srgb8(_p(x), C.c_uint64(x.size), _p(out, u8p)); return out


def preset_arrays(preset, dst_ch):
    pf = np.array([preset["center"], preset["focus"], preset["pitch"], preset["tilt"], preset["subp"], *preset["view"]], np.float32)
    pi = np.array([*preset["grid"], int(preset.get("flip_rgb", dst_ch == 3)), int(preset.get("quilt", False)), int(preset.get("nearest", False))], np.int32)
    return pf, pi


def interlace(src, preset, dst_w, dst_h, dst_ch=3):
    src = np.ascontiguousarray(src, np.uint8); h, w, c = src.shape; pf, pi = preset_arrays(preset, dst_ch)
    out = np.zeros((dst_h, dst_w, dst_ch), np.uint8)
    orc().orc_interlace(_p(src, u8p), C.c_uint32(w), C.c_uint32(h), C.c_uint32(c), _p(pf), _p(pi, C.POINTER(C.c_int32)), _p(out, u8p), C.c_uint32(dst_w), C.c_uint32(dst_h), C.c_uint32(dst_ch))
    return out


def math1(op, x):
    x = np.ascontiguousarray(x, np.float32); out = np.zeros_like(x); orc().orc_math1(C.c_int(op), _p(x), C.c_int(len(x)), _p(out)); return out


def sampler_stream(seed_value, lane, n):
    out = np.zeros(n, np.float32); orc().orc_sampler_stream(C.c_uint32(seed_value), C.c_uint32(lane), C.c_int(n), _p(out)); return out


def filter_eval(type, param, x, param2=0.0):
    x = np.ascontiguousarray(x, np.float32); out = np.zeros_like(x)
    orc().orc_filter_eval(C.c_int(RFILTERS[type]), C.c_float(param), C.c_float(param2), _p(x), C.c_int(len(x)), _p(out)); return out
