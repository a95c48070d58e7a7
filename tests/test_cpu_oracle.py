"""CPU-only tests (no GPU): the oracle against the reference's own golden vectors and against the reference's headers
compiled here (oracle/_ref), the host-side logic of the product (sensor matrices, BVH structure, BSDF tables, errors),
and the C-ABI surface (every symbol include/amvpt.h declares is exported)."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")))


def bits(a):
    a = np.ascontiguousarray(a, np.float32); u = a.view(np.uint32).copy(); u[np.isnan(a)] = 0x7fc00000; return u


# ------------------------------------------------------------------ oracle vs the reference's published known answers
def test_pcg32_and_tea_golden(ob):
    # ext/drjit/tests/test_rand.py:6-10 (PCG32 default state), src/core/tests/test_random.py:8-16 (TEA)
    o = ob.orc()
    out = np.zeros(1, np.float32)
    o.orc_pcg32_f32(C.c_uint64(0x853c49e6748fea9b), C.c_uint64(0xda3e39cb94b95bdb), C.c_int(1), out.ctypes.data_as(ob.f32p))
    assert float(out[0]) == 0.10837864875793457
    for (v0, v1), want in zip([(1, 1), (1, 2), (1, 3), (1, 4), (1, 5), (2, 1), (3, 1), (4, 1)], GOLDEN["tea_float32"]):
        r = np.zeros(2, np.uint32); o.orc_tea32(C.c_uint32(v0), C.c_uint32(v1), C.c_int(4), r.ctypes.data_as(ob.u32p))
        f = np.array([(int(r[1]) >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - np.float32(1)      # sample_tea_float32 uses .second, random.h:136-139
        assert float(f) == want


def test_independent_sampler_is_pcg32_initstate0(ob):
    # src/samplers/tests/test_independent.py:21-33: scalar sampler.seed(0) == PCG32(initstate=0); checks the PCG32 restatement
    o = ob.orc(); a = np.zeros(30, np.float32)
    o.orc_pcg32_f32(C.c_uint64(0), C.c_uint64(0xda3e39cb94b95bdb), C.c_int(30), a.ctypes.data_as(ob.f32p))
    assert np.allclose(a[:len(GOLDEN["pcg32_initstate0_f32"])], GOLDEN["pcg32_initstate0_f32"], rtol=0, atol=0)


def test_microfacet_golden(ob):
    # src/render/tests/test_microfacet.py:15-74 (Beckmann eval / pdf arrays from Mitsuba 0.6)
    o = ob.orc(); steps = 20
    theta = np.linspace(0, np.pi, steps); phi = np.full(steps, np.pi / 2)
    v = np.stack([np.cos(phi) * np.sin(theta), np.sin(phi) * np.sin(theta), np.cos(theta)], 1).astype(np.float32)
    wi = np.tile(np.array([0, 0, 1], np.float32), (steps, 1))
    for (au, av), key in [((0.1, 0.3), "beckmann_aniso"), ((0.1, 0.1), "beckmann_iso")]:
        ev = np.zeros(steps, np.float32); o.orc_mf_eval(C.c_int(0), C.c_float(au), C.c_float(av), v.ctypes.data_as(ob.f32p), C.c_int(steps), ev.ctypes.data_as(ob.f32p))
        assert np.allclose(ev, GOLDEN[key + "_eval"], rtol=1e-5, atol=1e-8)


def test_dielectric_golden(ob, scenes):
    # src/bsdfs/tests/test_dielectric.py:31-58: int_ior 1.5 / ext_ior 1, R = 0.04; radiance-mode transmission weight 0.6 / 1.5^2
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=8)
    desc["bsdfs"] = [dict(type="dielectric", int_ior=1.5, ext_ior=1.0, specular_reflectance=0.3, specular_transmittance=0.6)]
    for m in desc["meshes"]:
        m["bsdf"] = 0
    o = ob.OracleScene(desc)
    wi = np.array([[0, 0, 1], [0, 0, 1]], np.float32); wo = np.zeros((2, 3), np.float32)
    r = o.bsdf_eval_pdf_sample(0, wi, wo, np.array([0.0, 0.05], np.float32), np.zeros((2, 2), np.float32))
    assert np.allclose(r["weight"][0], 0.3) and np.isclose(r["bs_pdf"][0], 0.04) and r["bs_eta"][0] == 1.0 and np.allclose(r["bs_wo"][0], [0, 0, 1])
    assert np.allclose(r["weight"][1], 0.6 / 1.5 ** 2) and np.isclose(r["bs_pdf"][1], 0.96) and np.isclose(r["bs_eta"][1], 1.5) and np.allclose(r["bs_wo"][1], [0, 0, -1])
    assert r["bs_type"][0] == 16 and r["bs_type"][1] == 32


def test_perspective_center_ray(ob, scenes):
    # src/sensors/tests/test_perspective.py:79-84: the (0.5, 0.5) sample looks along the camera direction to 1e-7
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=64, cam_dist=0.0)
    o = ob.OracleScene(desc)
    s2c, tw = o.sensor(0)
    orc = ob.orc(); pos = np.array([[0.5, 0.5]], np.float32); ro = np.zeros((1, 3), np.float32); rd = np.zeros((1, 3), np.float32); mt = np.zeros(1, np.float32)
    orc.orc_perspective_sample_ray(s2c.ctypes.data_as(ob.f32p), tw.ctypes.data_as(ob.f32p), C.c_float(10), C.c_float(2800), pos.ctypes.data_as(ob.f32p), C.c_int(1),
                                   ro.ctypes.data_as(ob.f32p), rd.ctypes.data_as(ob.f32p), mt.ctypes.data_as(ob.f32p))
    assert np.allclose(rd[0], [0, 0, 1], atol=1e-7) and np.allclose(ro[0], [278, 273, -790], atol=1e-3)


# ------------------------------------------------------------------ oracle vs vectors produced by the reference's own headers (oracle/_ref)
def U(key, *path):
    d = GOLDEN[key]
    for p in path:
        d = d[p]
    return np.array(d, np.uint32)


def F(key, *path):
    return U(key, *path).view(np.float32)


def same(a, want, what):
    a = bits(a).reshape(-1); w = np.array(want, np.uint32).reshape(-1)
    wf = w.view(np.float32); w = w.copy(); w[np.isnan(wf)] = 0x7fc00000
    bad = np.flatnonzero(a != w)
    assert bad.size == 0, f"{what}: {bad.size} of {a.size} values differ from the reference headers, first {bad[:5]}"


def test_oracle_leaf_math_matches_reference_headers(ob):
    for op, v in GOLDEN["math1"].items():
        x = np.array(v["x"], np.uint32).view(np.float32)
        same(ob.math1(int(op), x), v["y"], f"math op {op}")
    for s in GOLDEN["sampler_streams"]:
        same(ob.sampler_stream(s["seed"], s["lane"], 16), s["bits"], f"sampler stream {s['seed']},{s['lane']}")


def test_oracle_vector_warp_fresnel_match_reference_headers(ob):
    o = ob.orc(); P = lambda x, t=ob.f32p: x.ctypes.data_as(t)
    a, b = F("dot", "a").reshape(-1, 3).copy(), F("dot", "b").reshape(-1, 3).copy(); n = len(a)
    o1 = np.zeros(n, np.float32); o3 = np.zeros((n, 3), np.float32); o3b = np.zeros((n, 3), np.float32); o2 = np.zeros((n, 2), np.float32); o4 = np.zeros((n, 4), np.float32)
    o.orc_dot(P(a), P(b), C.c_int(n), P(o1)); same(o1, GOLDEN["dot"]["y"], "dot")
    o.orc_cross(P(a), P(b), C.c_int(n), P(o3)); same(o3, GOLDEN["cross"]["y"], "cross")
    sc = F("normalize", "a").reshape(-1, 3).copy(); o.orc_normalize(P(sc), C.c_int(n), P(o3)); same(o3, GOLDEN["normalize"]["y"], "normalize")
    o.orc_coordinate_system(P(a), C.c_int(n), P(o3), P(o3b)); same(o3, GOLDEN["coordinate_system"]["s"], "coordinate_system s"); same(o3b, GOLDEN["coordinate_system"]["t"], "coordinate_system t")
    u2 = F("disk", "u").reshape(-1, 2).copy()
    o.orc_square_to_uniform_disk_concentric(P(u2), C.c_int(n), P(o2)); same(o2, GOLDEN["disk"]["y"], "concentric disk")
    o.orc_square_to_cosine_hemisphere(P(u2), C.c_int(n), P(o3), P(o1)); same(o3, GOLDEN["coshemi"]["y"], "cosine hemisphere"); same(o1, GOLDEN["coshemi"]["pdf"], "cosine hemisphere pdf")
    o.orc_square_to_uniform_triangle(P(u2), C.c_int(n), P(o2)); same(o2, GOLDEN["triangle"]["y"], "uniform triangle")
    ci = F("fresnel", "ci").copy()
    o.orc_fresnel(P(ci), C.c_float(GOLDEN["fresnel"]["eta"]), C.c_int(n), P(o4)); same(o4, GOLDEN["fresnel"]["y"], "fresnel")
    o.orc_fresnel_conductor(P(ci), C.c_float(0.2), C.c_float(3.9), C.c_int(n), P(o1)); same(o1, GOLDEN["fresnel_conductor"]["y"], "fresnel_conductor")
    o.orc_reflect_m(P(a), P(b), C.c_int(n), P(o3)); same(o3, GOLDEN["reflect"]["y"], "reflect")
    ct, et = F("refract", "ct").copy(), F("refract", "et").copy()
    o.orc_refract_m(P(a), P(b), P(ct), P(et), C.c_int(n), P(o3)); same(o3, GOLDEN["refract"]["y"], "refract")
    cs = [ob.ref_const_oracle(i) if hasattr(ob, "ref_const_oracle") else None for i in range(0)]
    assert GOLDEN["constants"][0] == pytest.approx(1500 * 2.0 ** -24) and GOLDEN["constants"][1] == pytest.approx(15000 * 2.0 ** -24)


def test_oracle_microfacet_matches_reference_headers(ob):
    o = ob.orc(); P = lambda x, t=ob.f32p: x.ctypes.data_as(t)
    wi, m, u2 = (F("mf_in", k).copy() for k in ("wi", "m", "u")); wi = wi.reshape(-1, 3); m = m.reshape(-1, 3); u2 = u2.reshape(-1, 2); n = len(wi)
    o1 = np.zeros(n, np.float32); o3 = np.zeros((n, 3), np.float32)
    for t in (0, 1):
        g = GOLDEN[f"mf_{t}"]; au, av = C.c_float(g["au"]), C.c_float(g["av"])
        o.orc_mf_eval(C.c_int(t), au, av, P(m), C.c_int(n), P(o1)); same(o1, g["eval"], f"microfacet {t} eval")
        o.orc_mf_smith_g1(C.c_int(t), au, av, P(wi), P(m), C.c_int(n), P(o1)); same(o1, g["g1"], f"microfacet {t} smith_g1")
        o.orc_mf_pdf(C.c_int(t), au, av, P(wi), P(m), C.c_int(n), P(o1)); same(o1, g["pdf"], f"microfacet {t} pdf")
        o.orc_mf_sample(C.c_int(t), au, av, P(wi), P(u2), C.c_int(n), P(o3), P(o1)); same(o3, g["sample"], f"microfacet {t} sample"); same(o1, g["sample_pdf"], f"microfacet {t} sample pdf")
    nodes = np.zeros(32, np.float32); w = np.zeros(32, np.float32); o.orc_gauss_legendre(C.c_int(32), P(nodes), P(w))
    same(nodes, GOLDEN["gauss_legendre32"]["nodes"], "gauss-legendre nodes"); same(w, GOLDEN["gauss_legendre32"]["weights"], "gauss-legendre weights")
    ext = np.zeros(64, np.float32); ir = np.zeros(1, np.float32)
    o.orc_rough_tables(C.c_int(0), C.c_float(0.1), C.c_float(np.float32(1.5046 / 1.000277)), C.c_int(64), P(ext), P(ir))
    # the reference evaluates the quadrature with 8/16-wide packets and a horizontal sum: same samples, different summation order
    assert np.allclose(ext, F("rough_tables_beckmann_0.1", "ext"), rtol=2e-5, atol=1e-6) and np.allclose(ir, F("rough_tables_beckmann_0.1", "int_refl"), rtol=2e-5)


def test_oracle_camera_and_triangle_match_reference_headers(ob):
    o = ob.orc(); P = lambda x, t=ob.f32p: x.ctypes.data_as(t)
    c2s = np.zeros(16, np.float32); s2c = np.zeros(16, np.float32)
    o.orc_sample_to_camera(C.c_int(256), C.c_int(256), C.c_float(39.3077), C.c_float(10), C.c_float(2800), C.c_float(0), P(c2s), P(s2c))
    same(c2s, GOLDEN["s2c_cbox"]["c2s"], "camera_to_sample (cbox)"); same(s2c, GOLDEN["s2c_cbox"]["s2c"], "sample_to_camera (cbox)")
    o.orc_sample_to_camera(C.c_int(512), C.c_int(384), C.c_float(55.5), C.c_float(0.1), C.c_float(100), C.c_float(0.125), P(c2s), P(s2c))
    same(c2s, GOLDEN["s2c_shift"]["c2s"], "camera_to_sample (lens shift)"); same(s2c, GOLDEN["s2c_shift"]["s2c"], "sample_to_camera (lens shift)")
    m = F("mat_inverse", "m").copy(); inv = np.zeros(16, np.float32); it = np.zeros(16, np.float32)
    o.orc_mat_inverse(P(m), P(inv)); o.orc_mat_inverse_transpose(P(m), P(it)); same(inv, GOLDEN["mat_inverse"]["inv"], "matrix inverse"); same(it, GOLDEN["mat_inverse"]["inv_t"], "matrix inverse transpose")
    g = GOLDEN["sample_ray"]; pos = F("sample_ray", "pos").copy(); n = len(pos) // 2
    ro = np.zeros((n, 3), np.float32); rd = np.zeros((n, 3), np.float32); mt = np.zeros(n, np.float32)
    o.orc_perspective_sample_ray(P(F("sample_ray", "s2c").copy()), P(F("sample_ray", "to_world").copy()), C.c_float(g["near"]), C.c_float(g["far"]), P(pos), C.c_int(n), P(ro), P(rd), P(mt))
    same(ro, g["o"], "sample_ray origin"); same(rd, g["d"], "sample_ray direction"); same(mt, g["maxt"], "sample_ray maxt")
    g = GOLDEN["moeller_trumbore"]; arrs = [F("moeller_trumbore", k).copy() for k in ("o", "d", "p0", "p1", "p2")]; n = len(arrs[0]) // 3
    mx = np.full(n, 1e30, np.float32); tuv = np.zeros((n, 3), np.float32); hit = np.zeros(n, np.int32)
    o.orc_moeller_trumbore(P(arrs[0]), P(arrs[1]), P(mx), P(arrs[2]), P(arrs[3]), P(arrs[4]), C.c_int(n), P(tuv), hit.ctypes.data_as(ob.i32p))
    assert hit.tolist() == g["hit"] and sum(g["hit"]) > 5
    same(tuv, g["tuv"], "moeller_trumbore t,u,v")


def test_golden_vectors_still_match_reference_headers_when_available(ob):
    """Where oracle/_ref is present (this container, and the GPU box through the snapshot) re-derive a few vectors live."""
    ref = ob.ref()
    if ref is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    for s in GOLDEN["sampler_streams"]:
        out = np.zeros(16, np.float32); ref.ref_sampler_stream(C.c_uint32(s["seed"]), C.c_uint32(s["lane"]), C.c_int(16), out.ctypes.data_as(ob.f32p))
        same(out, s["bits"], "live reference sampler stream")
    rng = np.random.default_rng(5); x = rng.uniform(-3, 3, 5000).astype(np.float32)
    for op in (0, 1, 2, 4, 6):
        out = np.zeros_like(x); ref.ref_math1(C.c_int(op), x.ctypes.data_as(ob.f32p), C.c_int(len(x)), out.ctypes.data_as(ob.f32p))
        same(ob.math1(op, x), out.view(np.uint32), f"live reference math op {op}")


def test_reconstruction_filters_known_answers_and_film(ob, scenes):
    """Known answers of the reference's src/rfilters/tests/test_rfilter.py:14-18, 42-46, and the footprint rule of the coalesced
    ImageBlock::put (src/render/imageblock.cpp:433-505): the weight channel of the film equals the sum over samples of the
    separable filter weights that fall inside the film."""
    g = ob.filter_eval("gaussian", 0.5, [0.2, 2.1, 0.0, 2.0, -0.2]); t = ob.filter_eval("tent", 1.0, [0.1, 1.1, -0.5])
    assert abs(g[0] - 0.9227) < 8e-3 and g[1] == 0 and abs(g[2] - 1.0) < 2e-3 and g[3] < 1e-5 and g[0] == g[4]
    assert abs(t[0] - 0.903) < 5e-2 and t[1] == 0 and t[2] == 0.5
    # test_rfilter.py:48-52 (mitchell), :54-58 (catmullrom), :60-64 (lanczos)
    m = ob.filter_eval("mitchell", 1 / 3, [0.0, 2.1, 1.0, -1.5], 1 / 3); c = ob.filter_eval("catmullrom", 0, [0.0, 2.1, 1.0, 0.5])
    z = ob.filter_eval("lanczos", 3, [1.4, 3.1, 0.0, 1.0, -1.4])
    assert abs(m[0] - 0.8888) < 1e-3 and m[1] == 0 and abs(m[2] - 1 / 18) < 1e-6 and abs(m[3] - (-7 / 3 * 3.375 + 12 * 2.25 - 20 * 1.5 + 32 / 3) / 6) < 1e-6
    assert abs(c[0] - 0.9765) < 5e-2 and c[0] == 1.0 and c[1] == 0 and c[2] == 0 and c[3] == np.float32(0.5625)
    assert abs(z[0] - (-0.14668)) < 1e-2 and z[1] == 0 and z[2] == 1 and abs(z[3]) < 1e-6 and z[0] == z[4]
    x = np.linspace(0, 3, 7, dtype=np.float32)[1:-1]; want = np.sinc(x.astype(np.float64)) * np.sinc(x.astype(np.float64) / 3)
    assert np.allclose(ob.filter_eval("lanczos", 3, x), want, atol=2e-6)
    radii = {"gaussian": 2.0, "tent": 1.0, "mitchell": 2.0, "catmullrom": 2.0, "lanczos": 2.0}
    for rf in (("gaussian", 0.5), ("tent", 1.0), ("mitchell", 1 / 3, 1 / 3), ("catmullrom", 0.0), ("lanczos", 2)):
        desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=12, spp_pass_lim=4, sample_count=4); desc["rfilter"] = rf
        o = ob.OracleScene(desc); film, _, _ = o.render(seed=0, spp=4, brute=True, nthreads=3)
        film1, _, _ = o.render(seed=0, spp=4, brute=True, nthreads=1)
        assert np.allclose(film, film1, rtol=0, atol=1e-5 * film1.max())            # per-thread films summed = sequential film
        W, H = o.film_size(); pos = o.trace_lanes(0, 4, 0, W * H * 4, trace=False)["sample_pos"]
        radius = radii[rf[0]]; n = int(np.ceil(radius - 0.5)); want = np.zeros((H, W)); p2 = rf[2] if len(rf) > 2 else 0.0
        for sx, sy in pos.astype(np.float64):
            ix, iy = int(np.floor(sx)) - n, int(np.floor(sy)) - n
            wx = ob.filter_eval(rf[0], rf[1], np.float32(ix + 0.5 - sx) + np.arange(2 * n + 1, dtype=np.float32), p2)
            wy = ob.filter_eval(rf[0], rf[1], np.float32(iy + 0.5 - sy) + np.arange(2 * n + 1, dtype=np.float32), p2)
            for j in range(2 * n + 1):
                for i in range(2 * n + 1):
                    if 0 <= ix + i < W and 0 <= iy + j < H:
                        want[iy + j, ix + i] += float(wy[j]) * float(wx[i])
        assert np.allclose(film[..., 3], want, rtol=1e-4, atol=1e-5)
        box = dict(desc); box["rfilter"] = ("box", 0.5); fb, _, _ = ob.OracleScene(box).render(seed=0, spp=4, brute=True)
        a, b = ob.develop(film), ob.develop(fb)
        assert abs(a.mean() - b.mean()) < 0.05 * b.mean()                           # the filter redistributes, it does not add energy


# ------------------------------------------------------------------ display step: oracle/o_display.h
DISPLAY = json.load(open(os.path.join(ROOT, "tests", "golden", "display_vectors.json")))
LKG4x8 = dict(center=-0.5, focus=-0.05, pitch=354.548, tilt=-0.114, subp=0.00013, view=(1.0, 1.0), grid=(4, 8))   # src/mitsuba/program.h:74-75


def _display_leaves(lib, prefix, g):
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    x = np.array(g["linear_to_srgb"]["x"], np.uint32).view(np.float32); y = np.zeros_like(x)
    if prefix == "orc": lib.orc_linear_to_srgb(P(x), C.c_uint64(len(x)), P(y))
    else: lib.ref_linear_to_srgb(P(x), C.c_int(len(x)), P(y))
    assert np.array_equal(y.view(np.uint32), np.array(g["linear_to_srgb"]["y"], np.uint32)), "linear_to_srgb"
    a = np.array(g["clip_mod"]["x"], np.uint32).view(np.float32); b = np.array(g["clip_mod"]["y"], np.uint32).view(np.float32)
    c = np.zeros_like(a); m = np.zeros_like(a); getattr(lib, prefix + "_clip_mod")(P(a), P(b), C.c_int(len(a)), P(c), P(m))
    assert np.array_equal(c.view(np.uint32), np.array(g["clip_mod"]["clip"], np.uint32)) and np.array_equal(m.view(np.uint32), np.array(g["clip_mod"]["mod"], np.uint32))
    gi = g["interpolate2d"]; img = np.array(gi["img"], np.uint8); xy = np.array(gi["xy"], np.uint32).view(np.float32); n = len(xy) // 2
    for key, want in gi["out"].items():
        near, ch = map(int, key.split("_")); o = np.zeros(n, np.uint8)
        getattr(lib, prefix + "_interpolate2d_u8")(P(img), gi["w"], gi["h"], gi["nch"], P(xy), n, ch, near, P(o))
        assert np.array_equal(o, np.array(want, np.uint8)), f"interpolate2d {key}"


def test_display_leaves_match_reference_golden(ob):
    """dr::linear_to_srgb, clip, mod and interpolate2d<uint8_t> of the oracle, bit for bit against vectors made from the reference's headers."""
    _display_leaves(ob.orc(), "orc", DISPLAY)


def test_display_golden_still_matches_reference_headers_when_available(ob):
    ref = ob.ref()
    if ref is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    _display_leaves(ref, "ref", DISPLAY)


def test_srgb8_known_values_and_rounding(ob):
    """Bitmap::convert(RGB, UInt8, srgb): 0 -> 0, 1 -> 255, mid grey 0.2140 -> 127/128 boundary, clamping above 1; monotone."""
    x = np.array([0.0, 1.0, 2.5, 0.0031308, 0.001, 0.18, 0.5], np.float32)
    got = ob.to_srgb8(x)
    want = np.rint(np.clip(np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055), 0, 1) * 255).astype(np.uint8)
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1 and got[0] == 0 and got[1] == 255 and got[2] == 255
    ramp = ob.to_srgb8(np.linspace(0, 1.2, 5000, dtype=np.float32))
    assert (np.diff(ramp.astype(int)) >= 0).all() and set(np.unique(ramp)) == set(range(256))


def test_interlace_oracle_properties(ob):
    rng = np.random.default_rng(5)
    quilt = rng.integers(0, 256, (64, 48, 3), dtype=np.uint8)
    # "Display quilt" + nearest scaling at the quilt's own size: rows map to themselves; x0 = dx / (row_bytes - 1) with dx the BYTE offset
    # (program.cpp:213-215, 241), so only the first column is the identity -- a quirk of the reference that the oracle keeps
    q = ob.interlace(quilt, dict(LKG4x8, quilt=True, nearest=True, flip_rgb=True), 48, 64, 3)
    assert q.shape == (64, 48, 3) and np.array_equal(q[:, 0], quilt[:, 0])
    assert np.array_equal(q[0], ob.interlace(quilt[:1].repeat(2, 0), dict(LKG4x8, quilt=True, nearest=True, flip_rgb=True), 48, 2, 3)[0])
    # flip_rgb = false writes channel c to 2 - c (BGR surface, program.cpp:245); the fourth channel is opaque
    a = ob.interlace(quilt, dict(LKG4x8, flip_rgb=True), 40, 30, 4); b = ob.interlace(quilt, dict(LKG4x8, flip_rgb=False), 40, 30, 4)
    assert np.array_equal(a[..., :3], b[..., 2::-1]) and (a[..., 3] == 255).all()
    # a quilt whose views are flat colours: each interlaced sub-pixel is exactly one view's colour (nearest), and several views show up
    gx, gy = LKG4x8["grid"]; flat = np.zeros((8 * gy, 8 * gx, 3), np.uint8)
    for v in range(gx * gy):
        flat[(v // gx) * 8:(v // gx + 1) * 8, (v % gx) * 8:(v % gx + 1) * 8] = 8 * v + 3
    il = ob.interlace(flat, dict(LKG4x8, nearest=True, focus=0.0), 200, 120, 3)
    assert set(np.unique(il)) <= {8 * v + 3 for v in range(gx * gy)} and len(np.unique(il)) > gx * gy // 2
