"""Generates tests/golden/reference_vectors.json.  Run in the build container only (needs /root/reference):
  python tests/golden/make_golden.py
Two kinds of vectors: (1) known answers transcribed by regex from the reference's own test files; (2) outputs of the
reference's own header-only code compiled here (oracle/_ref, see oracle/Makefile) on seeded inputs, stored as uint32
bit patterns together with the inputs, so that the oracle can be pinned bit-for-bit where /root/reference is absent."""
import ctypes as C
import json
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_binding as ob  # noqa: E402

REF = "/root/reference"
f32p, u32p, i32p = ob.f32p, ob.u32p, ob.i32p


def floats_after(path, anchor, count):
    txt = open(os.path.join(REF, path)).read()
    txt = txt[txt.index(anchor):]
    nums = re.findall(r"-?\d+\.\d+(?:e[-+]?\d+)?", txt)
    return [float(x) for x in nums[:count]]


def u(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32).reshape(-1).tolist()


def main():
    ref = ob.ref()
    assert ref is not None, "oracle/_ref could not be built (needs /root/reference)"
    rng = np.random.default_rng(20261104)
    g = {}
    # ---- (1) known answers from the reference's tests
    txt = open(os.path.join(REF, "src/core/tests/test_random.py")).read()
    g["tea_float32"] = [float(x) for x in re.findall(r"sample_tea_float32\(\d, \d, 4\) == ([0-9.]+)", txt)]
    assert len(g["tea_float32"]) == 8
    g["beckmann_aniso_eval"] = floats_after("src/render/tests/test_microfacet.py", "assert dr.allclose(mdf.eval(v)", 20)
    g["beckmann_iso_eval"] = floats_after("src/render/tests/test_microfacet.py", "assert dr.allclose(mdf_i.eval(v)", 20)
    # ---- (2) reference headers on seeded inputs
    n = 200
    out = np.zeros(30, np.float32); ref.ref_pcg32_f32(C.c_uint64(0), C.c_uint64(0xda3e39cb94b95bdb), C.c_int(30), out.ctypes.data_as(f32p))
    g["pcg32_initstate0_f32"] = [float(x) for x in out]
    streams = []
    for seed, lane in [(0, 0), (0, 1), (16, 4194303), (0xdeadbeef, 0xffffffff), (48, 123456789)]:
        o = np.zeros(16, np.float32); ref.ref_sampler_stream(C.c_uint32(seed), C.c_uint32(lane), C.c_int(16), o.ctypes.data_as(f32p))
        streams.append(dict(seed=seed, lane=lane, bits=u(o)))
    g["sampler_streams"] = streams
    ranges = {0: (-20, 20), 1: (-20, 20), 2: (-80, 80), 3: None, 4: (-4, 4), 5: (-0.999999, 0.999999), 6: (-120, 120), 7: (0, 1e6), 8: (-1e3, 1e3), 9: (1e-6, 1e6), 10: (-1, 1)}
    g["math1"] = {}
    for op, r in ranges.items():
        x = (np.exp(rng.uniform(-60, 60, n)) if r is None else rng.uniform(r[0], r[1], n)).astype(np.float32)
        o = np.zeros(n, np.float32); ref.ref_math1(C.c_int(op), x.ctypes.data_as(f32p), C.c_int(n), o.ctypes.data_as(f32p))
        g["math1"][str(op)] = dict(x=u(x), y=u(o))

    def dirs(k):
        v = rng.normal(size=(k, 3)); v /= np.linalg.norm(v, axis=1, keepdims=True); return np.ascontiguousarray(v, np.float32)
    a3, b3 = dirs(n), dirs(n); u2 = np.ascontiguousarray(rng.uniform(size=(n, 2)), np.float32)
    o1 = np.zeros(n, np.float32); o3 = np.zeros((n, 3), np.float32); o3b = np.zeros((n, 3), np.float32); o2 = np.zeros((n, 2), np.float32)
    P = lambda x, t=f32p: x.ctypes.data_as(t)
    ref.ref_dot(P(a3), P(b3), C.c_int(n), P(o1)); g["dot"] = dict(a=u(a3), b=u(b3), y=u(o1))
    ref.ref_cross(P(a3), P(b3), C.c_int(n), P(o3)); g["cross"] = dict(y=u(o3))
    sc = (a3 * np.float32(3.7)).astype(np.float32); ref.ref_normalize(P(sc), C.c_int(n), P(o3)); g["normalize"] = dict(a=u(sc), y=u(o3))
    ref.ref_coordinate_system(P(a3), C.c_int(n), P(o3), P(o3b)); g["coordinate_system"] = dict(s=u(o3), t=u(o3b))
    ref.ref_square_to_uniform_disk_concentric(P(u2), C.c_int(n), P(o2)); g["disk"] = dict(u=u(u2), y=u(o2))
    ref.ref_square_to_cosine_hemisphere(P(u2), C.c_int(n), P(o3), P(o1)); g["coshemi"] = dict(y=u(o3), pdf=u(o1))
    ref.ref_square_to_uniform_triangle(P(u2), C.c_int(n), P(o2)); g["triangle"] = dict(y=u(o2))
    ci = np.ascontiguousarray(rng.uniform(-1, 1, n), np.float32); o4 = np.zeros((n, 4), np.float32)
    ref.ref_fresnel(P(ci), C.c_float(1.5046 / 1.000277), C.c_int(n), P(o4)); g["fresnel"] = dict(ci=u(ci), eta=float(np.float32(1.5046 / 1.000277)), y=u(o4))
    ref.ref_fresnel_conductor(P(ci), C.c_float(0.2), C.c_float(3.9), C.c_int(n), P(o1)); g["fresnel_conductor"] = dict(y=u(o1))
    ref.ref_reflect_m(P(a3), P(b3), C.c_int(n), P(o3)); g["reflect"] = dict(y=u(o3))
    ct = np.ascontiguousarray(rng.uniform(-1, 1, n), np.float32); et = np.ascontiguousarray(rng.uniform(0.5, 1.6, n), np.float32)
    ref.ref_refract_m(P(a3), P(b3), P(ct), P(et), C.c_int(n), P(o3)); g["refract"] = dict(ct=u(ct), et=u(et), y=u(o3))
    up = a3.copy(); up[:, 2] = np.abs(up[:, 2]); mm = b3.copy(); mm[:, 2] = np.abs(mm[:, 2])
    g["mf_in"] = dict(wi=u(up), m=u(mm), u=u(u2))
    for t, (au, av) in [(0, (0.1, 0.3)), (1, (0.25, 0.25))]:
        ref.ref_mf_eval(C.c_int(t), C.c_float(au), C.c_float(av), P(mm), C.c_int(n), P(o1)); e = u(o1)
        ref.ref_mf_smith_g1(C.c_int(t), C.c_float(au), C.c_float(av), P(up), P(mm), C.c_int(n), P(o1)); g1 = u(o1)
        ref.ref_mf_pdf(C.c_int(t), C.c_float(au), C.c_float(av), P(up), P(mm), C.c_int(n), P(o1)); pdf = u(o1)
        ref.ref_mf_sample(C.c_int(t), C.c_float(au), C.c_float(av), P(up), P(u2), C.c_int(n), P(o3), P(o1))
        g[f"mf_{t}"] = dict(au=au, av=av, eval=e, g1=g1, pdf=pdf, sample=u(o3), sample_pdf=u(o1))
    nodes = np.zeros(32, np.float32); wts = np.zeros(32, np.float32); ref.ref_gauss_legendre(C.c_int(32), P(nodes), P(wts)); g["gauss_legendre32"] = dict(nodes=u(nodes), weights=u(wts))
    ext = np.zeros(64, np.float32); ir = np.zeros(1, np.float32)
    ref.ref_rough_tables(C.c_int(0), C.c_float(0.1), C.c_float(1.5046 / 1.000277), C.c_int(64), P(ext), P(ir)); g["rough_tables_beckmann_0.1"] = dict(ext=u(ext), int_refl=u(ir))
    # camera set-up: sample_to_camera for the cbox sub-camera, grid placement, sample_ray
    c2s = np.zeros(16, np.float32); s2c = np.zeros(16, np.float32)
    ref.ref_sample_to_camera(C.c_int(256), C.c_int(256), C.c_float(39.3077), C.c_float(10), C.c_float(2800), C.c_float(0), P(c2s), P(s2c)); g["s2c_cbox"] = dict(c2s=u(c2s), s2c=u(s2c))
    ref.ref_sample_to_camera(C.c_int(512), C.c_int(384), C.c_float(55.5), C.c_float(0.1), C.c_float(100), C.c_float(0.125), P(c2s), P(s2c)); g["s2c_shift"] = dict(c2s=u(c2s), s2c=u(s2c))
    m = np.ascontiguousarray(rng.normal(size=(4, 4)) + 3 * np.eye(4), np.float32); inv = np.zeros(16, np.float32); it = np.zeros(16, np.float32)
    ref.ref_mat_inverse(P(m), P(inv)); ref.ref_mat_inverse_transpose(P(m), P(it)); g["mat_inverse"] = dict(m=u(m), inv=u(inv), inv_t=u(it))
    pos = np.ascontiguousarray(rng.uniform(size=(n, 2)), np.float32); tw = np.eye(4, dtype=np.float32); tw[:3, 3] = [278, 273, -800]
    ro = np.zeros((n, 3), np.float32); rd = np.zeros((n, 3), np.float32); mt = np.zeros(n, np.float32)
    ref.ref_perspective_sample_ray(P(s2c), P(tw), C.c_float(0.1), C.c_float(100), P(pos), C.c_int(n), P(ro), P(rd), P(mt))
    g["sample_ray"] = dict(s2c=u(s2c), to_world=u(tw), near=0.1, far=100.0, pos=u(pos), o=u(ro), d=u(rd), maxt=u(mt))
    p0, p1, p2 = (np.ascontiguousarray(rng.uniform(-1, 1, (n, 3)), np.float32) for _ in range(3))
    oo = np.ascontiguousarray(rng.uniform(-2, 2, (n, 3)), np.float32)
    aim = (p0 + p1 + p2) / 3 + rng.normal(scale=0.35, size=(n, 3)); dd = aim - oo; dd = np.ascontiguousarray(dd / np.linalg.norm(dd, axis=1, keepdims=True), np.float32); mx = np.full(n, 1e30, np.float32)
    tuv = np.zeros((n, 3), np.float32); hit = np.zeros(n, np.int32)
    ref.ref_moeller_trumbore(P(oo), P(dd), P(mx), P(p0), P(p1), P(p2), C.c_int(n), P(tuv), P(hit, i32p))
    g["moeller_trumbore"] = dict(o=u(oo), d=u(dd), p0=u(p0), p1=u(p1), p2=u(p2), tuv=u(tuv), hit=hit.tolist())
    g["constants"] = [float(ref.ref_const(C.c_int(i))) for i in range(6)]
    # ---- round 2 additions (appended so that the seeded inputs above stay what they were)
    for op, r in {12: (-1, 1), 13: (-3, 3), 14: (-3, 3), 15: (-3, 3), 16: (-1.5, 1.5)}.items():       # acos, atan2 (three argument patterns), safe_acos
        x = rng.uniform(r[0], r[1], n).astype(np.float32); x[:6] = [0, 1, -1, 0.5, -0.5, 0.3]
        o = np.zeros(n, np.float32); ref.ref_math1(C.c_int(op), P(x), C.c_int(n), P(o)); g["math1"][str(op)] = dict(x=u(x), y=u(o))
    ref.ref_square_to_uniform_sphere(P(u2), C.c_int(n), P(o3)); g["uniform_sphere"] = dict(y=u(o3))
    v4 = np.ascontiguousarray(rng.uniform(0.01, 2, (n, 4)), np.float32); v4[:20, 1] = v4[:20, 0]; v4[:20, 3] = v4[:20, 2]      # some flat patches
    pe = np.zeros(n, np.float32); ref.ref_square_to_bilinear(P(v4), P(u2), C.c_int(n), P(o2), P(o1), P(pe)); g["bilinear"] = dict(v=u(v4), y=u(o2), pdf=u(o1), pdf_eval=u(pe))
    # Hierarchical2D<Float, 0> over an odd-sized positive array (the envmap's importance sampler): sample, its pdf, eval at the sample
    sx, sy = 37, 21; data = np.ascontiguousarray(rng.uniform(0, 1, (sy, sx)) ** 4, np.float32); hu = np.ascontiguousarray(rng.uniform(size=(400, 2)), np.float32)
    huv = np.zeros((400, 2), np.float32); hp = np.zeros(400, np.float32); he = np.zeros(400, np.float32)
    ref.ref_hier2d(P(data), C.c_uint32(sx), C.c_uint32(sy), P(hu), C.c_int(400), P(huv), P(hp), P(he))
    g["hier2d"] = dict(sx=sx, sy=sy, data=u(data), u=u(hu), uv=u(huv), pdf=u(hp), eval=u(he))
    # DiscreteDistribution<Float>::sample_reuse_pmf (the area CDF of Mesh::sample_position), unnormalised pmf with zero entries
    pmf = np.ascontiguousarray(rng.uniform(0, 3, 73) ** 2, np.float32); pmf[[0, 7, 8, 72]] = 0; du = np.ascontiguousarray(rng.uniform(size=n), np.float32); du[:3] = [0, 0.5, 0.99999994]
    di = np.zeros(n, np.uint32); dr_ = np.zeros(n, np.float32); dp = np.zeros(n, np.float32); ds = np.zeros(1, np.float32); dn = np.zeros(1, np.float32)
    ref.ref_discrete_sample_reuse(P(pmf), C.c_uint32(len(pmf)), P(du), C.c_int(n), P(di, u32p), P(dr_), P(dp), P(ds), P(dn))
    g["discrete"] = dict(pmf=u(pmf), u=u(du), index=di.tolist(), reused=u(dr_), pmf_out=u(dp), sum=u(ds), normalization=u(dn))
    # GridSensor placement (src/sensors/grid.cpp:208-226): inverse(transpose(inverse_transpose + offset row)) for 8 cameras on a line
    tw = np.eye(4, dtype=np.float32); tw[:3, 3] = [278, 273, -800]; tw[0, 0] = tw[2, 2] = -1
    it = np.ascontiguousarray(np.linalg.inv(tw.astype(np.float64)).T, np.float32); mats = []
    for i in range(8):
        dt = np.float32(i) / np.float32(7); off = np.array([np.float32(1) * (np.float32(400) * (dt - np.float32(0.5))), 0, 0], np.float32); om = np.zeros(16, np.float32)   # grid.cpp:209-210 in float32
        ref.ref_grid_to_world(P(tw), P(it), P(off), P(om)); mats.append(dict(offset=u(off), m=u(om)))
    g["grid_to_world"] = dict(inv_t=u(it), mats=mats)
    json.dump(g, open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json"), "w"))
    print("wrote", len(g), "entries")


if __name__ == "__main__":
    main()
