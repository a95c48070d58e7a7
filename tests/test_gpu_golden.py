"""GPU tests of round 2 that pin the CUDA render loop to fixtures the REFERENCE owns (through the C ABI):

* the Z-test of src/render/tests/test_renders.py:152-235 on the reference's golden renders of its own scene files
  (bsdf_spheres/test_{roughplastic_linear,roughplastic_nonlinear,roughdielectric}_path.xml, infinite_mirrors/scene_path.xml: area + envmap +
  point emitters, wavelength spectra, roughplastic / roughdielectric / conductor BSDFs), loaded unmodified by amvpt_scene_from_xml,
  rendered with the reference's 2 M-sample budget by both production schedulers;
* per-lane bit-exactness against the CPU oracle on those scenes (rays, hit primitives, radiance) -- the emitter code, acos / atan2 and the
  Hierarchical2D warp run on both sides in the same canonical arithmetic;
* parity at the sizes that are benchmarked: the 1 002 528-triangle height field (closest + any hit, 200 k rays vs the oracle's BVH, a few
  thousand vs brute force) and lane windows at the start / middle / end of a cbox8x8 512^2 pass and of an hfield16 pass.
"""
import os

import numpy as np
import pytest

from test_cpu_golden import RENDER_SCENES, SCENE_DIR, load_render_case, z_test_passes
from test_gpu_parity import assert_bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden(pkg, ob, gpu):
    cache = {}

    def get(name):
        if name not in cache:
            scene, ref, var, spp = load_render_case(pkg, name)
            o = ob.OracleScene.from_product(scene); scene.commit(gpu)
            cache[name] = (scene, o, ref, var, spp)
        return cache[name]
    return get


@pytest.mark.parametrize("sched", ["wavefront", "regenerate"])
@pytest.mark.parametrize("name", sorted(RENDER_SCENES))
def test_gpu_render_passes_reference_z_test(golden, name, sched):
    scene, _, ref, var, spp = golden(name)
    img = scene.render(seed=0, spp=spp, scheduler=sched)
    st = scene.last_stats; spp_done = st["spp_done"]
    assert spp_done == spp // 16 * 16 and st["scheduler_used"] == (3 if sched == "wavefront" else 1)
    frac, pmin = z_test_passes(img, spp_done, ref, var)
    assert frac >= 0.9975, f"{name} [{sched}]: only {frac:.4%} of the pixels pass the reference's Z-test (min p-value {pmin:.3g})"


@pytest.mark.parametrize("name", sorted(RENDER_SCENES))
def test_golden_scene_lanes_bit_exact(golden, name):
    scene, o, ref, _, _ = golden(name)
    n = 20000; total = ref.shape[0] * ref.shape[1] * 16
    for seed_value, begin in [(0, 0), (32, total - n)]:
        a = scene.trace_lanes(seed_value, 16, begin, n, max_hits=8); b = o.trace_lanes(seed_value, 16, begin, n, max_hits=8)
        for k in ("ray_o", "ray_d", "hit_prim", "hit_t", "radiance"):
            assert_bits(a[k], b[k], f"{name}: {k}")
        assert (a["radiance"][:, 3] == 1).all()            # an environment emitter makes every ray valid (mvpath_single.h:98)
        assert (a["radiance"][:, :3].sum(1) > 0).mean() > 0.3           # the museum map has black regions


def test_golden_scene_film_and_schedulers(golden):
    """Film against the oracle's film, and the two schedulers against each other sample by sample (1 spp: the film is a per-lane tap)."""
    scene, o, _, _, _ = golden("roughdielectric")
    scene.render(seed=3, spp=16, develop=False, scheduler="wavefront"); film = scene.film_raw()
    ofilm, _, _ = o.render(seed=3, spp=16)
    assert np.array_equal(film[..., 3], ofilm[..., 3])
    assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= 2e-5 * ofilm[..., :3].max()
    scene.render(seed=9, spp=1, develop=False, scheduler="wavefront"); a = scene.film_raw().copy()
    scene.render(seed=9, spp=1, develop=False, scheduler="regenerate"); b = scene.film_raw()
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_env_warp_and_emitter_math_bit_exact(pkg, ob, golden, gpu):
    scene, o, _, _, _ = golden("infinite_mirrors")
    rng = np.random.default_rng(7); u = rng.uniform(size=(100000, 2)).astype(np.float32); u[:4] = [[0, 0], [1, 1], [0.5, 0.999999], [1e-7, 0.3]]
    uv, pdf, ev = scene.env_warp(u)
    ouv = np.zeros_like(u); opdf = np.zeros(len(u), np.float32); oev = np.zeros(len(u), np.float32)
    import ctypes as C
    P = lambda x: x.ctypes.data_as(ob.f32p)
    o.o.orc_env_warp(o.h, P(u), C.c_int(len(u)), P(ouv), P(opdf), P(oev))
    assert_bits(uv, ouv, "Hierarchical2D sample"); assert_bits(pdf, opdf, "Hierarchical2D pdf"); assert_bits(ev, oev, "Hierarchical2D eval")
    assert np.allclose(pdf, ev, rtol=2e-3, atol=1e-6)          # the density returned with a sample is the density eval() reports there
    for op, lo, hi in [(12, -1, 1), (13, -3, 3), (14, -3, 3), (15, -3, 3), (16, -1.5, 1.5)]:
        x = rng.uniform(lo, hi, 50000).astype(np.float32); x[:5] = [0, 1, -1, 0.5, -0.5]
        assert_bits(pkg.math1(op, x, gpu), ob.math1(op, x), f"math op {op}")


def test_point_and_constant_emitters_programmatic(pkg, scenes, ob, gpu):
    """PointLight + ConstantBackgroundEmitter + an area light in one scene built through the C ABI; shutter-time draw enabled: lanes bit-exact
    against the oracle, hide_emitters zeroes the rays that see only the background."""
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=32, sample_count=16)
    s = scenes.build(desc, commit=False)
    s.add_emitter_point((278, 400, 279.5), (8e4, 6e4, 3e4)); s.add_emitter_constant((0.3, 0.4, 0.6)); s.set_shutter(0.0, 0.5)
    assert [e["type"] for e in s.emitters()] == ["area", "point", "constant"] and s.shutter_draw()
    s.commit(gpu); o = ob.OracleScene.from_product(s)
    a = s.trace_lanes(0, 16, 0, 30000, max_hits=6); b = o.trace_lanes(0, 16, 0, 30000, max_hits=6, brute=True)
    for k in ("sample_pos", "ray_o", "ray_d", "hit_prim", "radiance"):
        assert_bits(a[k], b[k], k)
    # the time draw shifts the stream: without it the jitter is the same but everything after differs
    s.set_shutter(0.0, 0.0); c = s.trace_lanes(0, 16, 0, 30000, max_hits=6)
    assert np.array_equal(a["sample_pos"], c["sample_pos"]) and not np.array_equal(a["radiance"], c["radiance"])
    s.set_integrator(max_depth=6, hide_emitters=True); d = s.trace_lanes(0, 16, 0, 30000, max_hits=6)
    miss = d["hit_prim"][:, 0] == 0xffffffff
    assert miss.any() and (d["radiance"][miss] == 0).all() and (c["radiance"][miss, 3] == 1).all()
    with pytest.raises(pkg.AmvptError, match="Only one environment emitter"):
        s.add_emitter_constant(1.0)


# ------------------------------------------------------------------ parity at the benchmarked sizes
@pytest.fixture(scope="module")
def big_hfield(scenes, ob, gpu):
    desc = scenes.height_field(n=708, grid_x=4, grid_y=4, view_res=1024, max_depth=8, sample_count=16)      # bench.py `hfield16`
    return desc, scenes.build(desc, device=gpu), ob.OracleScene(desc)


def test_million_triangle_bvh_closest_and_any_hit(big_hfield):
    desc, s, o = big_hfield
    assert s.info()["n_prims"] == 1002528 + 14            # 708 x 708 quads + the enclosing box (12) and its light (2)
    depth, leaves = s.bvh_validate(); assert depth < 64
    rng = np.random.default_rng(11); n = 200000
    # rays from the cameras' side and from inside the box, towards points on / near the terrain, plus grazing and axis-parallel ones
    origin = np.stack([rng.uniform(-1.4, 1.4, n), rng.uniform(-1.4, 1.4, n), rng.uniform(-0.2, 2.4, n)], 1).astype(np.float32)
    target = np.stack([rng.uniform(-1.2, 1.2, n), rng.uniform(-1.2, 1.2, n), rng.uniform(-0.25, 0.25, n)], 1).astype(np.float32)
    d = target - origin; d[:2000, 2] = 0; d[2000:3000, :2] = 0; d[2000:3000, 2] = -1
    d = (d / np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-20)).astype(np.float32)
    mt = np.full(n, np.finfo(np.float32).max, np.float32); mt[::7] = rng.uniform(0.05, 3, len(mt[::7])).astype(np.float32)
    prim, tuv, _ = s.intersect(origin, d, mt); rprim, rtuv, _ = o.intersect(origin, d, mt)
    assert_bits(prim, rprim, "closest-hit primitive (1M triangles)"); assert_bits(tuv[prim != 0xffffffff], rtuv[rprim != 0xffffffff], "closest-hit t, u, v")
    assert 0.5 < (prim != 0xffffffff).mean() < 1.0
    occ, _, _ = s.intersect(origin, d, mt, any_hit=True); rocc = o.ray_test(origin, d, mt)
    assert np.array_equal(occ != 0xffffffff, rocc != 0)
    # the compressed 4-wide tree of the wavefront trace stage finds the same hits
    nq, dq, lq = s.bvh_q4_validate(); assert nq > 200000 and lq == leaves and dq <= 20
    qprim, qtuv, qstats = s.intersect(origin, d, mt, q4=True)
    assert_bits(qprim, rprim, "closest-hit primitive, compressed 4-wide tree"); assert_bits(qtuv[qprim != 0xffffffff], rtuv[rprim != 0xffffffff], "t, u, v, compressed 4-wide tree")
    qocc, _, _ = s.intersect(origin, d, mt, any_hit=True, q4=True); assert np.array_equal(qocc != 0xffffffff, rocc != 0)
    k = 3000; bprim, btuv, _ = o.intersect(origin[:k], d[:k], mt[:k], brute=True)
    assert_bits(prim[:k], bprim, "closest hit vs brute force (1M triangles)"); assert_bits(tuv[:k][bprim != 0xffffffff], btuv[bprim != 0xffffffff], "t, u, v vs brute force")


def test_lane_windows_of_hfield16_pass(big_hfield):
    desc, s, o = big_hfield
    total = 4096 * 4096 * 16; n = 4096
    for begin in (0, total // 2 + 12345, total - n):
        a = s.trace_lanes(0, 16, begin, n, max_hits=8); b = o.trace_lanes(0, 16, begin, n, max_hits=8)
        for k in ("ray_o", "ray_d", "view", "pixel", "hit_prim", "hit_t", "radiance"):
            assert_bits(a[k], b[k], f"hfield16 lanes at {begin}: {k}")


def test_lane_windows_of_cbox8x8_pass(scenes, ob, gpu):
    desc = scenes.cornell_box(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, max_depth=6, sample_count=256)      # bench.py `cbox8x8`
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    total = 4096 * 4096 * 16; n = 20000
    assert total == 268435456
    for begin in (0, 130000000, total - n):
        a = s.trace_lanes(0, 16, begin, n, max_hits=6); b = o.trace_lanes(0, 16, begin, n, max_hits=6, brute=True)
        for k in ("sample_pos", "ray_o", "ray_d", "view", "pixel", "hit_prim", "hit_t", "radiance"):
            assert_bits(a[k], b[k], f"cbox8x8 lanes at {begin}: {k}")
    # the pass seeds of the full render: seed_value = sampler seed + spp_per_pass * pass + render seed (mvpath.cpp:226-231)
    a = s.trace_lanes(16 * 15, 16, total - n, n); b = o.trace_lanes(16 * 15, 16, total - n, n, brute=True)
    assert_bits(a["radiance"], b["radiance"], "cbox8x8 last pass, last lanes")


# ------------------------------------------------------------------ chi-square tests of BSDF sampling through the C ABI (reference method, tests/chi2.py)
from test_cpu_golden import CHI2_CASES, chi2_scene


@pytest.mark.parametrize("name", sorted(CHI2_CASES))
def test_gpu_bsdf_sampling_chi2(scenes, gpu, name):
    from chi2 import bsdf_chi2
    cfg, wi, kw = CHI2_CASES[name]
    s = scenes.build(chi2_scene(scenes, cfg), device=gpu)
    p, info = bsdf_chi2(lambda a, b, c, d: s.bsdf_eval_pdf_sample(0, a, b, c, d), wi, **kw)
    alpha = 1.0 - (1.0 - 0.01) ** (1.0 / len(CHI2_CASES))
    assert p > alpha, f"{name}: chi^2 rejects amvpt_bsdf_eval_pdf_sample's sample against its pdf: p = {p:.3g}, {info}"
    assert abs(info["hist_sum"] - info["pdf_sum"]) < 5e-3
