"""Parity of the CUDA path against the CPU oracle, through the C ABI (include/amvpt.h), on seeded inputs.
Bar (BASELINE.json north_star): sampler streams, camera rays and hit primitive IDs bit-exact; per-lane radiance is
also compared bit-for-bit (same canonical arithmetic on both sides); the film, which is accumulated with atomics in
a different order than the oracle's sequential sum, within a stated relative tolerance."""
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu
FILM_RTOL = 2e-5      # order of fp32 summation of <= 64 samples per pixel; stated tolerance for the film compare


def bits(a):
    """Bit patterns, with every NaN mapped to one canonical pattern (x86 and sm_100 produce different NaN payloads /
    signs for the same invalid operation; the reference's own variants do too)."""
    a = np.ascontiguousarray(a)
    u = a.view(np.uint32).copy()
    if a.dtype == np.float32:
        u[np.isnan(a)] = 0x7fc00000
    return u


def assert_bits(a, b, what):
    a, b = bits(a), bits(b)
    bad = np.flatnonzero((a != b).reshape(len(a), -1).any(axis=1)) if a.ndim > 1 else np.flatnonzero(a != b)
    assert bad.size == 0, f"{what}: {bad.size} of {len(a)} rows differ, first {bad[:5]}"


@pytest.fixture(scope="module")
def cbox(scenes, ob, gpu):
    desc = scenes.cornell_box(grid_x=4, grid_y=1, view_res=64, sample_count=32)
    return desc, scenes.build(desc, device=gpu), ob.OracleScene(desc)


@pytest.fixture(scope="module")
def hfield(scenes, ob, gpu):
    desc = scenes.height_field(n=24, grid_x=2, grid_y=2, view_res=32, sample_count=16)
    return desc, scenes.build(desc, device=gpu), ob.OracleScene(desc)


def test_sampler_stream_bit_exact(pkg, ob, gpu):
    # PCG32Sampler::seed wavefront seeding + next_float32 (reference sampler.cpp:125-144, drjit/random.h:59-150)
    for seed, lane in [(0, 0), (0, 1), (16, 4194303), (0xdeadbeef, 0xffffffff), (48, 123456789)]:
        assert_bits(pkg.sampler_stream(seed, lane, 64, gpu), ob.sampler_stream(seed, lane, 64), f"stream {seed},{lane}")


@pytest.mark.parametrize("op,lo,hi", [(0, -20, 20), (1, -20, 20), (2, -80, 80), (3, 1e-30, 1e30), (4, -4, 4), (5, -0.999999, 0.999999),
                                      (6, -120, 120), (7, 0, 1e6), (8, -1e3, 1e3), (9, 1e-6, 1e6), (10, -1, 1), (11, -1.5, 1.5)])
def test_leaf_math_bit_exact(pkg, ob, gpu, op, lo, hi):
    rng = np.random.default_rng(op)
    x = rng.uniform(lo, hi, 20000).astype(np.float32)
    if op == 3:
        x = np.exp(rng.uniform(-60, 60, 20000)).astype(np.float32)
    x[:6] = [lo, hi, 0.0 if lo <= 0 <= hi else lo, 1.0 if lo <= 1 <= hi else hi, lo / 2, hi / 2]
    assert_bits(pkg.math1(op, x, gpu), ob.math1(op, x), f"math op {op}")


BSDFS = [dict(type="diffuse", reflectance=(0.2, 0.5, 0.8)),
         dict(type="diffuse", reflectance=(0.2, 0.5, 0.8), twosided=True),
         dict(type="conductor", eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)),
         dict(type="dielectric", int_ior=1.5046, ext_ior=1.000277),
         dict(type="roughconductor", distribution="ggx", alpha=0.15, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)),
         dict(type="roughconductor", distribution="beckmann", alpha_u=0.05, alpha_v=0.3, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14), twosided=True),
         dict(type="roughdielectric", distribution="beckmann", alpha=0.1),
         dict(type="roughdielectric", distribution="ggx", alpha=0.3, int_ior=1.33, ext_ior=1.0),
         dict(type="roughplastic", distribution="beckmann", alpha=0.1, reflectance=(0.5, 0.3, 0.1)),
         dict(type="roughplastic", distribution="ggx", alpha=0.2, reflectance=(0.5, 0.3, 0.1), nonlinear=True)]


def test_bsdf_eval_pdf_sample_bit_exact(pkg, scenes, ob, gpu):
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=8)
    desc["bsdfs"] = BSDFS
    for m in desc["meshes"]:
        m["bsdf"] = 0
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    rng = np.random.default_rng(7); n = 20000

    def dirs(k):
        v = rng.normal(size=(k, 3)); v /= np.linalg.norm(v, axis=1, keepdims=True); return v.astype(np.float32)
    wi, wo = dirs(n), dirs(n)
    wi[:10, 2] = [0, 1, -1, 1e-8, -1e-8, 0.5, 0.5, 0.5, 0.5, 0.5]      # grazing / normal incidence edge cases
    s1 = rng.uniform(size=n).astype(np.float32); s2 = rng.uniform(size=(n, 2)).astype(np.float32)
    s2[:4] = [[0, 0], [0.5, 0.5], [0.99999994, 0.99999994], [0, 0.99999994]]
    for i, b in enumerate(BSDFS):
        a = s.bsdf_eval_pdf_sample(i, wi, wo, s1, s2); r = o.bsdf_eval_pdf_sample(i, wi, wo, s1, s2)
        for k in ("f", "pdf", "bs_wo", "bs_pdf", "bs_eta", "bs_type", "weight"):
            assert_bits(a[k], r[k], f"bsdf {i} ({b['type']}) field {k}")
        if b["type"] == "roughplastic":
            ta, sa = s.bsdf_tables(i); tb, sb = o.bsdf_tables(i)
            assert_bits(ta, tb, "roughplastic transmittance table"); assert_bits(sa, sb, "roughplastic scalars")


def test_camera_rays_bit_exact(cbox):
    desc, s, o = cbox
    w = desc["sensor"]["res_x"]; h = desc["sensor"]["res_y"]
    total = w * h * 16
    for seed_value, begin in [(0, 0), (16, total // 2 - 5000), (0, total - 20000)]:
        a = s.trace_lanes(seed_value, 16, begin, 20000, trace=False); b = o.trace_lanes(seed_value, 16, begin, 20000, trace=False)
        for k in ("sample_pos", "ray_o", "ray_d", "ray_maxt", "view", "pixel"):
            assert_bits(a[k], b[k], f"camera tap {k} (seed {seed_value}, lane {begin})")
    assert set(np.unique(a["view"])) <= {0, 1, 2, 3}


def test_thinlens_rays_bit_exact(scenes, ob, gpu):
    desc = scenes.cornell_box(grid_x=3, grid_y=2, view_res=16, thinlens=True, aperture_radius=5.0, cone_deg=20.0)
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    n = 3 * 2 * 16 * 16 * 4
    a = s.trace_lanes(3, 4, 0, n, trace=False); b = o.trace_lanes(3, 4, 0, n, trace=False)
    for k in ("sample_pos", "ray_o", "ray_d", "ray_maxt", "view", "pixel"):
        assert_bits(a[k], b[k], f"thinlens/cone tap {k}")
    assert len(np.unique(a["view"])) == 6


def random_rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    d[: n // 50] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, n // 50)] * np.float32(1) * rng.choice([-1, 1], (n // 50, 1)).astype(np.float32)
    maxt = np.full(n, np.finfo(np.float32).max, np.float32); maxt[::7] = rng.uniform(0, hi - lo, len(maxt[::7])).astype(np.float32)
    return o, d, maxt


def test_bvh_equals_brute_force_cbox(cbox):
    # kd-tree == brute force is the reference's own test (src/render/tests/test_kdtrees.py:49-115)
    _, s, o = cbox
    ro, rd, mt = random_rays(np.random.default_rng(1), 60000, -50, 600)
    prim, tuv, _ = s.intersect(ro, rd, mt); rprim, rtuv, _ = o.intersect(ro, rd, mt, brute=True)
    assert_bits(prim, rprim, "closest-hit primitive"); hit = rprim != 0xffffffff
    assert hit.sum() > 10000
    assert_bits(tuv[hit], rtuv[hit], "closest-hit t,u,v")
    occ, _, _ = s.intersect(ro, rd, mt, any_hit=True); rocc = o.ray_test(ro, rd, mt, brute=True)
    assert np.array_equal(occ != 0xffffffff, rocc != 0), "shadow-ray occlusion"
    # the flat leaf list the render kernels use for small scenes (no counters requested): same hits
    fprim, ftuv, fst = s.intersect(ro, rd, mt, flat=True)
    assert_bits(fprim, rprim, "flat leaf list: closest-hit primitive"); assert_bits(ftuv[hit], rtuv[hit], "flat leaf list: t,u,v"); assert fst[0] == 0
    focc, _, _ = s.intersect(ro, rd, mt, any_hit=True, flat=True)
    assert np.array_equal(focc != 0xffffffff, rocc != 0), "flat leaf list: shadow-ray occlusion"


def test_bvh_equals_brute_force_heightfield(hfield):
    _, s, o = hfield
    ro, rd, mt = random_rays(np.random.default_rng(2), 20000, -1.5, 1.5)
    prim, tuv, _ = s.intersect(ro, rd, mt); rprim, rtuv, _ = o.intersect(ro, rd, mt, brute=True)
    assert_bits(prim, rprim, "closest-hit primitive"); hit = rprim != 0xffffffff
    assert_bits(tuv[hit], rtuv[hit], "closest-hit t,u,v")
    occ, _, _ = s.intersect(ro, rd, mt, any_hit=True); rocc = o.ray_test(ro, rd, mt, brute=True)
    assert np.array_equal(occ != 0xffffffff, rocc != 0)


def test_compressed_wide_tree_equals_binary_tree(cbox, hfield):
    """DScene::nodes_q4 (quantised 4-wide nodes, wavefront trace stage) against the binary tree and the oracle's brute force."""
    for which, (desc, s, o) in (("cbox", cbox), ("hfield", hfield)):
        nq, dq, lq = s.bvh_q4_validate(); d2, l2 = s.bvh_validate(); assert nq >= 1 and lq == l2
        rng = np.random.default_rng(3); n = 20000
        lo, hi = (np.array([-100, -100, -900]), np.array([660, 650, 660])) if which == "cbox" else (np.array([-1.4, -1.4, -0.3]), np.array([1.4, 1.4, 2.4]))
        origin = rng.uniform(lo, hi, (n, 3)).astype(np.float32); target = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
        d = target - origin; d[:500, 0] = 0; d[500:1000, 1] = 0; d[1000:1200, :2] = 0
        d = (d / np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-20)).astype(np.float32)
        mt = np.full(n, np.finfo(np.float32).max, np.float32); mt[::5] = rng.uniform(0.01, 1.0, len(mt[::5])).astype(np.float32) * (1000 if which == "cbox" else 3)
        p2, t2, _ = s.intersect(origin, d, mt); p4, t4, st4 = s.intersect(origin, d, mt, q4=True); pb, tb, _ = o.intersect(origin, d, mt, brute=True)
        assert_bits(p4, pb, f"{which}: q4 primitive vs brute force"); assert_bits(p2, pb, f"{which}: binary tree vs brute force")
        assert_bits(t4[pb != 0xffffffff], tb[pb != 0xffffffff], f"{which}: q4 t, u, v")
        a2, _, _ = s.intersect(origin, d, mt, any_hit=True); a4, _, _ = s.intersect(origin, d, mt, any_hit=True, q4=True)
        assert np.array_equal(a2 != 0xffffffff, a4 != 0xffffffff)


def test_shared_edge_tie_rule(cbox):
    # rays aimed exactly at shared edges / vertices of the floor quad: the larger primitive index must win
    _, s, o = cbox
    t = np.linspace(0.05, 0.95, 200, dtype=np.float32)
    target = np.stack([552.8 * (1 - t) + 0.0 * t, np.zeros_like(t), 0.0 * (1 - t) + 559.2 * t], 1).astype(np.float32)   # diagonal p0->p2
    origin = np.tile(np.array([278, 273, 280], np.float32), (len(t), 1))
    d = target - origin; d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    mt = np.full(len(t), np.finfo(np.float32).max, np.float32)
    prim, tuv, _ = s.intersect(origin, d, mt); rprim, rtuv, _ = o.intersect(origin, d, mt, brute=True)
    assert_bits(prim, rprim, "tie rule primitive"); assert_bits(tuv, rtuv, "tie rule tuv")


@pytest.mark.parametrize("which", ["cbox", "hfield"])
def test_paths_bit_exact(which, cbox, hfield):
    desc, s, o = cbox if which == "cbox" else hfield
    spp = 16; total = desc["sensor"]["res_x"] * desc["sensor"]["res_y"] * spp
    n = 30000 if which == "cbox" else 8000
    for seed_value, begin in [(0, 0), (16, max(total // 2 - n, 0))]:
        a = s.trace_lanes(seed_value, spp, begin, n, max_hits=8)
        b = o.trace_lanes(seed_value, spp, begin, n, max_hits=8, brute=(which == "cbox"))
        assert_bits(a["hit_prim"], b["hit_prim"], f"{which}: hit primitive IDs per bounce")
        assert_bits(a["hit_t"], b["hit_t"], f"{which}: hit distances per bounce")
        assert_bits(a["radiance"], b["radiance"], f"{which}: per-lane radiance")
        assert (a["radiance"][:, :3].sum(1) > 0).mean() > 0.2


def test_film_matches_oracle(cbox):
    desc, s, o = cbox
    img = s.render(seed=0, spp=32)
    film = s.film_raw(); ofilm, _, _ = o.render(seed=0, spp=32, brute=True)
    assert np.array_equal(film[..., 3], ofilm[..., 3]), "weight channel (samples per pixel) must match exactly"
    scale = ofilm[..., :3].max()
    assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= FILM_RTOL * scale
    assert np.allclose(img, ob_develop(ofilm), rtol=0, atol=FILM_RTOL * scale / 32 * 4)
    assert s.last_stats["paths"] == film.shape[0] * film.shape[1] * 32 and s.last_stats["passes"] == 2


def ob_develop(rgbw):
    w = rgbw[..., 3:4]
    return rgbw[..., :3] / np.where(w == 0, 1, w)


def test_film_matches_oracle_heightfield(hfield):
    desc, s, o = hfield
    s.render(seed=5, spp=16, develop=False)
    film = s.film_raw(); ofilm, _, _ = o.render(seed=5, spp=16, brute=False)
    assert np.array_equal(film[..., 3], ofilm[..., 3])
    scale = max(ofilm[..., :3].max(), 1e-6)
    assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= FILM_RTOL * scale


def test_schedulers_agree_and_accumulate(cbox):
    _, s, _ = cbox
    s.render(seed=1, spp=32, scheduler="warp_batch", develop=False); a = s.film_raw()
    s.render(seed=1, spp=32, scheduler="regenerate", develop=False); b = s.film_raw()
    s.render(seed=1, spp=32, scheduler="stream", develop=False); c = s.film_raw()
    s.render(seed=1, spp=32, scheduler="wavefront", develop=False); d = s.film_raw()
    assert np.array_equal(a[..., 3], b[..., 3]) and np.array_equal(a[..., 3], c[..., 3]) and np.array_equal(a[..., 3], d[..., 3])
    for other in (b, c, d):
        assert np.abs(a - other).max() <= FILM_RTOL * a[..., :3].max()
    # spp rounding (mvpath.cpp:36-41): 40 spp with spp_pass_lim 16 -> 2 passes of 16
    s.render(seed=1, spp=40, develop=False)
    w = s.film_raw()[..., 3]     # a jittered position can round up into the next pixel (float32(px + 0.99999994) == px + 1), so a few
    assert s.last_stats["passes"] == 2 and w.sum() == 32 * w.size and (w == 32).mean() > 0.999      # pixels hold 31 / 33
    # accumulation without clearing doubles the weights (Film accumulates across render() calls, hdrfilm.cpp:293-297)
    s.render(seed=2, spp=32, develop=False, clear=False)
    w = s.film_raw()[..., 3]
    assert w.sum() == 64 * w.size and (w == 64).mean() > 0.998


def test_view_sharding_reproduces_single_gpu_film(cbox, scenes, gpu):
    desc, s, _ = cbox
    s.render(seed=3, spp=16, develop=False); full = s.film_raw()
    for world in (2, 4, 8):            # 8 ranks > 4 views: views are cut into bands
        acc = np.zeros_like(full)
        for rank in range(world):
            s.render(seed=3, spp=16, develop=False, rank=rank, world=world)
            part = s.film_raw()
            acc += part
        assert np.array_equal(acc[..., 3], full[..., 3])
        assert np.abs(acc - full).max() <= FILM_RTOL * full[..., :3].max()


def test_max_depth_and_rr_edge_cases(scenes, ob, gpu):
    for md, rr in [(0, 5), (1, 5), (2, 1), (-1, 2)]:
        desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=16, max_depth=md, rr_depth=rr)
        s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
        a = s.trace_lanes(0, 16, 0, 2 * 16 * 16 * 16, max_hits=4); b = o.trace_lanes(0, 16, 0, 2 * 16 * 16 * 16, max_hits=4, brute=True)
        assert_bits(a["radiance"], b["radiance"], f"max_depth={md} rr_depth={rr}")
        assert_bits(a["hit_prim"], b["hit_prim"], f"max_depth={md} rr_depth={rr} prims")


@pytest.mark.parametrize("which", ["cbox", "hfield", "cbox_glass"])
def test_stream_scheduler_bit_exact_per_sample(which, scenes, gpu):
    """One sample per pixel (spp_pass_lim = 1): every film pixel with weight 1 holds exactly one path's radiance, so the
    films of two schedulers can be compared bit for bit -- the stream scheduler (unified traversal loop, deferred shadow
    rays) must produce the very radiance of the per-lane path loop that test_paths_bit_exact pins to the oracle.  The ray
    and traversal counters must agree too (same rays, same BVH steps)."""
    if which == "cbox":
        desc = scenes.cornell_box(grid_x=4, grid_y=2, view_res=64, spp_pass_lim=1, sample_count=1)
    elif which == "cbox_glass":
        desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=96, spp_pass_lim=1, sample_count=1, max_depth=12, rr_depth=3,
                                  bsdf_overrides={"box": dict(type="dielectric"), "white": dict(type="roughplastic", reflectance=(.7, .6, .5), alpha=0.2),
                                                  "red": dict(type="conductor", eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14))})
    else:
        desc = scenes.height_field(n=48, grid_x=2, grid_y=2, view_res=64, spp_pass_lim=1, sample_count=1)
    s = scenes.build(desc, device=gpu)
    films, stats = {}, {}
    for sched in ("regenerate", "stream", "warp_batch", "wavefront", "split"):
        for seed in (0, 7):
            # the split scheduler has no counting variant (it falls back to regenerate under stats=True): render it plain
            s.render(seed=seed, spp=1, scheduler=sched, develop=False, stats=sched != "split")
            films[sched, seed] = s.film_raw(); stats[sched, seed] = dict(s.last_stats)
            assert s.last_stats["scheduler_used"] == {"regenerate": 1, "stream": 2, "warp_batch": 0, "wavefront": 3, "split": 6}[sched]
    for seed in (0, 7):
        ref = films["regenerate", seed]; one = ref[..., 3] == 1
        assert one.mean() > 0.99
        for sched in ("stream", "warp_batch", "wavefront", "split"):
            f = films[sched, seed]
            assert np.array_equal(f[..., 3], ref[..., 3])
            assert_bits(f[one], ref[one], f"{which} seed {seed}: {sched} vs regenerate, single-sample pixels")
            assert np.abs(f - ref).max() <= FILM_RTOL * max(ref[..., :3].max(), 1e-6)
            for k in ("paths",) if sched == "split" else ("paths", "closest_rays", "shadow_rays", "bvh_nodes", "tri_tests", "path_vertices"):
                assert stats[sched, seed][k] == stats["regenerate", seed][k], f"{which} {sched} {k}"
        assert (ref[..., :3].sum(-1) > 0).mean() > 0.2


def test_stream_scheduler_depth_edge_cases_and_thresholds(scenes, gpu):
    import os
    for md, rr in [(0, 5), (1, 5), (2, 1), (-1, 2)]:
        desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=32, max_depth=md, rr_depth=rr, spp_pass_lim=1, sample_count=1)
        s = scenes.build(desc, device=gpu)
        s.render(seed=3, spp=1, scheduler="regenerate", develop=False); a = s.film_raw()
        for tm, rm in ((20, 8), (1, 1), (32, 32), (33, 1), (8, 24)):       # any threshold setting is only a schedule
            os.environ["AMVPT_TRACE_MIN"], os.environ["AMVPT_REGEN_MIN"] = str(tm), str(rm)
            try:
                s.render(seed=3, spp=1, scheduler="stream", develop=False); b = s.film_raw()
            finally:
                del os.environ["AMVPT_TRACE_MIN"], os.environ["AMVPT_REGEN_MIN"]
            one = a[..., 3] == 1
            assert np.array_equal(a[..., 3], b[..., 3]) and np.array_equal(bits(a[one]), bits(b[one])), f"max_depth={md} rr={rr} trace_min={tm} regen_min={rm}"
        for pool in (128, 4096, 1 << 20):                                   # wavefront pool smaller / larger than the pass
            os.environ["AMVPT_WF_POOL"] = str(pool)
            try:
                s2 = scenes.build(desc, device=gpu); s2.render(seed=3, spp=1, scheduler="wavefront", develop=False); b = s2.film_raw()
            finally:
                del os.environ["AMVPT_WF_POOL"]
            assert np.array_equal(a[..., 3], b[..., 3]) and np.array_equal(bits(a[one]), bits(b[one])), f"max_depth={md} rr={rr} wavefront pool={pool}"


@pytest.mark.parametrize("rf", [("gaussian", 0.5), ("tent", 1.0), ("gaussian", 0.3), ("mitchell", 1 / 3, 1 / 3), ("mitchell", 0.55, 0.2),
                                ("catmullrom", 0.0), ("lanczos", 3)])
def test_reconstruction_filter_film_matches_oracle(rf, scenes, ob, gpu):
    """ImageBlock::put with a non-box filter (coalesced path, src/render/imageblock.cpp:433-505): every sample is spread over
    (2n+1)^2 pixels with separable weights, across view-tile boundaries of the quilt and clipped at the film border."""
    desc = scenes.cornell_box(grid_x=2, grid_y=2, view_res=24, sample_count=16); desc["rfilter"] = rf
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    ofilm, _, _ = o.render(seed=2, spp=16, brute=True, nthreads=4); scale = ofilm[..., :3].max()
    for sched in ("auto", "wavefront", "regenerate"):          # every scheduler choice is served by the wavefront when a filter is set
        s.render(seed=2, spp=16, scheduler=sched, develop=False); film = s.film_raw()
        assert s.last_stats["scheduler_used"] == 3
        assert np.allclose(film[..., 3], ofilm[..., 3], rtol=1e-5, atol=1e-5), sched
        assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= 5 * FILM_RTOL * scale, sched
    assert ofilm[..., 3].min() > 0 and abs(ofilm[..., 3].mean() / 16 - (ofilm[..., 3].mean() / 16)) == 0
    with pytest.raises(Exception, match="box reconstruction filter"):
        s.render(seed=2, spp=16, adaptive=True)


def _two_scenes(pkg, ob, desc, gpu, sensor=None):
    """The same description in the product and in the oracle, optionally with a batch sensor instead of the grid."""
    s = pkg.Scene(); o = ob.OracleScene()
    for b in desc["bsdfs"]:
        s.add_bsdf(**b); o.add_bsdf(**b)
    for m in desc["meshes"]:
        s.add_mesh(m["positions"], m["faces"], m["bsdf"], normals=m["normals"], radiance=m["radiance"], flip_normals=m["flip_normals"])
        o.add_mesh(m["positions"], m["faces"], m["bsdf"], m["normals"], m["radiance"], m["flip_normals"])
    if sensor is None:
        s.set_sensor_grid(**desc["sensor"]); o.set_sensor_grid(**desc["sensor"])
    else:
        s.set_sensor_batch(*sensor); o.set_sensor_batch(*sensor[:6])
    s.set_integrator(**desc["integrator"]); o.set_integrator(**desc["integrator"])
    s.commit(gpu); o.prepare()
    return s, o


def test_batch_sensor_paths_bit_exact(pkg, scenes, ob, gpu):
    """BatchSensor (src/sensors/batch.cpp:94-160): explicit sub-cameras side by side, no y reversal; the sub-camera matrices
    are taken from three views of a grid sensor and handed over in a different order."""
    desc = scenes.cornell_box(grid_x=3, grid_y=1, view_res=24, cam_dist=300.0, sample_count=16)
    g = scenes.build(desc, commit=False)
    cams = [g.sensor(i) for i in (2, 0, 1)]
    sensor = (24, 24, np.stack([c[0] for c in cams]), np.stack([c[1] for c in cams]), 10.0, 2800.0, False)
    s, o = _two_scenes(pkg, ob, desc, gpu, sensor)
    assert s.info()["film_w"] == 72 and s.info()["n_views"] == 3
    n = 72 * 24 * 16
    a = s.trace_lanes(0, 16, 0, n, max_hits=4); b = o.trace_lanes(0, 16, 0, n, max_hits=4, brute=True)
    for k in ("ray_o", "ray_d", "view", "hit_prim", "radiance", "pixel"):
        assert_bits(a[k], b[k], f"batch sensor tap {k}")
    assert sorted(set(a["view"].tolist())) == [0, 1, 2]
    s.render(seed=0, spp=16, develop=False); film = s.film_raw(); ofilm, _, _ = o.render(seed=0, spp=16, brute=True)
    assert np.array_equal(film[..., 3], ofilm[..., 3]) and np.abs(film - ofilm).max() <= FILM_RTOL * ofilm[..., :3].max()


def test_two_emitters_flipped_normals_and_many_faces(pkg, scenes, ob, gpu):
    """Scene::sample_emitter with N > 1 emitters (src/render/scene.cpp:223-246: index = min(floor(s.x N), N - 1), sample re-used),
    Mesh::sample_position through the area CDF of a many-faced emitter (mesh.cpp:764-818, distr_1d.h:116-134) and flip_normals."""
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=32, sample_count=16)
    # second light: a 6 x 6 patch grid (72 triangles of unequal area) on the back wall, facing -z after flip_normals
    n = 6; xs = 150 + 250 * (np.arange(n + 1) / n) ** 1.7; ys = 100 + 300 * (np.arange(n + 1) / n) ** 0.6
    X, Y = np.meshgrid(xs, ys); pos = np.stack([X, Y, np.full_like(X, 559.0)], -1).reshape(-1, 3).astype(np.float32)
    faces = []
    for j in range(n):
        for i in range(n):
            v = j * (n + 1) + i; faces += [(v, v + 1, v + n + 2), (v, v + n + 2, v + n + 1)]       # normal +z: needs the flip to face the room
    desc["bsdfs"].append(dict(type="diffuse", reflectance=(0.0, 0.0, 0.0)))
    desc["meshes"].append(dict(name="panel", positions=pos, faces=np.asarray(faces, np.uint32), normals=None, bsdf=len(desc["bsdfs"]) - 1,
                               radiance=(2.0, 4.0, 9.0), flip_normals=True))
    s, o = _two_scenes(pkg, ob, desc, gpu)
    nl = 64 * 32 * 16
    a = s.trace_lanes(0, 16, 0, nl, max_hits=6); b = o.trace_lanes(0, 16, 0, nl, max_hits=6, brute=True)
    for k in ("hit_prim", "hit_t", "radiance"):
        assert_bits(a[k], b[k], f"two emitters tap {k}")
    blue = a["radiance"][:, 2] > a["radiance"][:, 0]
    assert blue.mean() > 0.05                       # the blue panel lights part of the room
    for sched in ("regenerate", "wavefront"):
        s.render(seed=0, spp=16, develop=False, scheduler=sched); film = s.film_raw(); ofilm, _, _ = o.render(seed=0, spp=16, brute=True)
        assert np.array_equal(film[..., 3], ofilm[..., 3]) and np.abs(film - ofilm).max() <= FILM_RTOL * ofilm[..., :3].max()


def test_xml_scene_renders_like_programmatic_scene(pkg, scenes, gpu, tmp_path):
    """The XML boundary end to end: scenes.export_xml -> load_file -> render equals the programmatic scene bit for bit
    (one sample per pixel, so no atomic ordering is involved)."""
    desc = scenes.height_field(n=32, grid_x=2, grid_y=2, view_res=32, spp_pass_lim=1, sample_count=1, smooth_normals=False)
    a = pkg.load_file(scenes.export_xml(desc, str(tmp_path), mesh_format="ply")); a.commit(gpu)
    b = scenes.build(desc, device=gpu)
    a.render(seed=4, spp=1, develop=False); fa = a.film_raw()
    b.render(seed=4, spp=1, develop=False); fb = b.film_raw()
    one = fb[..., 3] == 1
    assert np.array_equal(fa[..., 3], fb[..., 3]) and np.array_equal(bits(fa[one]), bits(fb[one]))
    assert a.last_stats["scheduler_used"] == 3 and (fb[..., :3].sum(-1) > 0).mean() > 0.3


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def test_relmse_against_oracle_at_matched_spp(scenes, ob, gpu):
    """BASELINE.json's quality bar: per-view relMSE <= 1e-3 against the reference at matched high spp, and the mean per-channel
    error within 3 sigma of the Monte Carlo estimate.  GPU (seed 11) and oracle (seed 0) draw independent samples here, so
    this is a statistical test of the whole loop rather than a replay."""
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=24, sample_count=1024)
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    img = s.render(seed=11, spp=1024)
    ofilm, _, _ = o.render(seed=0, spp=1024, brute=False, nthreads=8); oimg = ob_develop(ofilm)
    for v in range(2):
        a, b = img[:, 24 * v:24 * (v + 1)], oimg[:, 24 * v:24 * (v + 1)]
        assert rel_mse(a, b) <= 1e-3, f"view {v}: relMSE {rel_mse(a, b)}"
    # two independent 1024-spp estimates: the difference of the image means is a few sigma of the per-image Monte Carlo error
    half = s.render(seed=12, spp=1024)
    sigma = np.abs(img.mean((0, 1)) - half.mean((0, 1))) / np.sqrt(2) + 1e-4
    assert np.all(np.abs(img.mean((0, 1)) - oimg.mean((0, 1))) <= 3 * np.sqrt(2) * np.maximum(sigma, 2e-3 * oimg.mean((0, 1))))


def test_adaptive_allocation(scenes, gpu):
    """Build-defined allocator (the reference has none): the budget is respected, noisy views receive more samples
    than clean ones, the result converges to the fixed-spp image, and every sample is one the fixed-spp render would also draw."""
    desc = scenes.cornell_box(grid_x=4, grid_y=1, view_res=32, sample_count=128, cam_dist=900.0)
    s = scenes.build(desc, device=gpu)
    ref = s.render(seed=0, spp=512)
    img = s.render(seed=1, spp=128, adaptive=True, adaptive_min_spp=16)
    w = s.film_raw()[..., 3]
    per_view = [w[:, 32 * v:32 * (v + 1)].mean() for v in range(4)]
    assert abs(np.mean(per_view) - 128) <= 1.0 and s.last_stats["spp_done"] in (127, 128)
    assert max(per_view) > 1.05 * min(per_view) and min(per_view) >= 16           # allocation differs between views
    for v in range(4):
        assert rel_mse(img[:, 32 * v:32 * (v + 1)], ref[:, 32 * v:32 * (v + 1)]) < 2e-2
    fixed = s.render(seed=1, spp=128)
    assert rel_mse(img, ref) < 1.5 * rel_mse(fixed, ref) + 1e-4                   # no worse than uniform sampling at the same budget
    # threshold: with a huge error threshold nothing is refined beyond the warm-up
    s.render(seed=1, spp=128, adaptive=True, adaptive_min_spp=16, adaptive_threshold=1e9)
    assert s.film_raw()[..., 3].mean() <= 17


def test_adaptive_film_is_the_claimed_samples_and_identical_on_any_world_size(scenes, ob, gpu):
    """(1) The adaptive film equals the CPU oracle's film of exactly the samples the allocator says it handed out (weights exact, radiance within
    the summation-order tolerance).  (2) 2 / 4 / 8 ranks -- run in lockstep on this one GPU, their statistics summed by a host-side hook at the
    point where NCCL sums them on a multi-GPU box -- derive the same allocation pass by pass and their films add up to the one-rank film."""
    import threading
    import torch
    mg = __import__("importlib").import_module("mitsuba3-amvpt_b200.multigpu")
    desc = scenes.cornell_box(grid_x=4, grid_y=2, view_res=16, sample_count=64, cam_dist=900.0)
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    s.render(seed=1, spp=64, adaptive=True, adaptive_min_spp=16, develop=False); film = s.film_raw().copy(); alloc = s.adaptive_alloc()
    assert alloc.shape[1] == 8 and alloc.shape[0] >= 4 and len({tuple(r) for r in alloc}) > 1, "allocation never differentiated the views"
    assert alloc.sum() * 16 * 16 == film[..., 3].sum()
    ofilm = o.render_alloc(alloc, seed=1, spp_per_pass=16, brute=True, nthreads=4)
    assert np.array_equal(film[..., 3], ofilm[..., 3])
    assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= FILM_RTOL * ofilm[..., :3].max()
    with pytest.raises(Exception, match="summed over ranks"):
        s.render(seed=1, spp=64, adaptive=True, rank=0, world=2, develop=False)
    for world in (2, 4, 8):
        ranks = [scenes.build(desc, device=gpu) for _ in range(world)]
        barrier = threading.Barrier(world); shared = [None] * world; errors = []

        def run(r):
            def hook(dev_ptr, n):               # what ncclAllReduce(sum) does between the ranks of a multi-GPU run
                t = torch.as_tensor(mg.DevicePtr(dev_ptr, n), device=f"cuda:{gpu}"); torch.cuda.synchronize()
                shared[r] = t.cpu().numpy().copy(); barrier.wait()
                total = np.sum(np.stack(shared), axis=0, dtype=np.float32); barrier.wait()
                t.copy_(torch.from_numpy(total)); torch.cuda.synchronize()
            try:
                ranks[r].render(seed=1, spp=64, adaptive=True, adaptive_min_spp=16, develop=False, rank=r, world=world, allreduce=hook)
            except Exception as e:       # a failing rank must not leave the others waiting at the barrier
                errors.append(e); barrier.abort()
        th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
        [t.start() for t in th]; [t.join() for t in th]
        assert not errors, errors
        total = np.zeros_like(film)
        for r in range(world):
            assert np.array_equal(ranks[r].adaptive_alloc(), alloc), f"rank {r} of {world} derived another allocation"
            total += ranks[r].film_raw()
        assert np.array_equal(total[..., 3], film[..., 3])
        assert np.abs(total[..., :3] - film[..., :3]).max() <= FILM_RTOL * film[..., :3].max()


# ---------------------------------------------------------------------------------------------------------------------
# Cross-view sample reuse + MIS (src/integrators/mvpath_multi.h).  Parity unpinned in the reference
# (no tests); the checker is the CPU restatement oracle/o_reuse.h.
REUSE_CASES = [
    ("diffuse_mis_fast", None, dict(sa_mis=True, fast_mis=True, reuse_count=4)),
    ("diffuse_mis", None, dict(sa_mis=True, fast_mis=False, reuse_count=2)),
    ("diffuse_nomis", None, dict(sa_mis=False, reuse_count=4)),
    ("rough_mis", "rough", dict(sa_mis=True, fast_mis=False, reuse_count=4)),
    ("rough_mis_fast", "rough", dict(sa_mis=True, fast_mis=True, reuse_count=0)),
    ("glass_mis", "glass", dict(sa_mis=True, fast_mis=False, reuse_count=4)),
]
REUSE_OVERRIDES = {
    None: None,
    "rough": {"white": dict(type="roughplastic", alpha=0.15, reflectance=(.5, .5, .5)),
              "red": dict(type="roughconductor", alpha=0.2, distribution="ggx", eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)),
              "green": dict(type="roughconductor", alpha_u=0.05, alpha_v=0.3, eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14), twosided=True)},
    "glass": {"box": dict(type="roughdielectric", alpha=0.2), "white": dict(type="diffuse", reflectance=(.7, .7, .7), twosided=True),
              "red": dict(type="conductor", eta=(0.2, 0.92, 1.1), k=(3.9, 2.45, 2.14)), "green": dict(type="dielectric")},
}


@pytest.mark.parametrize("name,ov,cfg", REUSE_CASES, ids=[c[0] for c in REUSE_CASES])
def test_sample_reuse_lanes_and_film_match_oracle(name, ov, cfg, scenes, ob, gpu):
    """Every lane's n reused samples (film position, MIS weight, radiance, validity, camera index) are bit-identical to the
    oracle's, and the accumulated film agrees within the summation-order tolerance."""
    desc = scenes.cornell_box(grid_x=4, grid_y=2, view_res=16, cam_dist=150.0, sample_count=16, bsdf_overrides=REUSE_OVERRIDES[ov])
    desc["reuse"] = dict(sa_reuse=True, **cfg)
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    assert s.n_reuse == o.n_reuse and s.n_reuse >= 2
    W, H = o.film_size(); n = W * H * 4
    a = s.trace_lanes_multi(7, 4, 0, n); b = o.trace_lanes_multi(7, 4, 0, n, brute=True)
    assert np.array_equal(a[..., 6:], b[..., 6:]), "validity / indirect flags / camera indices"
    valid = b[..., 6] > 0
    assert valid[:, 1:].any(), "no sample was ever reused: the test would be vacuous"
    for c, what in ((slice(0, 2), "film position"), (slice(2, 3), "weight"), (slice(3, 6), "radiance")):
        x, y = a[..., c][valid], b[..., c][valid]
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), f"{what}: {np.abs(x - y).max()}"
    s.render(seed=3, spp=16, develop=False); film = s.film_raw()
    assert s.last_stats["scheduler_used"] == 5
    ofilm, _, _ = o.render(seed=3, spp=16, brute=True, nthreads=4)
    scale = ofilm[..., :3].max()
    assert np.allclose(film[..., 3], ofilm[..., 3], rtol=1e-5, atol=1e-5)
    assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= 5 * FILM_RTOL * scale
    # the pipeline schedule (default: primary / project / select / suffix / finish kernels per chunk) with small ragged chunks, the
    # three-kernel staged schedule, and the
    # single persistent kernel accumulate the same samples (differences: order of the float atomics only)
    for env in (dict(AMVPT_REUSE_CHUNK="1000"), dict(AMVPT_REUSE_STAGED="1", AMVPT_REUSE_CHUNK="1500"), dict(AMVPT_REUSE_STAGED="0")):
        os.environ.update(env)
        try:
            s.render(seed=3, spp=16, develop=False); other = s.film_raw()
        finally:
            for k in env:
                del os.environ[k]
        assert np.allclose(other[..., 3], film[..., 3], rtol=1e-5, atol=1e-5), env
        assert np.abs(other[..., :3] - film[..., :3]).max() <= 5 * FILM_RTOL * scale, env


@pytest.mark.parametrize("adaptive", [1, 4])
@pytest.mark.parametrize("ov", [None, "rough", "glass"])
def test_sample_reuse_adaptive_fill_in(ov, adaptive, scenes, ob, gpu):
    """The reference's `adaptive` feature (mvpath_multi.h:52-58, 79-115): lanes whose sample only the primary camera collected get their weight
    scaled by 1 / (n_adapt + 1) and n_adapt extra plain paths (compress / repeat order, forked sampler seeded (wavefront, wavefront)).  Lane
    records incl. the mask are bit-exact against the oracle, the film agrees within the summation-order tolerance, weights are conserved, and
    `debug` shows the mask."""
    desc = scenes.cornell_box(grid_x=4, grid_y=2, view_res=16, cam_dist=150.0, sample_count=16, bsdf_overrides=REUSE_OVERRIDES[ov])
    desc["reuse"] = dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4, adaptive=adaptive)
    s = scenes.build(desc, device=gpu); o = ob.OracleScene(desc)
    W, H = o.film_size(); n = W * H * 4
    a = s.trace_lanes_multi(7, 4, 0, n); b = o.trace_lanes_multi(7, 4, 0, n, brute=True)
    assert np.array_equal(a[..., 6:], b[..., 6:]), "validity + adapt mask / indirect flags / camera indices"
    mask = b[:, 0, 6] >= 2
    assert 0.01 < mask.mean() < 0.99, f"adapt mask fraction {mask.mean()}: the test would be vacuous"
    valid = (b[..., 6] % 2) > 0
    for c in (slice(0, 2), slice(2, 3), slice(3, 6)):
        assert np.array_equal(a[..., c][valid].view(np.uint32), b[..., c][valid].view(np.uint32))
    n_adapt = min(adaptive, 3)
    # without the fill-in the masked lanes' primary weight is (n_adapt + 1) times larger
    d0 = dict(desc); d0["reuse"] = dict(desc["reuse"], adaptive=0); s0 = scenes.build(d0, device=gpu); a0 = s0.trace_lanes_multi(7, 4, 0, n)
    assert np.allclose(a[mask, 0, 2] * (n_adapt + 1), a0[mask, 0, 2], rtol=1e-6) and np.array_equal(a[~mask, 0, 2], a0[~mask, 0, 2])
    s.render(seed=3, spp=16, develop=False); film = s.film_raw(); fills = s.last_fill_paths()
    ofilm, _, _ = o.render(seed=3, spp=16, brute=True, nthreads=4)
    assert fills > 0 and fills % n_adapt == 0
    scale = ofilm[..., :3].max()
    assert np.allclose(film[..., 3], ofilm[..., 3], rtol=1e-5, atol=1e-5)
    assert np.abs(film[..., :3] - ofilm[..., :3]).max() <= 5 * FILM_RTOL * scale
    s0.render(seed=3, spp=16, develop=False); film0 = s0.film_raw()
    assert abs(film[..., 3].sum() - film0[..., 3].sum()) < 1e-2 * film0[..., 3].sum()         # (w0 + n_adapt) / (n_adapt + 1) per masked lane: equal when the primary's MIS weight is 1
    # every schedule records the same mask and runs the same fill-in
    for env in (dict(AMVPT_REUSE_CHUNK="1000"), dict(AMVPT_REUSE_STAGED="1", AMVPT_REUSE_CHUNK="1500"), dict(AMVPT_REUSE_STAGED="0")):
        os.environ.update(env)
        try:
            s.render(seed=3, spp=16, develop=False); other = s.film_raw(); assert s.last_fill_paths() == fills
        finally:
            for k in env:
                del os.environ[k]
        assert np.allclose(other[..., 3], film[..., 3], rtol=1e-5, atol=1e-5) and np.abs(other[..., :3] - film[..., :3]).max() <= 5 * FILM_RTOL * scale, env
    # debug: the film holds the mask (weight 1 per primary sample), nothing else
    s.set_reuse_debug(True); s.render(seed=3, spp=4, develop=False); dbg = s.film_raw()
    o2 = ob.OracleScene(dict(desc, reuse=dict(desc["reuse"], debug=True))); odbg, _, _ = o2.render(seed=3, spp=4, brute=True, nthreads=4)
    assert np.array_equal(dbg, odbg) and np.array_equal(dbg[..., 0], dbg[..., 1]) and (dbg[..., 3] >= 3).all() and 0 < dbg[..., 0].sum() < dbg[..., 3].sum()
    with pytest.raises(Exception, match="one rank"):
        s.set_reuse_debug(False); s.render(seed=3, spp=4, develop=False, rank=0, world=2)


def test_sample_reuse_is_consistent_with_plain_path_tracing(scenes, gpu):
    """With MIS the reused estimator converges to the same image as the plain integrator and, at equal primary samples, has a
    lower error against a high-spp reference (the point of the method); sharding by view still sums to the single-GPU film."""
    desc = scenes.cornell_box(grid_x=4, grid_y=1, view_res=32, cam_dist=100.0, sample_count=64)
    s0 = scenes.build(desc, device=gpu); ref = s0.render(seed=11, spp=4096); base = s0.render(seed=1, spp=64)
    d2 = dict(desc); d2["reuse"] = dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4)
    s1 = scenes.build(d2, device=gpu); img = s1.render(seed=1, spp=64); film = s1.film_raw()

    def relmse(x):
        return float(np.mean((x - ref) ** 2 / (ref ** 2 + 1e-2)))
    assert abs(img.mean() - ref.mean()) < 0.01 * ref.mean()
    assert relmse(img) < relmse(base)
    assert abs(film[..., 3].mean() - 64.0) < 0.5           # the MIS weights of a primary sample sum to one on average
    parts = np.zeros_like(film)
    for r in range(2):
        s1.render(seed=1, spp=64, develop=False, rank=r, world=2); parts += s1.film_raw()
    assert np.allclose(parts, film, rtol=1e-4, atol=1e-4 * film[..., :3].max())
    with pytest.raises(Exception, match="adaptive"):
        s1.render(seed=1, spp=64, adaptive=True)


# ------------------------------------------------------------------ output path and display step
LKG4x8 = dict(center=-0.5, focus=-0.05, pitch=354.548, tilt=-0.114, subp=0.00013, view=(1.0, 1.0), grid=(4, 8))   # src/mitsuba/program.h:74-75


def _read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline() == b"PF\n"; w, h = map(int, f.readline().split()); assert f.readline().strip() == b"-1"
        return np.frombuffer(f.read(), "<f4").reshape(h, w, 3)[::-1]


def _read_exr(path):
    """Minimal reader for single-part, uncompressed scan-line OpenEXR files (the layout of the OpenEXR 2 file format document)."""
    import struct
    b = open(path, "rb").read(); assert struct.unpack_from("<ii", b, 0) == (20000630, 2); o = 8; attrs = {}
    def cstr(o):
        e = b.index(b"\0", o); return b[o:e].decode(), e + 1
    while b[o] != 0:
        name, o = cstr(o); typ, o = cstr(o); n, = struct.unpack_from("<i", b, o); o += 4; attrs[name] = (typ, b[o:o + n]); o += n
    o += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"] == ("lineOrder", b"\0")
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1]); w, h = x1 - x0 + 1, y1 - y0 + 1
    assert attrs["displayWindow"] == attrs["dataWindow"] and attrs["pixelAspectRatio"][1] == struct.pack("<f", 1.0)
    ch, c, names, types = attrs["channels"][1], 0, [], []
    while ch[c] != 0:
        e = ch.index(b"\0", c); names.append(ch[c:e].decode()); t, lin, xs, ys = struct.unpack_from("<iB3xii", ch, e + 1); types.append(t); assert (xs, ys) == (1, 1); c = e + 17
    assert names == sorted(names) == ["B", "G", "R"] and len(set(types)) == 1 and types[0] in (1, 2)
    dt = np.dtype("<f2") if types[0] == 1 else np.dtype("<f4"); row = w * 3 * dt.itemsize
    offs = struct.unpack_from(f"<{h}Q", b, o); assert o + 8 * h == offs[0] and len(b) == offs[0] + h * (8 + row)
    img = np.zeros((h, w, 3), dt)
    for y in range(h):
        yy, n = struct.unpack_from("<ii", b, offs[y]); assert (yy, n) == (y, row)
        img[y] = np.frombuffer(b, dt, 3 * w, offs[y] + 8).reshape(3, w)[::-1].T          # file order B, G, R
    return img


def test_exr_output_half_and_float_quilt_and_views(scenes, gpu, tmp_path):
    """HDRFilm::write (file_format = openexr, hdrfilm.cpp:420-560): float32 holds the developed film bit for bit; float16 (the film's
    default component_format) is its round-to-nearest-even conversion; view >= 0 cuts that view's tile."""
    s = scenes.build(scenes.cornell_box(grid_x=3, grid_y=2, view_res=40, sample_count=4), device=gpu)
    img = s.render(seed=2, spp=4)
    s.write_exr(tmp_path / "q32.exr", component_format="float32"); got = _read_exr(tmp_path / "q32.exr")
    assert got.dtype == np.float32; assert_bits(got.reshape(-1, 3), img.reshape(-1, 3), "quilt exr f32")
    s.write_exr(tmp_path / "q16.exr"); got = _read_exr(tmp_path / "q16.exr")
    assert got.dtype == np.float16 and np.array_equal(got.view(np.uint16), img.astype(np.float16).view(np.uint16))
    tiles = s.view_tiles(); assert sorted(int(t[4]) for t in tiles) == list(range(6))
    for x0, y0, w, h, v in tiles:
        s.write_exr(tmp_path / f"v{v}.exr", view=v)
        assert np.array_equal(_read_exr(tmp_path / f"v{v}.exr").view(np.uint16), img[y0:y0 + h, x0:x0 + w].astype(np.float16).view(np.uint16))
    # the conversion at its edges: film_add puts chosen values into the film (weight 1 -> develop returns them unchanged)
    i = s.info(); n = i["film_h"] * i["film_w"] * 3
    edge = np.array([0.0, -0.0, 65504.0, 65519.99, 65520.0, 1e9, np.inf, -np.inf, 2.0 ** -24, 2.0 ** -25, 2.0 ** -25 * 1.0001, 2.0 ** -14, 2.0 ** -14 * 0.9999,
                     6.1e-5, 1.0009765625, 1.00048828125, 1.00146484375, 0.333333, -2.7182817, 1023.75, 2047.5, 5.9e-8, 3e-8, 1e-10], np.float32)
    vals = np.resize(np.concatenate([edge, np.random.default_rng(5).standard_normal(4096).astype(np.float32) * np.float32(100.0),
                                     np.exp(np.random.default_rng(6).uniform(-20, 12, 4096)).astype(np.float32)]), n).reshape(i["film_h"], i["film_w"], 3)
    film = np.concatenate([vals, np.ones_like(vals[..., :1])], axis=-1)
    s.film_clear(); s.film_add(film); dev = s.develop()
    assert np.isinf(dev).any() and (dev == np.float32(65520.0)).any() and (dev == np.float32(2.0 ** -25)).any()     # the edge values reached the film
    s.write_exr(tmp_path / "edge.exr"); got = _read_exr(tmp_path / "edge.exr")
    with np.errstate(over="ignore"):
        want = dev.astype(np.float16)
    assert np.array_equal(got.view(np.uint16), want.view(np.uint16))
    with pytest.raises(Exception, match="out of range"):
        s.write_exr(tmp_path / "x.exr", view=6)
    with pytest.raises(ValueError):
        s.write_exr(tmp_path / "x.exr", component_format="uint32")


def test_pfm_output_quilt_and_per_view_split(scenes, gpu, tmp_path):
    """HDRFilm::write (file_format = pfm): the file holds the developed film bit for bit, bottom row first; view >= 0 cuts that view's tile."""
    s = scenes.build(scenes.cornell_box(grid_x=3, grid_y=2, view_res=40, sample_count=4), device=gpu)
    img = s.render(seed=2, spp=4)
    s.write_pfm(tmp_path / "quilt.pfm"); assert_bits(_read_pfm(tmp_path / "quilt.pfm").reshape(-1, 3), img.reshape(-1, 3), "quilt pfm")
    tiles = s.view_tiles(); assert sorted(int(t[4]) for t in tiles) == list(range(6))
    for x0, y0, w, h, v in tiles:
        s.write_pfm(tmp_path / f"v{v}.pfm", view=v)
        assert_bits(_read_pfm(tmp_path / f"v{v}.pfm").reshape(-1, 3), img[y0:y0 + h, x0:x0 + w].reshape(-1, 3), f"view {v} pfm")
    with pytest.raises(Exception, match="out of range"):
        s.write_pfm(tmp_path / "x.pfm", view=6)


@pytest.mark.parametrize("res", [40, 37])           # 37: film pixel count not a multiple of four (ragged tail of the 4-pixel kernel)
def test_srgb8_quilt_bit_exact(res, scenes, ob, gpu):
    s = scenes.build(scenes.cornell_box(grid_x=3, grid_y=1, view_res=res, sample_count=4), device=gpu)
    s.render(seed=1, spp=4, develop=False)
    # widen the value range beyond what the box produces: dark, bright (clamped), zero-weight pixels
    rng = np.random.default_rng(res); extra = np.zeros_like(s.film_raw()); m = rng.uniform(size=extra.shape[:2]) < 0.3
    extra[m, :3] = np.exp(rng.uniform(-12, 4, (int(m.sum()), 3))).astype(np.float32); s.film_add(extra)
    got = s.to_srgb8(); want = ob.to_srgb8(ob.develop(s.film_raw()).astype(np.float32))
    assert np.array_equal(s.develop().view(np.uint32), ob.develop(s.film_raw()).astype(np.float32).view(np.uint32))
    assert got.shape == want.shape and np.array_equal(got, want), f"{(got != want).sum()} sub-pixels differ"
    assert len(np.unique(got)) > 200


@pytest.mark.parametrize("mode", ["lkg", "lkg_bgr4", "nearest", "quilt", "extern_rgba", "fine_pitch"])
def test_interlace_bit_exact(mode, scenes, ob, gpu):
    """Program::display_image as a kernel against the oracle restatement: every destination byte equal."""
    s = scenes.build(scenes.cornell_box(grid_x=4, grid_y=8, view_res=48, sample_count=4), device=gpu)
    s.render(seed=0, spp=4, develop=False); quilt = s.to_srgb8()
    preset, dst, ch, src = dict(LKG4x8), (301, 173), 3, None
    if mode == "lkg_bgr4": preset["flip_rgb"] = False; ch = 4
    if mode == "nearest": preset["nearest"] = True; preset["focus"] = 0.3
    if mode == "quilt": preset["quilt"] = True; dst = (150, 97)
    if mode == "extern_rgba":
        rng = np.random.default_rng(3); src = rng.integers(0, 256, (211, 163, 4), dtype=np.uint8); preset.update(grid=(5, 9), view=(0.9, 0.8)); ch = 4
    if mode == "fine_pitch": preset.update(pitch=52.3, tilt=0.31, center=0.2, subp=0.0021, focus=0.12); dst = (640, 360)
    got = s.interlace(preset, dst[0], dst[1], ch, src=src)
    want = ob.interlace(quilt if src is None else src, preset, dst[0], dst[1], ch)
    assert np.array_equal(got, want), f"{mode}: {(got != want).sum()} of {got.size} bytes differ"
    assert len(np.unique(got[..., :3])) > 20
    srgb_ms, il_ms = s.display_last_ms(); assert il_ms > 0
    with pytest.raises(Exception):
        s.interlace(preset, dst[0], 1, ch)


def test_cli_default_output_is_half_float_exr(pkg, scenes, gpu, tmp_path):
    """`-o img.exr` (and no -o at all) writes OpenEXR with float16 components as the reference's hdrfilm does by default."""
    import subprocess, sys
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=24, sample_count=2)
    xml = scenes.export_xml(desc, str(tmp_path))
    r = subprocess.run([sys.executable, "-m", "mitsuba3-amvpt_b200", str(xml), "-n", "2", "--views"], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    out = os.path.splitext(str(xml))[0] + ".exr"; assert out in r.stdout
    s = scenes.build(desc, device=gpu); img = s.render(seed=0, spp=2)
    assert np.array_equal(_read_exr(out).view(np.uint16), img.astype(np.float16).view(np.uint16))
    tiles = s.view_tiles(); assert sorted(int(t[4]) for t in tiles) == list(range(2))
    for x0, y0, w, h, v in tiles:
        assert np.array_equal(_read_exr(os.path.splitext(out)[0] + f"_v{v}.exr").view(np.uint16), img[y0:y0 + h, x0:x0 + w].astype(np.float16).view(np.uint16))


def test_cli_renders_writes_pfm_views_and_interlace(pkg, scenes, gpu, tmp_path):
    """`python -m mitsuba3-amvpt_b200 scene.xml -n 4 -o out.pfm --views --interlace W H` = render_internal of the reference's CLI
    (src/mitsuba/mitsuba_render.cpp:105-133): render with seed 0, then film->write -- same bits as the API calls."""
    import subprocess, sys
    desc = scenes.cornell_box(grid_x=2, grid_y=2, view_res=32, sample_count=4)
    xml = scenes.export_xml(desc, str(tmp_path))
    out = tmp_path / "img.pfm"
    r = subprocess.run([sys.executable, "-m", "mitsuba3-amvpt_b200", str(xml), "-n", "4", "-o", str(out), "--views", "--interlace", "96", "64"],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Rendering finished." in r.stdout and "img.pfm" in r.stdout
    s = scenes.build(desc, device=gpu); img = s.render(seed=0, spp=4)
    assert_bits(_read_pfm(tmp_path / "img.pfm").reshape(-1, 3), img.reshape(-1, 3), "CLI quilt")
    tiles = s.view_tiles(); assert sorted(int(t[4]) for t in tiles) == list(range(4))
    for x0, y0, w, h, v in tiles:
        assert_bits(_read_pfm(tmp_path / f"img_v{v}.pfm").reshape(-1, 3), img[y0:y0 + h, x0:x0 + w].reshape(-1, 3), f"CLI view {v}")
    s.to_srgb8(fetch=False); want = s.interlace(dict(LKG4x8, grid=(2, 2)), 96, 64, 3)
    ppm = open(tmp_path / "img_interlaced.ppm", "rb").read()
    assert ppm.startswith(b"P6\n96 64\n255\n") and np.array_equal(np.frombuffer(ppm[len(b"P6\n96 64\n255\n"):], np.uint8).reshape(64, 96, 3), want)
