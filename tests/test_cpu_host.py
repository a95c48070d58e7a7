"""CPU-only tests of the product's host side: C-ABI surface, scene construction logic, error behaviour, view sharding
(with a world_size-2 gloo run of the film reduction used by bench.py).  No compute entry point is called: without a GPU
they must fail loudly, which is itself tested."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(pkg):
    hdr = open(os.path.join(ROOT, "include", "amvpt.h")).read()
    declared = sorted(set(re.findall(r"\b(amvpt_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 29
    for name in declared:
        assert hasattr(pkg.lib, name), f"{name} is declared in include/amvpt.h but not exported by libamvpt_b200.so"
    assert set(pkg.EXPORTS) == set(declared)


def test_sensor_matrices_match_oracle(pkg, scenes, ob):
    for kw in (dict(grid_x=4, grid_y=1, view_res=64), dict(grid_x=8, grid_y=8, view_res=32, cam_dist=400.0),
               dict(grid_x=3, grid_y=2, view_res=16, thinlens=True, aperture_radius=5.0, cone_deg=20.0)):
        desc = scenes.cornell_box(**kw)
        s = scenes.build(desc, commit=False); o = ob.OracleScene(desc)
        info = s.info(); assert (info["film_w"], info["film_h"]) == o.film_size()
        for i in range(info["n_views"]):
            a, b = s.sensor(i); c, d = o.sensor(i)
            assert np.array_equal(a.view(np.uint32), c.view(np.uint32)), f"sample_to_camera of view {i}"
            assert np.array_equal(b[:3].view(np.uint32), d[:3].view(np.uint32)), f"to_world of view {i}"
    # the camera line: sub-cameras sit on the x axis, cam_dist apart end to end, centred (grid.cpp:208-226)
    desc = scenes.cornell_box(grid_x=4, grid_y=1, view_res=8, cam_dist=200.0); s = scenes.build(desc, commit=False)
    xs = [s.sensor(i)[1][0, 3] for i in range(4)]
    assert np.allclose(np.diff(xs), -200.0 / 3, atol=1e-3) and np.isclose(np.mean(xs), 278.0, atol=1e-3)


def test_roughplastic_tables_match_oracle(pkg, scenes, ob):
    desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=8)
    desc["bsdfs"] = [dict(type="roughplastic", distribution="beckmann", alpha=0.1, reflectance=(0.5, 0.3, 0.1)),
                     dict(type="roughplastic", distribution="ggx", alpha=0.25, reflectance=(0.2, 0.2, 0.7), int_ior=1.33, ext_ior=1.0, nonlinear=True)]
    for m in desc["meshes"]:
        m["bsdf"] = 0
    s = scenes.build(desc, commit=False); o = ob.OracleScene(desc)
    for i in range(2):
        ta, sa = s.bsdf_tables(i); tb, sb = o.bsdf_tables(i)
        assert np.array_equal(ta.view(np.uint32), tb.view(np.uint32)) and np.array_equal(sa.view(np.uint32), sb.view(np.uint32))
        assert 0 < ta.min() and ta.max() < 1 and np.all(np.diff(ta) > -1e-3)      # transmittance grows with cos(theta)
    # the tables are computed on several threads and then kept per (distribution, alpha, eta): a second scene gets the cached copy -- same bits
    s2 = scenes.build(desc, commit=False)
    for i in range(2):
        (ta, sa), (tb, sb) = s.bsdf_tables(i), s2.bsdf_tables(i)
        assert np.array_equal(ta.view(np.uint32), tb.view(np.uint32)) and np.array_equal(sa.view(np.uint32), sb.view(np.uint32))


def test_bvh_structure(pkg, scenes):
    for desc in (scenes.cornell_box(grid_x=2, grid_y=1, view_res=8), scenes.height_field(n=64, grid_x=2, grid_y=2, view_res=8)):
        s = scenes.build(desc, commit=False)
        depth, leaves = s.bvh_validate()       # throws if a triangle leaves its leaf box, a child leaves its parent, or a triangle is lost
        n = s.info()["n_prims"]
        assert leaves * 4 >= n and depth <= 62 and depth >= int(np.log2(max(n / 4, 1)) / 2)      # nodes are 4-wide (or 2-wide) 
    # degenerate inputs: one triangle, coincident triangles, zero-area triangle
    s = pkg.Scene(); b = s.add_bsdf(type="diffuse")
    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    s.add_mesh(tri, [[0, 1, 2]], b)
    s.set_sensor_grid(2, 1, 16, 8, np.eye(4), 45.0); assert s.bvh_validate()[1] >= 1
    for _ in range(9):
        s.add_mesh(tri, [[0, 1, 2]], b)
    s.add_mesh(np.zeros((3, 3), np.float32), [[0, 1, 2]], b)
    assert s.bvh_validate()[1] >= 3 and s.info()["n_prims"] == 11


def test_bvh_build_does_not_depend_on_the_thread_count(pkg, scenes, monkeypatch):
    """The host build runs on several threads above 65 536 primitives (top of the tree sequentially, disjoint subtrees in parallel, emitted in
    the pre-order of the plain recursion): node numbering, leaf order and the compressed tree must be identical whatever the thread count."""
    sums = []
    for threads in ("1", "3", "8"):
        monkeypatch.setenv("AMVPT_BVH_THREADS", threads)
        s = scenes.build(scenes.height_field(n=200, grid_x=2, grid_y=2, view_res=8), commit=False)       # 80 000 triangles
        assert s.info()["n_prims"] >= 65536
        sums.append(s.bvh_checksum()); s.bvh_validate()
    assert sums[0] == sums[1] == sums[2], sums


def test_error_behaviour(pkg, scenes):
    s = pkg.Scene(); b = s.add_bsdf(type="diffuse")
    with pytest.raises(pkg.AmvptError, match="out of range"):
        s.add_mesh(np.zeros((3, 3), np.float32), [[0, 1, 3]], b)
    with pytest.raises(pkg.AmvptError, match="BSDF"):
        s.add_mesh(np.zeros((3, 3), np.float32), [[0, 1, 2]], 7)
    with pytest.raises(pkg.AmvptError, match="divisible"):          # grid.cpp:100-101
        s.set_sensor_grid(3, 1, 64, 64, np.eye(4), 45.0)
    with pytest.raises(pkg.AmvptError, match="max_depth"):          # integrator.cpp:510-512
        s.set_integrator(max_depth=-2)
    with pytest.raises(pkg.AmvptError, match="rr_depth"):           # integrator.cpp:517-518
        s.set_integrator(rr_depth=0)
    with pytest.raises(pkg.AmvptError, match="not committed"):
        s.render(spp=16)
    with pytest.raises(KeyError):
        s.add_bsdf(type="plastic")                                   # not on the path: rejected, not approximated


def test_compute_fails_loudly_without_gpu(pkg, scenes):
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    s = scenes.build(scenes.cornell_box(grid_x=2, grid_y=1, view_res=8), commit=False)
    with pytest.raises(pkg.AmvptError, match="no CUDA device"):
        s.commit(0)
    with pytest.raises(pkg.AmvptError, match="no CUDA device"):
        pkg.math1(0, np.zeros(4, np.float32))


def test_shard_tiles_partition_the_quilt(pkg, scenes):
    for gx, gy, res in ((8, 8, 16), (4, 1, 32), (3, 2, 8)):
        s = scenes.build(scenes.cornell_box(grid_x=gx, grid_y=gy, view_res=res), commit=False)
        for world in (1, 2, 3, 4, 8, 16):
            cover = np.zeros((gy * res, gx * res), np.int32); px = []
            for rank in range(world):
                t = s.shard_tiles(rank, world); n = 0
                for x0, y0, w, h, view in t:
                    cover[y0:y0 + h, x0:x0 + w] += 1; n += w * h
                    # the tile of view v sits where GridSensor::sample_ray routes it (reverse_y = true, grid.cpp:246-255)
                    if world > 1:
                        assert view == (x0 // res) + gx * (gy - 1 - y0 // res)
                px.append(n)
            assert np.all(cover == 1), f"world {world}: tiles must cover every pixel exactly once"
            assert max(px) - min(px) <= (gx * gy * res * res) // world // 2 + res * res, f"world {world}: unbalanced {px}"


GLOO_WORKER = r"""
import os, sys, importlib, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
mg = importlib.import_module('mitsuba3-amvpt_b200.multigpu')
dist.init_process_group('gloo', rank=int(os.environ['RANK']), world_size=int(os.environ['WORLD_SIZE']))
rank, world = dist.get_rank(), dist.get_world_size()
# (1) the rendezvous of the NCCL unique id: rank 0 makes it, everybody ends up with the same 128 bytes
uid = mg.exchange_unique_id(lambda: bytes(range(128)), rank)
assert uid == bytes(range(128))
# (2) the film hand-off with the library's own rectangles (amvpt_gather_tiles: owned view tiles grown by one pixel), moved here with
# send / recv the way comm.cpp moves them with ncclSend / ncclRecv: pack -> send -> add into the root's film
s = scenes.build(scenes.cornell_box(grid_x=4, grid_y=2, view_res=8), commit=False)
H, W = 16, 32
film = torch.zeros(H, W, 4)
for x0, y0, w, h, view in s.shard_tiles(rank, world):
    film[y0:y0 + h, x0:x0 + w] = float(view + 1)
    if x0 + w < W: film[y0, x0 + w, 3] += 0.25           # a jittered sample that rounded into the neighbouring view's pixel
full = film.clone(); dist.all_reduce(full)                # what a whole-quilt sum would give
def pack(f, tiles): return torch.cat([f[y0:y0 + h, x0:x0 + w].reshape(-1, 4) for x0, y0, w, h, off in tiles])
if rank != 0:
    t = s.gather_tiles(rank, world); buf = pack(film, t); assert int(t[-1][4] + t[-1][2] * t[-1][3]) == len(buf)
    dist.send(buf.contiguous(), dst=0)
else:
    for r in range(1, world):
        t = s.gather_tiles(r, world); n = int(t[-1][4] + t[-1][2] * t[-1][3]); buf = torch.zeros(n, 4); dist.recv(buf, src=r)
        for x0, y0, w, h, off in t:
            film[y0:y0 + h, x0:x0 + w] += buf[int(off):int(off) + int(w) * int(h)].reshape(int(h), int(w), 4)
    assert torch.equal(film, full), 'tile gather must equal the whole-quilt sum, strays included'
    sent = sum(int(t[2] * t[3]) for r in range(1, world) for t in s.gather_tiles(r, world)); assert sent < H * W
dist.destroy_process_group()
sys.stdout.write('ok %d\n' % rank); sys.stdout.flush()
"""


def test_film_reduce_world_size_2_gloo(pkg, tmp_path):
    script = tmp_path / "worker.py"; script.write_text(GLOO_WORKER.format(root=ROOT))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29591", str(script)], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2 and "0" in r.stdout and "1" in r.stdout


def test_cli_arguments_mirror_the_reference_executable(pkg):
    """python -m mitsuba3-amvpt_b200: the options of src/mitsuba/mitsuba_render.cpp:165-184 that reach this path (-D -s -n -o -m -t)."""
    import importlib
    cli = importlib.import_module("mitsuba3-amvpt_b200.__main__")
    a = cli.parse_args(["a.xml", "b.xml", "-D", "spp=64", "--define", "res=256", "-n", "32", "-s", "0", "-m", "cuda_rgb", "-t", "8", "-o", "x.exr"])
    assert a.scenes == ["a.xml", "b.xml"] and a.define == {"spp": "64", "res": "256"} and a.nspp == 32 and a.sensor == 0
    assert cli.output_path("dir/scene.xml", None, 1) == ("dir/scene.exr", False)        # hdrfilm's default file_format is openexr
    assert cli.output_path("scene.xml", "out/img.exr", 1) == ("out/img.exr", False)
    assert cli.output_path("scene.xml", "out/img.pfm", 1) == ("out/img.pfm", False)
    assert cli.output_path("scene.xml", "out/img.png", 1) == ("out/img.exr", True)      # film->write replaces a foreign extension
    assert cli.output_path("a.xml", "ignored.pfm", 2) == ("a.exr", False)               # several scenes: one output per scene name
    for bad in (["a.xml", "-D", "novalue"], ["a.xml", "-m", "scalar_spectral"], []):
        with pytest.raises(SystemExit):
            cli.parse_args(bad)


def test_exr_half_conversion_is_round_to_nearest_even(pkg):
    """amvpt_f32_to_f16 (the component conversion of amvpt_film_write_exr) against numpy's IEEE conversion: every float16 value, the
    midpoints between neighbours and their float32 neighbours, the overflow / underflow thresholds, infinities, NaN stays NaN."""
    h = np.arange(0x10000, dtype=np.uint32).astype(np.uint16); f = h.view(np.float16).astype(np.float32)
    fin = np.isfinite(f); assert np.array_equal(pkg.f32_to_f16(f)[fin | np.isinf(f)], h[fin | np.isinf(f)])
    nan = pkg.f32_to_f16(f[np.isnan(f)]).view(np.float16); assert np.isnan(nan).all()
    pos = np.sort(f[fin & (f >= 0)]); mid = ((pos[:-1].astype(np.float64) + pos[1:].astype(np.float64)) * 0.5).astype(np.float32)   # exact in float32
    cases = np.concatenate([mid, np.nextafter(mid, np.float32(np.inf)), np.nextafter(mid, np.float32(-np.inf)),
                            np.array([65519.996, 65520.0, 65536.0, 3.4e38, 2.0 ** -25, 2.0 ** -26, 1e-45, 0.0], np.float32),
                            np.random.default_rng(3).standard_normal(200000).astype(np.float32) * np.float32(300.0),
                            np.exp(np.random.default_rng(4).uniform(-25.0, 13.0, 200000)).astype(np.float32)])
    cases = np.concatenate([cases, -cases])
    with np.errstate(over="ignore"):
        want = cases.astype(np.float16).view(np.uint16)
    assert np.array_equal(pkg.f32_to_f16(cases), want)
    assert pkg.f32_to_f16(np.zeros(0, np.float32)).size == 0


def test_view_tiles_are_the_per_view_split_of_the_quilt(pkg, scenes):
    """amvpt_view_tiles: one rectangle per sub-sensor, view indices 0..n-1 in order, disjoint cover of the quilt, and the same
    rectangles the view sharding hands out when there are as many ranks as views."""
    for gx, gy, res in ((2, 2, 32), (3, 2, 40), (1, 1, 16), (8, 1, 8)):
        s = scenes.build(scenes.cornell_box(grid_x=gx, grid_y=gy, view_res=res), commit=False)
        t = s.view_tiles(); n = gx * gy
        assert t.shape == (n, 5) and t[:, 4].tolist() == list(range(n))
        cover = np.zeros((gy * res, gx * res), np.int32)
        for x0, y0, w, h, v in t:
            assert (w, h) == (res, res); cover[y0:y0 + h, x0:x0 + w] += 1
        assert (cover == 1).all()
        if n > 1:
            sharded = sorted(tuple(int(x) for x in r) for rank in range(n) for r in s.shard_tiles(rank, n))
            assert sharded == sorted(tuple(int(x) for x in r) for r in t)


def test_integration_shim_compiles_against_the_reference_headers():
    """integration/mvpath_b200.cpp -- the plugin a maintainer adds to the reference tree -- passes `g++ -fsyntax-only` against the reference's
    own headers (scalar_rgb config generated by oracle/gen_ref_config.py) and include/amvpt.h.  Skipped where /root/reference is absent."""
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "include", "mitsuba")):
        pytest.skip("reference tree not mounted")
    cfg = os.path.join(ROOT, "oracle", "_ref", "include", "mitsuba", "core", "config.h")
    if not os.path.exists(cfg):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/include/mitsuba/core/config.h"], check=True, capture_output=True)
    inc = [os.path.join(ROOT, "oracle", "_ref", "include"), os.path.join(ROOT, "include"), f"{ref}/include", f"{ref}/ext/drjit/include",
           f"{ref}/ext/drjit/ext/drjit-core/include", f"{ref}/ext/nanobind/include", f"{ref}/ext/tinyformat",
           f"{ref}/ext/drjit/ext/drjit-core/ext/nanothread/include", f"{ref}/ext/fastfloat/include"]
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-include", "cmath"] + [x for i in inc for x in ("-I", i)] + [os.path.join(ROOT, "integration", "mvpath_b200.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
