#!/usr/bin/env python
"""bench.py -- Mpaths/s of the AMVPT render loop on BASELINE.json's configuration 2 (Cornell box, 8x8 light-field
quilt of 512x512 views, 256 spp fixed), one JSON line on stdout (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cbox8x8|cbox4|hfield16]

A step = one MVPathIntegrator::render() of the whole quilt (16 passes of 16 spp).  `value` is timed on the device
(CUDA events, max over ranks) with the scene and film resident in HBM; `e2e` is the same job through the C ABI with
host buffers: scene upload (commit) + render + develop + device->host copy of the developed image every step.
N > 1: views are sharded over ranks (strong scaling, total work fixed), the owned tiles are gathered to rank 0 with NCCL.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    # name: (scene factory name, kwargs, spp)
    "cbox8x8": ("cornell_box", dict(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, max_depth=6, sample_count=256), 256),
    "cbox4": ("cornell_box", dict(grid_x=4, grid_y=1, view_res=256, cam_dist=200.0, max_depth=6, sample_count=64), 64),
    "hfield4": ("height_field", dict(n=708, grid_x=2, grid_y=2, view_res=512, max_depth=8, sample_count=16), 16),      # profiling size of hfield16
    "hfield16": ("height_field", dict(n=708, grid_x=4, grid_y=4, view_res=1024, max_depth=8, sample_count=64), 64),
    # BASELINE config 5 geometry (64 views of 1024^2 on the 1M-triangle scene); its 1024-spp adaptive budget is reached with --adaptive --spp 1024
    "hfield64": ("height_field", dict(n=708, grid_x=8, grid_y=8, view_res=1024, max_depth=8, sample_count=64), 64),
    # the same light field with cross-view sample reuse + MIS, 8 cameras per group (one row of the 8x8 array)
    "cbox8x8_reuse": ("cornell_box", dict(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, max_depth=6, sample_count=64), 64),
    "cbox4_reuse": ("cornell_box", dict(grid_x=4, grid_y=1, view_res=256, cam_dist=200.0, max_depth=6, sample_count=64), 64),   # profiling size
    "hfield16_reuse": ("height_field", dict(n=708, grid_x=4, grid_y=4, view_res=1024, max_depth=8, sample_count=16), 16),
}
EXTRA = {"cbox8x8_reuse": dict(reuse=dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=8)),
         "cbox4_reuse": dict(reuse=dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4)),
         "hfield16_reuse": dict(reuse=dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4))}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def cpu_reference_run(desc, spp_per_pass, target_seconds, threads):
    """Times the CPU oracle (the restatement of the reference's own algorithm, oracle/amvpt_oracle.cpp) on a bounded
    sample of the workload: every k-th row of the quilt (so that all views and image regions are represented) at one
    pass of spp_per_pass samples.  Returns Mpaths/s."""
    import oracle_binding as ob
    o = ob.OracleScene(desc)
    w, h = o.film_size()

    def odd(n):           # odd strides are coprime with the power-of-two tile heights, so every image region is sampled
        n = max(1, int(n)); return n if n % 2 or n == 1 else n + 1
    stride = odd(h / 48)
    _, sec, _ = o.render(seed=0, spp=spp_per_pass, brute=False, nthreads=threads, row_stride=stride)
    rate = len(range(0, h, stride)) * w * spp_per_pass / max(sec, 1e-9)
    rows = max(1.0, min(float(h), rate * target_seconds / (w * spp_per_pass)))
    stride = odd(h / rows) if rows < h else 1
    _, sec, _ = o.render(seed=0, spp=spp_per_pass, brute=False, nthreads=threads, row_stride=stride)
    rows = len(range(0, h, stride)); paths = rows * w * spp_per_pass
    return paths / sec / 1e6, (f"every {stride}th row of the {w}x{h} quilt ({rows} rows) x {spp_per_pass} spp = {paths} paths in {sec:.2f} s on {threads} threads "
                               "(CPU restatement of the reference render loop with its own BVH; closest-hit results identical to the GPU path)")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1); ap.add_argument("--steps", type=int, default=3); ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200"); ap.add_argument("--workload", default="cbox8x8")
    ap.add_argument("--scheduler", default="auto"); ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--adaptive", action="store_true", help="BASELINE configs 3 / 5: allocate spp per view from the variance statistics (all-reduced across ranks)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel (mean budget with --adaptive)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    scenes = None
    factory, kw, spp = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp; kw = dict(kw, sample_count=spp)
    threads = os.cpu_count() or 1

    if args.impl == "reference":
        # the reference's own CPU algorithm for the path (oracle port; llvm_rgb / scalar_rgb cannot be built here)
        if rank != 0:
            return
        spec = importlib.util.spec_from_file_location("amvpt_scenes", os.path.join(ROOT, "mitsuba3-amvpt_b200", "scenes.py"))
        scenes = importlib.util.module_from_spec(spec); spec.loader.exec_module(scenes)
        desc = getattr(scenes, factory)(**kw); desc.update(EXTRA.get(args.workload, {}))
        per_step = max(2.0, min(20.0, 120.0 / max(args.steps + args.warmup, 1)))
        vals = []
        for i in range(args.warmup + args.steps):
            v, sample = cpu_reference_run(desc, 16, per_step, threads)
            if i >= args.warmup:
                vals.append(v)
        v = float(np.mean(vals))
        print(json.dumps({"impl": "reference", "metric": "Mpaths/s (all views)", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                          "dtype": "f32", "data": "synthetic", "config": {"workload": args.workload, **{k: v_ for k, v_ in kw.items()}},
                          "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": threads, "kind": "port", "sample": sample},
                          "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    pkg = importlib.import_module("mitsuba3-amvpt_b200"); scenes = importlib.import_module("mitsuba3-amvpt_b200.scenes")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    desc = getattr(scenes, factory)(**kw); desc.update(EXTRA.get(args.workload, {}))
    s = scenes.build(desc, device=local)
    info = s.info(); W, H = info["film_w"], info["film_h"]
    n_paths = W * H * spp
    film_ptr, _ = s.film_device_ptr()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def gather_film():
        """NCCL film gather: ranks own disjoint pixels (box filter), so a sum-reduce to rank 0 equals the gather of tiles."""
        if dist is None:
            return
        t = film_tensor()
        dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)

    class _Ptr:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}

    def film_tensor():
        return torch.as_tensor(_Ptr(film_ptr, W * H * 4), device=f"cuda:{local}")

    hook = None
    if args.adaptive and world > 1:
        hook = importlib.import_module("mitsuba3-amvpt_b200.multigpu").make_allreduce_hook(f"cuda:{local}")

    def step():
        s.render(seed=0, spp=spp, develop=False, scheduler=args.scheduler, rank=rank, world=world, adaptive=args.adaptive, allreduce=hook)
        gather_film()
        return s.last_stats

    for _ in range(args.warmup):
        step()
    clocks = ClockSampler(local); barrier(); clocks.start()
    dev_s, launches, render_s, traced = 0.0, 0, 0.0, 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); st = step(); e1.record(); e1.synchronize()
        dev_s += e0.elapsed_time(e1) * 1e-3; launches += st["launches"]
        render_s += st["seconds"]; traced += st["paths"]
    barrier(); wall = time.perf_counter() - t0
    clk = clocks.stop()
    tmax = torch.tensor([dev_s], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    # per-rank breakdown (diagnostic): device seconds inside render() vs. the whole step (render + NCCL film reduce)
    mine = torch.tensor([render_s, dev_s], dtype=torch.float64, device=f"cuda:{local}")
    per_rank = [torch.zeros_like(mine) for _ in range(world)]
    if dist is not None:
        dist.all_gather(per_rank, mine)
    else:
        per_rank = [mine]
    rank_ms = [[round(float(x) / args.steps * 1e3, 2) for x in t.tolist()] for t in per_rank]
    dev_s = float(tmax.item())
    tr = torch.tensor([float(traced)], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(tr, op=dist.ReduceOp.SUM)
    if args.adaptive:                       # the allocator spends the budget in whole passes: count the paths actually traced
        n_paths = int(tr.item()) // max(args.steps, 1)
    value = n_paths * args.steps / dev_s / 1e6

    # ---- end to end through the C ABI with host buffers (scene upload + render + develop + D2H every step)
    def e2e_step():
        sc = scenes.build(desc, device=local)          # host description -> BVH build -> H2D
        img = sc.render(seed=0, spp=spp, develop=(world == 1), scheduler=args.scheduler, rank=rank, world=world, adaptive=args.adaptive, allreduce=hook)
        if world > 1:
            fp, _ = sc.film_device_ptr(); t = torch.as_tensor(_Ptr(fp, W * H * 4), device=f"cuda:{local}")
            dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                img = sc.develop()
        return img
    e2e_step(); barrier(); t0 = time.perf_counter()
    for _ in range(max(1, args.steps)):
        img = e2e_step()
    barrier(); e2e_s = time.perf_counter() - t0
    tm = torch.tensor([e2e_s], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    e2e_val = n_paths * max(1, args.steps) / float(tm.item()) / 1e6
    h2d = sum(m["positions"].nbytes + m["faces"].nbytes for m in desc["meshes"]) + info["n_bvh_nodes"] * 64 + info["n_prims"] * 48 + info["n_views"] * 128
    d2h = W * H * 3 * 4

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel: algorithmic bytes from the kernels' own counters (one 16-spp pass with counters)
    s.render(seed=0, spp=16, develop=False, scheduler=args.scheduler, stats=True)
    cs = s.last_stats
    rays = cs["closest_rays"] + cs["shadow_rays"]
    bvh_bytes_per_path = (cs["bvh_nodes"] * 64.0 + cs["tri_tests"] * 48.0) / cs["paths"]
    s.render(seed=0, spp=spp, develop=False, scheduler=args.scheduler, time_stages=True); ks = s.last_stats
    peak, peak_src = peaks()
    sched = {0: "warp_batch", 1: "regenerate", 2: "stream", 3: "wavefront", 5: "reuse", 6: "split"}[ks["scheduler_used"]]
    if sched == "wavefront":
        # dominant kernel: wf_trace_kernel.  Per ray: BVH nodes (64 B) + triangles (48 B) + ray read (32 B) + hit / flag traffic (24 B)
        kernel = "wf_trace_kernel"; b_path = bvh_bytes_per_path + 56.0 * rays / cs["paths"]
        per_launch_s = ks["trace_seconds"] / max(ks["trace_launches"], 1); paths_per_launch = ks["paths"] / max(ks["trace_launches"], 1)
        note = "staged wavefront: trace stage = %.0f%% of the device time; the BVH (~130 MB at 1M triangles) straddles L2, latency-bound traversal" % (100 * ks["trace_seconds"] / ks["seconds"])
    elif sched == "reuse":
        # one primary path = prefix (1 closest ray, 1 light ray, n - 1 camera visibility rays) + the shared suffix + up to n splats of 16 B
        film_w = float(torch.as_tensor(_Ptr(film_ptr, W * H * 4), device=f"cuda:{local}").view(-1, 4)[:, 3].sum().item())
        kernel = "render_reuse_kernel"; b_path = 16.0 * max(1.0, film_w / max(ks["paths"], 1) * s.n_reuse / 2) + bvh_bytes_per_path
        per_launch_s = ks["seconds"] / ks["launches"]; paths_per_launch = ks["paths"] / ks["launches"]
        note = "sample reuse: Mpaths/s counts PRIMARY paths; every primary path also feeds up to n - 1 other views (config.reuse_samples)"
    else:
        kernel = "render_pass_kernel"; b_path = 16.0 + bvh_bytes_per_path
        per_launch_s = ks["seconds"] / ks["launches"]; paths_per_launch = ks["paths"] / ks["launches"]
        note = "scene is cache-resident for the Cornell box: the megakernel is issue-bound (SIMD efficiency / issue slots), see profiles/"
    achieved = paths_per_launch * b_path / per_launch_s / 1e9
    # DRAM bytes per launch of that kernel from the committed ncu --set full capture, scaled from the profiled launch size
    traffic, traffic_src = None, None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))[kernel]
        units = ks["paths"] / ks["launches"] if kernel == "render_pass_kernel" else float(int(os.environ.get("AMVPT_WF_POOL", 1 << 22)))
        traffic = tr["dram_bytes_per_launch"] * units / tr.get("paths_per_launch", tr.get("slots_per_launch"))
        traffic_src = f'{tr["source"]} ({tr["workload"]}), scaled by launch size'
    except Exception:
        pass
    out = {"metric": "Mpaths/s (all views, device-timed)", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_s / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": args.workload, **kw, **({"reuse": EXTRA[args.workload]["reuse"], "reuse_samples": s.n_reuse} if args.workload in EXTRA else {}), "spp": spp, "adaptive": bool(args.adaptive), "scheduler": args.scheduler, "scheduler_used": sched, "quilt": [W, H], "paths_per_step": n_paths,
                      "l2_policy": f"film {W * H * 16 / 1e6:.0f} MB > 126 MB L2 rewritten every pass" if W * H * 16 > 126e6 else "film smaller than L2 (small workload)",
                      "parallelism": f"views sharded over {world} rank(s), NCCL film reduce to rank 0" if world > 1 else "single GPU"},
           "clocks": clk, "gpu_launches": launches, "wall_s": wall, "rank_ms_render_step": rank_ms,
           "e2e": {"value": e2e_val, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": paths_per_launch * b_path,
                        "peak_source": peak_src, "kernel": kernel, "bytes_per_path": b_path, "rays_per_path": rays / cs["paths"],
                        "nodes_per_ray": cs["bvh_nodes"] / max(rays, 1), "tris_per_ray": cs["tri_tests"] / max(rays, 1),
                        "launch_ms": per_launch_s * 1e3, "note": note}}
    if world == 1 and not args.no_cpu_baseline:
        v, sample = cpu_reference_run(desc, 16, 15.0, threads)
        out["cpu_baseline"] = {"value": v, "unit": "Mpaths/s", "cores": threads, "kind": "port", "sample": sample}
    print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
