#!/usr/bin/env python
"""bench.py -- Mpaths/s of the AMVPT render loop on BASELINE.json's configuration 2 (Cornell box, 8x8 light-field
quilt of 512x512 views, 256 spp fixed), one JSON line on stdout (contract: see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cbox8x8|cbox4|hfield16|hfield64] [--adaptive] [--spp S]

A step = one MVPathIntegrator::render() of the whole quilt (16 passes of 16 spp).  `value` is timed on the device
(CUDA events, max over ranks) with the scene and film resident in HBM; `e2e` is the same job through the C ABI with
host buffers: scene upload (commit) + render + develop + device->host copy of the developed image every step.
N > 1: views are sharded over ranks (strong scaling, total work fixed), the owned tiles are gathered to rank 0 with NCCL.

BASELINE configs: 1 = --workload cbox4, 2 = default, 3 = --adaptive [--gpus N], 4 = --workload hfield16,
5 = --workload hfield64 --adaptive --spp 1024 --gpus 8.  The default single-GPU run also carries short lines for
configs 1 and 4 under "other_workloads" so that the driver's record holds them.
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

METRIC = "Mpaths/s (all views, device-timed)"     # both arms print this string (the driver divides only equal metrics)

WORKLOADS = {
    # name: (scene factory name, kwargs, spp).  cornell_box_ref = geometry / materials of the reference's resources/data/scenes/cbox/cbox-rgb.xml
    #  read through the product's XML loader, camera replaced by the `grid` multi-sensor; cbox8x8_synth keeps round 1's
    # re-derived quads (32 triangles, no vertex normals) for comparison with round-1 numbers
    "cbox8x8": ("cornell_box_ref", dict(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, max_depth=6, sample_count=256), 256),
    "cbox4": ("cornell_box_ref", dict(grid_x=4, grid_y=1, view_res=256, cam_dist=200.0, max_depth=6, sample_count=64), 64),
    "cbox8x8_synth": ("cornell_box", dict(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, max_depth=6, sample_count=256), 256),
    "hfield4": ("height_field", dict(n=708, grid_x=2, grid_y=2, view_res=512, max_depth=8, sample_count=16), 16),      # profiling size of hfield16
    "hfield16": ("height_field", dict(n=708, grid_x=4, grid_y=4, view_res=1024, max_depth=8, sample_count=64), 64),
    # BASELINE config 5 geometry (64 views of 1024^2 on the 1M-triangle scene); its 1024-spp adaptive budget is reached with --adaptive --spp 1024
    "hfield64": ("height_field", dict(n=708, grid_x=8, grid_y=8, view_res=1024, max_depth=8, sample_count=64), 64),
    # the same light field with cross-view sample reuse + MIS, 8 cameras per group (one row of the 8x8 array)
    "cbox8x8_reuse": ("cornell_box_ref", dict(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, max_depth=6, sample_count=64), 64),
    "cbox4_reuse": ("cornell_box_ref", dict(grid_x=4, grid_y=1, view_res=256, cam_dist=200.0, max_depth=6, sample_count=64), 64),   # profiling size
    "hfield16_reuse": ("height_field", dict(n=708, grid_x=4, grid_y=4, view_res=1024, max_depth=8, sample_count=16), 16),
}
EXTRA = {"cbox8x8_reuse": dict(reuse=dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=8)),
         "cbox4_reuse": dict(reuse=dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4)),
         "hfield16_reuse": dict(reuse=dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4))}
SCHED_NAMES = {0: "warp_batch", 1: "regenerate", 2: "stream", 3: "wavefront", 5: "reuse", 6: "split"}
NODE_BYTES = 64       # one BVH node record as the kernels fetch it (csrc/dtypes.h); SURVEY sec.8d planned 32 B, the build uses 64 B two-child nodes
TRI_BYTES = 48


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_metrics():
    """Per-launch ncu counters of the dominant kernels at the benchmarked launch size (profiles/r02_ncu_metrics.json, written by
    scripts/ncu_metrics.py from the committed captures; instruction counts per path are a property of (kernel, workload, seed))."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")))
    except Exception:
        return {}


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
    pass of spp_per_pass samples.  Returns (Mpaths/s, description, seconds of the timed render)."""
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
                               "(CPU restatement of the reference render loop, oracle/, traversing the same BVH as the GPU build; closest-hit results identical to the GPU path)"), sec


def load_scenes_module():
    spec = importlib.util.spec_from_file_location("amvpt_scenes", os.path.join(ROOT, "mitsuba3-amvpt_b200", "scenes.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    return mod


def reference_arm(args, factory, kw, threads):
    """The reference's own CPU algorithm for the path on the box's host cores (oracle port; llvm_rgb / scalar_rgb cannot be
    built here).  Each step = one bounded sample of the workload; ms_per_step = measured seconds of that sample."""
    scenes = load_scenes_module()
    desc = getattr(scenes, factory)(**kw); desc.update(EXTRA.get(args.workload, {}))
    budget = max(2.0, min(20.0, 120.0 / max(args.steps + args.warmup, 1)))
    vals, secs, sample = [], [], ""
    for i in range(args.warmup + args.steps):
        v, sample, sec = cpu_reference_run(desc, 16, budget, threads)
        if i >= args.warmup:
            vals.append(v); secs.append(sec)
    v = float(np.mean(vals))
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": float(np.mean(secs)) * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                      "dtype": "f32", "data": "synthetic", "config": {"workload": args.workload, **kw, "timed_with": "steady clock around the CPU pass loop (no device involved)"},
                      "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": threads, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def roofline(pkg, s, args, spp, sched_arg, clk, device):
    """Roofline position of the dominant kernel.  Algorithmic bytes come from the kernels' own counters (one pass with counters),
    launch durations from CUDA events of a live render, instruction / DRAM / L2 counters per launch from the committed ncu capture of
    the same launch size, the L2 denominator from a live micro-benchmark, the HBM denominator from MEASURED_PEAKS.json."""
    s.render(seed=0, spp=16, develop=False, scheduler=sched_arg, stats=True)
    cs = s.last_stats
    rays = cs["closest_rays"] + cs["shadow_rays"]
    bvh_bytes_per_path = (cs["bvh_nodes"] * float(NODE_BYTES) + cs["tri_tests"] * float(TRI_BYTES)) / cs["paths"]
    # launch durations are per pass (16 spp), so a capped sample count gives the same per-launch figures; it bounds this single-GPU pass on the
    # multi-GPU / 1024-spp configurations
    s.render(seed=0, spp=min(spp, 64), develop=False, scheduler=sched_arg, time_stages=True); ks = s.last_stats
    hbm_peak, peak_src = peaks()
    sched = SCHED_NAMES[ks["scheduler_used"]]
    if sched == "wavefront":
        # dominant kernel: wf_trace_kernel.  Per ray: BVH nodes + triangles + ray read (32 B) + hit / flag traffic (24 B)
        kernel = "wf_trace_kernel"; b_path = bvh_bytes_per_path + 56.0 * rays / cs["paths"]; bound = "l2-latency"
        per_launch_s = ks["trace_seconds"] / max(ks["trace_launches"], 1); paths_per_launch = ks["paths"] / max(ks["trace_launches"], 1)
        share = ks["trace_seconds"] / ks["seconds"]
        note = "staged wavefront: the trace stage is %.0f%% of the device time; per-thread pointer chasing through the BVH, bound by L2 / HBM latency (ncu: long-scoreboard stalls), not by bandwidth" % (100 * share)
    elif sched == "reuse":
        kernel = "reuse pipeline"; b_path = 16.0 * max(1.0, s.n_reuse / 2) + bvh_bytes_per_path; bound = "issue"
        per_launch_s = ks["seconds"] / ks["launches"]; paths_per_launch = ks["paths"] / ks["launches"]; share = 1.0
        note = "sample reuse: Mpaths/s counts PRIMARY paths; every primary path also feeds up to n - 1 other views (config.reuse_samples)"
    else:
        kernel = "render_pass_kernel"; b_path = 16.0 + bvh_bytes_per_path; bound = "issue"
        per_launch_s = ks["seconds"] / ks["launches"]; paths_per_launch = ks["paths"] / ks["launches"]; share = 1.0
        note = ("persistent megakernel, path state in registers: the only HBM bytes are the film splat (16 B/path); BVH bytes are served by L1 for the Cornell box. "
                "Bound by SM issue slots x lane utilisation (ncu), so `achieved`/`peak` are warp instructions per second; the hbm / l2 objects give the byte lines of SURVEY sec.8d. "
                "Scenes whose tree has at most 32 leaves (the Cornell box) are traced through the flat leaf list (32 B per leaf box, all leaves tested per ray); "
                "bytes_per_path / nodes_per_ray come from the counter pass, which walks the tree")
    alg_bytes = paths_per_launch * b_path
    byte_rate = alg_bytes / per_launch_s / 1e9
    # L2 denominator: live micro-benchmark (32 MB buffer, 16-byte loads)
    try:
        l2_peak = pkg.membench(32 << 20, iters=200, device=device); hbm_read = pkg.membench(2 << 30, iters=3, device=device)
    except Exception:
        l2_peak, hbm_read = None, None
    out = {"bound": bound, "kernel": kernel, "kernel_share_of_step": share, "launch_ms": per_launch_s * 1e3, "paths_per_launch": paths_per_launch,
           "bytes_per_path": b_path, "algorithmic_bytes_per_launch": alg_bytes, "node_record_bytes": NODE_BYTES, "tri_record_bytes": TRI_BYTES,
           "rays_per_path": rays / cs["paths"], "nodes_per_ray": cs["bvh_nodes"] / max(rays, 1), "tris_per_ray": cs["tri_tests"] / max(rays, 1),
           "hbm": {"achieved": byte_rate, "peak": hbm_peak, "unit": "GB/s", "frac": byte_rate / hbm_peak, "peak_source": peak_src,
                   "hbm_read_gbs_membench": hbm_read},
           "l2": {"achieved": byte_rate, "peak": l2_peak, "unit": "GB/s", "frac": (byte_rate / l2_peak) if l2_peak else None,
                  "peak_source": "live amvpt_membench: 16-byte ld.global.cg loads over a 32 MB buffer"},
           "note": note}
    m = ncu_metrics().get(f"{kernel}:{args.workload}") or ncu_metrics().get(kernel)
    traffic = None
    if m:
        scale = paths_per_launch / m["units_per_launch"]            # counters are per launch of m["units_per_launch"] paths (or slots)
        traffic = m["dram_bytes"] * scale
        clock_hz = (clk.get("sm_mhz") or 1965.0) * 1e6
        issue_peak = 148 * 4 * clock_hz / 1e9                         # warp instructions per second, 4 schedulers per SM
        inst_rate = m["inst_executed"] * scale / per_launch_s / 1e9
        out.update({"issue_util": inst_rate / issue_peak, "issue_util_ncu": m.get("issue_active_pct", 0.0) / 100.0,
                    "warp_exec_eff": m["thread_inst_executed"] / m["inst_executed"] / 32.0, "lanes_per_inst": m["thread_inst_executed"] / m["inst_executed"],
                    "l2_bytes_per_launch": (m.get("lts_bytes") or 0) * scale, "l2_hit_pct": m.get("l2_hit_pct"), "l1_hit_pct": m.get("l1_hit_pct"),
                    "long_scoreboard_per_issue": m.get("long_scoreboard"), "ncu_source": m.get("source"), "ncu_workload": m.get("workload")})
        if bound == "issue":
            out.update({"achieved": inst_rate, "peak": issue_peak, "unit": "Gwarp-inst/s", "frac": inst_rate / issue_peak})
        else:
            out.update({"achieved": traffic / per_launch_s / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": traffic / per_launch_s / 1e9 / hbm_peak,
                        "achieved_is": "measured DRAM bytes per launch (ncu) / live launch time: the kernel is latency-bound, so this is far below 1"})
    else:
        out.update({"achieved": byte_rate, "peak": hbm_peak, "unit": "GB/s", "frac": byte_rate / hbm_peak, "achieved_is": "algorithmic bytes (no ncu capture for this kernel / workload)"})
    out["traffic"] = traffic
    return out


def b200_arm(args, factory, kw, spp, threads, rank, world, local, extra_workloads):
    import torch
    pkg = importlib.import_module("mitsuba3-amvpt_b200"); scenes = importlib.import_module("mitsuba3-amvpt_b200.scenes")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    class _Ptr:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # torch.distributed is the bootstrap (rendezvous, the timing max / sum over ranks); the two data exchanges of the path -- film gather
    # and the adaptive statistics all-reduce -- go through the library's own NCCL communicator (amvpt_comm, csrc/comm.cpp)
    comm = None
    if world > 1:
        comm = importlib.import_module("mitsuba3-amvpt_b200.multigpu").bootstrap_comm(pkg, rank, world, local)

    def measure(workload, factory, kw, spp, steps, warmup, adaptive, with_e2e=True):
        desc = getattr(scenes, factory)(**kw); desc.update(EXTRA.get(workload, {}))
        s = scenes.build(desc, device=local)
        info = s.info(); W, H = info["film_w"], info["film_h"]
        n_paths = W * H * spp
        film_ptr, _ = s.film_device_ptr()

        def step():
            s.render(seed=0, spp=spp, develop=False, scheduler=args.scheduler, rank=rank, world=world, adaptive=adaptive, comm=comm)
            if comm is not None:
                s.film_gather(comm, root=0)        # every rank sends the view tiles it rendered; rank 0 adds them into its film
            return s.last_stats

        for _ in range(warmup):
            step()
        clocks = ClockSampler(local); barrier(); clocks.start()
        dev_s, launches, render_s, traced = 0.0, 0, 0.0, 0
        t0 = time.perf_counter()
        for _ in range(steps):
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
        rank_ms = [[round(float(x) / steps * 1e3, 2) for x in t.tolist()] for t in per_rank]
        dev_s = float(tmax.item())
        tr = torch.tensor([float(traced)], dtype=torch.float64, device=f"cuda:{local}")
        if dist is not None:
            dist.all_reduce(tr, op=dist.ReduceOp.SUM)
        if adaptive:                       # the allocator spends the budget in whole passes: count the paths actually traced
            n_paths = int(tr.item()) // max(steps, 1)
        value = n_paths * steps / dev_s / 1e6
        res = {"s": s, "desc": desc, "value": value, "dev_s": dev_s, "launches": launches, "wall": wall, "clk": clk, "rank_ms": rank_ms, "n_paths": n_paths,
               "W": W, "H": H, "info": info, "spp": spp, "workload": workload, "kw": kw, "steps": steps}
        if with_e2e:
            # ---- end to end through the C ABI with host buffers (scene description -> BVH build -> upload + render + develop + D2H every step)
            def e2e_step():
                sc = scenes.build(desc, device=local)
                sc.render(seed=0, spp=spp, develop=False, scheduler=args.scheduler, rank=rank, world=world, adaptive=adaptive, comm=comm)
                if comm is not None:
                    sc.film_gather(comm, root=0)
                return sc.develop_pinned() if rank == 0 else None      # develop + D2H into the library's page-locked buffer (synchronises)
            e2e_steps = max(1, min(steps, 5))
            if dev_s / max(steps, 1) < 10.0:        # untimed first end-to-end step (page-locking, block cache); a 30-second step needs none
                e2e_step()
            barrier(); t0 = time.perf_counter()
            for _ in range(e2e_steps):
                e2e_step()
            barrier(); e2e_s = time.perf_counter() - t0
            tm = torch.tensor([e2e_s], dtype=torch.float64, device=f"cuda:{local}")
            if dist is not None:
                dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            res["e2e"] = {"value": n_paths * e2e_steps / float(tm.item()) / 1e6, "unit": "Mpaths/s", "steps": e2e_steps,
                          "h2d_bytes_per_step": int(sum(m["positions"].nbytes + m["faces"].nbytes for m in desc["meshes"]) + info["n_bvh_nodes"] * NODE_BYTES + info["n_prims"] * TRI_BYTES + info["n_views"] * 128),
                          "d2h_bytes_per_step": int(W * H * 3 * 4)}
        return res

    def line(r, with_roofline=True):
        s = r["s"]; W, H = r["W"], r["H"]
        sched = SCHED_NAMES[s.last_stats["scheduler_used"]]
        out = {"metric": METRIC, "value": r["value"], "unit": "Mpaths/s", "n_gpus": world, "steps": r["steps"], "warmup": args.warmup,
               "ms_per_step": r["dev_s"] / r["steps"] * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": r["workload"], **r["kw"], **({"reuse": EXTRA[r["workload"]]["reuse"], "reuse_samples": s.n_reuse} if r["workload"] in EXTRA else {}),
                          "spp": r["spp"], "adaptive": bool(args.adaptive), "scheduler": args.scheduler, "scheduler_used": sched, "quilt": [W, H], "paths_per_step": r["n_paths"],
                          "l2_policy": f"film {W * H * 16 / 1e6:.0f} MB > 126 MB L2 rewritten every pass" if W * H * 16 > 126e6 else "film smaller than L2 (small workload)",
                          "parallelism": f"views sharded over {world} rank(s); each rank sends its own view tiles to rank 0 (ncclSend/Recv behind the C ABI)" if world > 1 else "single GPU"},
               "clocks": r["clk"], "gpu_launches": r["launches"], "wall_s": r["wall"], "rank_ms_render_step": r["rank_ms"]}
        if "e2e" in r:
            out["e2e"] = r["e2e"]
        if with_roofline and rank == 0:
            out["roofline"] = roofline(pkg, s, argparse.Namespace(workload=r["workload"]), r["spp"], args.scheduler, r["clk"], local)
        return out

    main_r = measure(args.workload, factory, kw, spp, args.steps, args.warmup, args.adaptive)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    out = line(main_r)
    if world == 1 and not args.no_cpu_baseline:
        v, sample, _ = cpu_reference_run(main_r["desc"], 16, 15.0, threads)
        out["cpu_baseline"] = {"value": v, "unit": "Mpaths/s", "cores": threads, "kind": "port", "sample": sample}
    del main_r["s"]
    if extra_workloads:
        # BASELINE configs 1 and 4, short (1 warm-up + 2 / 1 timed steps), same measurement code; the headline stays config 2
        out["other_workloads"] = {}
        for wl, steps in extra_workloads:
            f2, k2, spp2 = WORKLOADS[wl]
            r2 = measure(wl, f2, k2, spp2, steps, 1, False)
            l2 = line(r2)
            out["other_workloads"][wl] = {k: l2[k] for k in ("value", "unit", "ms_per_step", "steps", "config", "gpu_launches", "e2e", "roofline", "clocks") if k in l2}
            del r2["s"]
    print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1); ap.add_argument("--steps", type=int, default=3); ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200"); ap.add_argument("--workload", default=None)
    ap.add_argument("--scheduler", default="auto"); ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the short config-1 / config-4 lines of the default single-GPU run")
    ap.add_argument("--adaptive", action="store_true", help="BASELINE configs 3 / 5: allocate spp per view from the variance statistics (all-reduced across ranks)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel (mean budget with --adaptive)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    default_run = args.workload is None and not args.adaptive and not args.spp
    args.workload = args.workload or "cbox8x8"
    factory, kw, spp = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp; kw = dict(kw, sample_count=spp)
    threads = os.cpu_count() or 1
    if args.impl == "reference":
        if rank == 0:
            reference_arm(args, factory, kw, threads)
        return
    extra = [("cbox4", 2), ("hfield16", 1)] if (default_run and world == 1 and not args.no_extra) else []
    b200_arm(args, factory, kw, spp, threads, rank, world, local, extra)


if __name__ == "__main__":
    main()
