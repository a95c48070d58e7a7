#!/usr/bin/env python
"""Scheduler / threshold sweep on the GPU box: Mpaths/s per setting.  Usage: python scripts/sweep.py [cbox|hfield] (env: TMS, RMS, SCHEDS, HF_N)"""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
wl = sys.argv[1] if len(sys.argv) > 1 else 'cbox'
if wl == 'cbox':
    desc = scenes.cornell_box(grid_x=4, grid_y=4, view_res=256, cam_dist=400.0, sample_count=64); spp = 64
else:
    desc = scenes.height_field(n=int(os.environ.get('HF_N', '708')), grid_x=2, grid_y=2, view_res=512, sample_count=16); spp = 16
t0 = time.time(); s = scenes.build(desc, device=0); print(f"build {time.time() - t0:.2f}s info {s.info()}", flush=True)


def run(tag, sched):
    global s
    s.render(seed=0, spp=spp, develop=False, scheduler=sched)
    best = 1e9
    for _ in range(3):
        s.render(seed=0, spp=spp, develop=False, scheduler=sched); best = min(best, s.last_stats['seconds'])
    print(f"{tag:44s} {s.last_stats['paths'] / best / 1e6:9.1f} Mpaths/s", flush=True)


for sched in os.environ.get('SCHEDS', 'regenerate,wavefront,stream').split(','):
    if sched == 'wavefront':
        for pool in os.environ.get('POOLS', '524288,1048576,2097152').split(','):
            os.environ['AMVPT_WF_POOL'] = pool
            s2 = scenes.build(desc, device=0); s, s_old = s2, s
            for fm in os.environ.get('FMS', '8,16').split(','):
                for lm in os.environ.get('LMS', '1,4,8,12,16,24,33').split(','):
                    os.environ['AMVPT_FETCH_MIN'] = fm; os.environ['AMVPT_LEAF_MIN'] = lm
                    run(f"wavefront pool {pool} fetch_min {fm} leaf_min {lm}", 'wavefront')
            s = s_old
        continue
    if sched != 'stream':
        run(sched, sched); continue
    for tm in os.environ.get('TMS', '12,16,20,24,28').split(','):
        for rm in os.environ.get('RMS', '4,8,16').split(','):
            os.environ['AMVPT_TRACE_MIN'] = tm; os.environ['AMVPT_REGEN_MIN'] = rm
            run(f"stream trace_min {tm} regen_min {rm}", 'stream')
