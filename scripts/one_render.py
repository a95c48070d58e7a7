#!/usr/bin/env python
"""One render for profiling: python scripts/one_render.py [workload of bench.py | cbox | hfield] [scheduler] [spp] -- REPS renders (default 2)"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
import bench
wl = sys.argv[1] if len(sys.argv) > 1 else 'cbox'; sched = sys.argv[2] if len(sys.argv) > 2 else 'auto'; spp = int(sys.argv[3]) if len(sys.argv) > 3 else 16
if wl in bench.WORKLOADS:
    factory, kw, _ = bench.WORKLOADS[wl]
    desc = getattr(scenes, factory)(**dict(kw, sample_count=spp)); desc.update(bench.EXTRA.get(wl, {}))
elif wl == 'cbox':
    desc = scenes.cornell_box(grid_x=4, grid_y=4, view_res=256, cam_dist=400.0, sample_count=spp)
else:
    desc = scenes.height_field(n=int(os.environ.get('HF_N', '708')), grid_x=2, grid_y=2, view_res=512, sample_count=spp)
s = scenes.build(desc, device=0)
for _ in range(int(os.environ.get('REPS', '2'))):
    s.render(seed=0, spp=spp, develop=False, scheduler=sched)
print(s.last_stats, s.last_stats['paths'] / s.last_stats['seconds'] / 1e6, 'Mpaths/s')
