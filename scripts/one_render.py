#!/usr/bin/env python
"""One render for profiling: python scripts/one_render.py [cbox|hfield] [scheduler] [spp]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
wl = sys.argv[1] if len(sys.argv) > 1 else 'cbox'; sched = sys.argv[2] if len(sys.argv) > 2 else 'wavefront'; spp = int(sys.argv[3]) if len(sys.argv) > 3 else 16
if wl == 'cbox':
    desc = scenes.cornell_box(grid_x=4, grid_y=4, view_res=256, cam_dist=400.0, sample_count=spp)
else:
    desc = scenes.height_field(n=int(os.environ.get('HF_N', '708')), grid_x=2, grid_y=2, view_res=512, sample_count=spp)
s = scenes.build(desc, device=0)
for _ in range(int(os.environ.get('REPS', '2'))):
    s.render(seed=0, spp=spp, develop=False, scheduler=sched)
print(s.last_stats, s.last_stats['paths'] / s.last_stats['seconds'] / 1e6, 'Mpaths/s')
