import sys, importlib, time, faulthandler, numpy as np
faulthandler.dump_traceback_later(50, exit=True)
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
mode = sys.argv[1]
desc = scenes.cornell_box(grid_x=2, grid_y=1, view_res=32, sample_count=16)
s = scenes.build(desc, device=0); print('built', s.info(), flush=True)
if mode == 'lanes':
    a = s.trace_lanes(0, 16, 0, 4096, max_hits=8); print('lanes ok', a['radiance'].mean(), a['stats'], flush=True)
else:
    t = time.time(); s.render(seed=0, spp=16, develop=False, scheduler=mode); print(mode, 'ok', time.time() - t, s.last_stats, flush=True)
    print(s.film_raw()[..., 3].mean(), flush=True)
