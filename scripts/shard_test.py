import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
desc = scenes.cornell_box(grid_x=8, grid_y=8, view_res=512, cam_dist=400.0, sample_count=64)
s = scenes.build(desc, device=0)
for _ in range(3):
    s.render(seed=0, spp=64, develop=False)
for world, rank in ((1, 0), (64, 5), (1, 0)):
    best = None
    for _ in range(3):
        s.render(seed=0, spp=64, develop=False, rank=rank, world=world)
        if best is None or s.last_stats['seconds'] < best['seconds']: best = dict(s.last_stats)
    st = best
    print(world, rank, st['paths'], round(st['seconds'] * 1e3, 2), 'ms', round(st['paths'] / st['seconds'] / 1e6, 1), 'Mpaths/s', flush=True)
