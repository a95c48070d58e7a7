import sys, os, importlib, time, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
wl = sys.argv[1] if len(sys.argv) > 1 else 'cbox'
def run(tag, sched='regenerate'):
    if wl == 'cbox': desc = scenes.cornell_box(grid_x=4, grid_y=4, view_res=256, cam_dist=400.0, sample_count=64); spp = 64
    else: desc = scenes.height_field(n=int(os.environ.get('HF_N', '708')), grid_x=2, grid_y=2, view_res=512, sample_count=16); spp = 16
    t0 = time.time(); s = scenes.build(desc, device=0); tb = time.time() - t0
    s.render(seed=0, spp=spp, develop=False, scheduler=sched)
    best = 1e9
    for _ in range(3):
        s.render(seed=0, spp=spp, develop=False, scheduler=sched); best = min(best, s.last_stats['seconds'])
    st = s.last_stats
    s.render(seed=0, spp=16, develop=False, scheduler=sched, stats=True); c = s.last_stats
    rays = c['closest_rays'] + c['shadow_rays']
    print(f"{tag:40s} {st['paths'] / best / 1e6:9.1f} Mpaths/s  build {tb:.2f}s nodes/ray {c['bvh_nodes'] / rays:.2f} tris/ray {c['tri_tests'] / rays:.2f} rays/path {rays / c['paths']:.2f} info {s.info()['n_bvh_nodes']}", flush=True)
for leaf in (os.environ.get('LEAFS', '4,2,1,3').split(',')):
    for ct in (os.environ.get('CTS', '1.5').split(',')):
        os.environ['AMVPT_MAX_LEAF'] = leaf; os.environ['AMVPT_SAH_CT'] = ct
        for rm in (os.environ.get('RMS', '8').split(',')):
            os.environ['AMVPT_REGEN_MIN'] = rm
            run(f"leaf {leaf} ct {ct} regen_min {rm}")
os.environ['AMVPT_MAX_LEAF'] = '4'; os.environ['AMVPT_SAH_CT'] = '1.5'
run("warp_batch leaf 4", 'warp_batch')
