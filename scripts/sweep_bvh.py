#!/usr/bin/env python
"""BVH builder parameter sweep (leaf size / SAH triangle cost): Mpaths/s and traversal counts.  Usage: sweep_bvh.py [cbox|hfield]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes')
wl = sys.argv[1] if len(sys.argv) > 1 else 'cbox'
for leaf in os.environ.get('LEAFS', '2,3,4,6').split(','):
    for ct in os.environ.get('CTS', '0.7,1.0,1.5,2.5').split(','):
        os.environ['AMVPT_MAX_LEAF'] = leaf; os.environ['AMVPT_SAH_CT'] = ct
        if wl == 'cbox':
            desc = scenes.cornell_box(grid_x=4, grid_y=4, view_res=256, cam_dist=400.0, sample_count=64); spp = 64
        else:
            desc = scenes.height_field(n=708, grid_x=2, grid_y=2, view_res=512, sample_count=16); spp = 16
        s = scenes.build(desc, device=0)
        best = 1e9
        for _ in range(3):
            s.render(seed=0, spp=spp, develop=False); best = min(best, s.last_stats['seconds'])
        paths = s.last_stats['paths']
        s.render(seed=0, spp=16, develop=False, stats=True); c = s.last_stats; rays = c['closest_rays'] + c['shadow_rays']
        print(f"leaf {leaf} ct {ct}: {paths / best / 1e6:8.1f} Mpaths/s  nodes/ray {c['bvh_nodes'] / rays:5.2f} tris/ray {c['tri_tests'] / rays:5.2f} n_nodes {s.info()['n_bvh_nodes']}", flush=True)
