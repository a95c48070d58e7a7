#!/usr/bin/env python
"""Multi-GPU check on a real NCCL box: torchrun --nproc-per-node N scripts/shard_test.py
Every rank renders its view tiles, amvpt_film_gather moves them to rank 0 (ncclSend / ncclRecv behind the C ABI), rank 0 compares the result
with its own single-rank render of the whole quilt; the same for the adaptive allocator (statistics all-reduced by the library) and for the
whole-quilt reduction that sample reuse needs."""
import importlib, os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
pkg = importlib.import_module('mitsuba3-amvpt_b200'); scenes = importlib.import_module('mitsuba3-amvpt_b200.scenes'); mg = importlib.import_module('mitsuba3-amvpt_b200.multigpu')
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local); dist.init_process_group('nccl', device_id=torch.device('cuda', local))
comm = mg.bootstrap_comm(pkg, rank, world, local)
desc = scenes.cornell_box_ref(grid_x=4, grid_y=4, view_res=64, cam_dist=400.0, sample_count=64)
s = scenes.build(desc, device=local)
for adaptive in (False, True):
    s.render(seed=2, spp=64, develop=False, rank=rank, world=world, adaptive=adaptive, comm=comm); s.film_gather(comm, root=0); torch.cuda.synchronize()
    got = s.film_raw().copy(); alloc = s.adaptive_alloc()
    if rank == 0:
        s.render(seed=2, spp=64, develop=False, adaptive=adaptive); ref = s.film_raw()
        assert np.array_equal(got[..., 3], ref[..., 3]), f'adaptive={adaptive}: weights differ'
        assert np.abs(got[..., :3] - ref[..., :3]).max() <= 2e-5 * ref[..., :3].max()
        if adaptive: assert np.array_equal(alloc, s.adaptive_alloc()) and len({tuple(r) for r in alloc}) > 1
        print(f'rank 0: world {world} adaptive={adaptive}: gathered film equals the single-rank film ({int(ref[..., 3].sum())} samples)', flush=True)
d2 = dict(desc); d2['reuse'] = dict(sa_reuse=True, sa_mis=True, fast_mis=False, reuse_count=4)
s2 = scenes.build(d2, device=local)
s2.render(seed=2, spp=16, develop=False, rank=rank, world=world); s2.film_gather(comm, root=0); torch.cuda.synchronize(); got = s2.film_raw().copy()
if rank == 0:
    s2.render(seed=2, spp=16, develop=False); ref = s2.film_raw()
    assert np.allclose(got[..., 3], ref[..., 3], rtol=1e-5, atol=1e-5) and np.abs(got[..., :3] - ref[..., :3]).max() <= 1e-4 * ref[..., :3].max()
    print(f'rank 0: world {world} sample reuse: reduced film equals the single-rank film', flush=True)
dist.barrier(); dist.destroy_process_group()
