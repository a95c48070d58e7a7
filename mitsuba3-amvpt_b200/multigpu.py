"""Multi-GPU plumbing (torch.distributed over NCCL on the GPU box, gloo in the CPU tests).  The render path itself has
no data-path collective: views are sharded, the scene is replicated, lane indices stay global.  The two
collectives are (1) the final film hand-off to rank 0 and (2) the all-reduce of per-view statistics in adaptive mode."""
import torch
import torch.distributed as dist


class DevicePtr:
    """Wraps a raw device pointer of the C ABI (amvpt_film_device_ptr) for torch.as_tensor without a copy."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def film_tensor(scene, device):
    info = scene.info(); ptr, _ = scene.film_device_ptr()
    return torch.as_tensor(DevicePtr(ptr, info["film_w"] * info["film_h"] * 4), device=device)


def reduce_film(film, dst=0):
    """Ranks own disjoint pixels (box filter; a jittered sample that rounds into a neighbour's pixel is the only overlap),
    so a sum-reduce to `dst` equals the gather of the owned tiles, including those strays."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def allreduce_stats(stats):
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    return stats


def make_allreduce_hook(device):
    """Callback for Scene.render(allreduce=...): sums `n` floats at a raw device pointer across ranks."""
    def hook(dev_ptr, n):
        t = torch.as_tensor(DevicePtr(dev_ptr, n), device=device)
        torch.cuda.current_stream().synchronize()
        allreduce_stats(t)
        torch.cuda.current_stream().synchronize()
    return hook
