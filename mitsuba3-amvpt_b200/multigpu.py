"""Multi-GPU bootstrap.  The path's two exchanges -- the film gather and the adaptive statistics all-reduce -- live behind the C ABI
(csrc/comm.cpp: amvpt_comm_*, amvpt_film_gather; NCCL bound at run time).  What remains here is the rendezvous: rank 0 creates the NCCL
unique id and torch.distributed (NCCL on the GPU box, gloo in the CPU tests) hands its 128 bytes to the other ranks.  DevicePtr wraps a
raw device pointer of the C ABI for tests that emulate the all-reduce on one GPU."""
import torch
import torch.distributed as dist


class DevicePtr:
    """Wraps a raw device pointer of the C ABI (amvpt_film_device_ptr, the allreduce hook's argument) for torch.as_tensor without a copy."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def exchange_unique_id(make_id, rank):
    """Rank 0 calls make_id() -> 128 bytes; every rank returns them (broadcast over the initialised process group)."""
    box = [make_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def bootstrap_comm(pkg, rank, world, device):
    """An amvpt Comm (ncclCommInitRank behind the C ABI) for this process; torch.distributed must be initialised."""
    uid = exchange_unique_id(pkg.Comm.unique_id, rank)
    return pkg.Comm(uid, rank, world, device)
