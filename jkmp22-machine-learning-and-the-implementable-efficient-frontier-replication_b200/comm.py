"""The NCCL communicator of the path behind the C ABI (pfml_comm_*, csrc/comm.cu).

One process per GPU.  ``Communicator.from_torch_distributed()`` bootstraps the library's own NCCL communicator from an
initialised ``torch.distributed`` process group: rank 0 draws the unique id (``pfml_comm_unique_id``), the group
broadcasts its 128 bytes, every rank calls ``pfml_comm_init``.  The collectives themselves -- the all-reduce of the
per-window partial sums and the all-gather of the ridge coefficients (PFML_Search_Coef.py:77-135) -- are then issued by
the library on the caller's CUDA stream; torch.distributed carries no tensor of the data path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check


class Communicator:
    def __init__(self, id_bytes: bytes, rank: int, world: int, device: int):
        lib = _lib.load()
        assert len(id_bytes) == 128
        self.rank, self.world, self.device = int(rank), int(world), int(device)
        buf = C.create_string_buffer(id_bytes, 128)
        h = C.c_void_p()
        check(lib.pfml_comm_init(buf, self.rank, self.world, self.device, C.byref(h)), "pfml_comm_init")
        self.handle = h

    @staticmethod
    def unique_id() -> bytes:
        lib = _lib.load()
        buf = C.create_string_buffer(128)
        check(lib.pfml_comm_unique_id(buf), "pfml_comm_unique_id")
        return buf.raw

    @classmethod
    def from_torch_distributed(cls, device=None, group=None):
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        device = torch.cuda.current_device() if device is None else torch.device(device).index
        t = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            t = torch.frombuffer(bytearray(cls.unique_id()), dtype=torch.uint8).clone()
        if dist.get_backend(group) == "nccl":
            t = t.to(f"cuda:{device}")
        dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls(bytes(t.cpu().numpy().tobytes()), rank, world, device)

    def allreduce_sums(self, r_sum, d_sum, counts=None):
        """In-place FP64 sum over ranks of r_sum [nb, P], d_sum [nb, P, P] and counts [nb] (device tensors)."""
        lib = _lib.load()
        nb, P = r_sum.shape
        with torch.cuda.device(r_sum.device):
            check(lib.pfml_comm_allreduce_sums(self.handle, C.c_void_p(r_sum.data_ptr()), C.c_void_p(d_sum.data_ptr()),
                                               C.c_void_p(counts.data_ptr()) if counts is not None else None, P, nb,
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream)), "pfml_comm_allreduce_sums")

    def shard_range(self, nprob, rank=None):
        return shard_range(nprob, self.rank if rank is None else rank, self.world)

    def close(self):
        if self.handle:
            _lib.load().pfml_comm_destroy(self.handle)
            self.handle = None


def shard_range(nprob, rank, world):
    """pfml_ridge_shard_range: the flat (year, lambda) problems [lo, hi) rank ``rank`` of ``world`` solves for every p."""
    lib = _lib.load()
    lo, hi = C.c_int(), C.c_int()
    lib.pfml_ridge_shard_range(int(nprob), int(rank), int(world), C.byref(lo), C.byref(hi))
    return lo.value, hi.value
