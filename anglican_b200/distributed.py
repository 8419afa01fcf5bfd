"""Host-side plumbing of a sharded population: one process per GPU.

`torch.distributed` is used only to hand rank 0's NCCL unique id to the other
ranks; everything after that (IPC heap exchange, per-step scalars, particle
migration) happens inside libanglican_pop.so (csrc/comm.cu, fabric.cuh)."""
import ctypes as C

from . import _lib as L


def shard_range(n_total, rank, world):
    """contiguous block of particles owned by `rank`:
    global id = rank * n_local + local"""
    if n_total % world:
        raise ValueError("number of particles must be divisible by the number of ranks")
    n_local = n_total // world
    return rank * n_local, (rank + 1) * n_local


def nccl_unique_id(rank, src=0, group=None):
    """128-byte ncclUniqueId created on `src` and broadcast to every rank
    (works with the gloo and the nccl backend)."""
    import torch
    import torch.distributed as dist
    backend = dist.get_backend(group)
    dev = "cuda" if backend == "nccl" else "cpu"
    buf = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == src:
        raw = C.create_string_buffer(128)
        rc = L.load().apop_get_unique_id(raw)
        if rc != L.APOP_OK:
            raise L.ApopError(rc, "apop_get_unique_id (is libnccl.so.2 loadable?)")
        buf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
    dist.broadcast(buf, src, group=group)
    return bytes(buf.cpu().numpy().tobytes())
