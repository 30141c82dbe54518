"""Patch decomposition of the periodic WRF domain over ranks (one process and one GPU per patch) and the
host-side plumbing around it: what WRF does with MPI (2-D x-y patches, module_wrf_top.F:505-510;
MPI_Cart_shift on the periodic communicator, interface/wrf_pmc_trans_aero.F90:153-159) is done here with
torch.distributed, which only carries the 128-byte NCCL id and the scalars of the benchmark; the particle
halos travel inside libpmcgpu.so (ncclSend / ncclRecv).  Works with the gloo backend on CPU (tests) and with
nccl on GPUs (bench.py).
"""
import numpy as np


def patch_grid(world):
    """px x py patches for `world` ranks: the most square factorisation with px >= py (WRF's default
    processor grid, nproc_x >= nproc_y for square domains)."""
    py = int(np.floor(np.sqrt(world)))
    while world % py:
        py -= 1
    return world // py, py


def patch_of_rank(rank, px, py):
    """(rx, ry) of a rank; rank = rx + px * ry like pmcgpu_comm_init (x fastest)."""
    assert 0 <= rank < px * py
    return rank % px, rank // px


def neighbours(rank, px, py):
    """ranks of the patches on the four sides, order y-, y+, x-, x+ (the side order of csrc/halo.cu), with
    periodic wrap: a direction that is not decomposed has the rank itself as neighbour."""
    rx, ry = patch_of_rank(rank, px, py)
    return [rx + px * ((ry - 1) % py), rx + px * ((ry + 1) % py), (rx - 1) % px + px * ry, (rx + 1) % px + px * ry]


def patch_extents(rank, px, py, nx, ny):
    """weak-scaling layout: every rank owns nx x ny columns; returns dict(global_nx, global_ny, i_offset,
    j_offset) for pmcgpu_config."""
    rx, ry = patch_of_rank(rank, px, py)
    return dict(global_nx=nx * px, global_ny=ny * py, i_offset=nx * rx, j_offset=ny * ry)


def broadcast_unique_id(make_id, rank, device="cpu"):
    """rank 0 makes the 128-byte NCCL id (make_id() -> bytes), everybody gets it (MPI_Bcast in WRF)."""
    import torch
    import torch.distributed as dist
    if rank == 0:
        raw = make_id()
        assert len(raw) == 128
        uid = torch.frombuffer(bytearray(raw), dtype=torch.uint8).clone()
    else:
        uid = torch.zeros(128, dtype=torch.uint8)
    uid = uid.to(device)
    dist.broadcast(uid, 0)
    return bytes(uid.cpu().numpy().tobytes())


def reduce_job(seconds, particles, device="cpu"):
    """whole-job figures of a timed region: max over ranks of each time, sum over ranks of each particle
    count (the contract of bench.py: value = units all ranks processed / slowest rank's time)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(seconds), dtype=torch.float64, device=device)
    n = torch.tensor(list(particles), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.cpu()], [float(x) for x in n.cpu()]
