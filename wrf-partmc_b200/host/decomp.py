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


def split(n, p):
    """(offset, length) of each of the p pieces an axis of n cells is cut into, the way WRF cuts a domain that
    does not divide evenly: the first n % p pieces are one cell longer."""
    q, r = divmod(n, p)
    out, off = [], 0
    for k in range(p):
        m = q + (1 if k < r else 0)
        out.append((off, m))
        off += m
    return out


def tile_of_rank(rank, px, py, gnx, gny):
    """strong-scaling layout: ONE gnx x gny domain cut into px x py tiles (module_wrf_top.F:505-510); returns
    (i0, nxl, j0, nyl) of the tile of `rank`."""
    rx, ry = patch_of_rank(rank, px, py)
    i0, nxl = split(gnx, px)[rx]
    j0, nyl = split(gny, py)[ry]
    return i0, nxl, j0, nyl


def tile_cells(gnx, nz, i0, nxl, j0, nyl):
    """global cell indices (WRF memory order: i fastest, then k, then j) of a tile, in the tile's own cell order"""
    import numpy as _np
    jl, k, il = _np.meshgrid(_np.arange(nyl), _np.arange(nz), _np.arange(nxl), indexing="ij")
    return ((i0 + il) + gnx * (k + nz * (j0 + jl))).ravel()


def patch_case(case, i0, nxl, j0, nyl):
    """the cells [i0, i0+nxl) x all k x [j0, j0+nyl) of a global case (synthetic.make_patch), in the patch's own
    cell order: what the rank owning that tile uploads"""
    import numpy as _np
    cells = tile_cells(case["nx"], case["nz"], i0, nxl, j0, nyl)
    off = _np.concatenate([[0], _np.cumsum(case["counts"])])
    idx = _np.concatenate([_np.arange(off[c], off[c + 1]) for c in cells])
    out = dict(case)
    out.update(n_cells=len(cells), nx=nxl, ny=nyl, counts=case["counts"][cells], vol=case["vol"][:, idx])
    for k in ("wclass", "wgroup", "id", "n_primary", "hyst", "lct", "gct"):
        out[k] = case[k][idx]
    for k in ("wtype", "wmag", "wexp", "T", "P", "RH", "rrho", "cell_volume"):
        if k in case:
            out[k] = case[k][cells]
    out["cells"] = cells
    return out


def gather_global(x_patch, gnx, nz, gny, px, py, device="cpu"):
    """every rank contributes the per-cell array of its tile ([n_cells_local, ...] in tile order); everybody gets
    the global array in WRF memory order (what MPI_Allgatherv + a reorder would do in WRF).  Tiles may differ
    in size, so every rank pads to the largest tile."""
    import numpy as _np
    import torch
    import torch.distributed as dist
    world = px * py
    x_patch = _np.ascontiguousarray(x_patch)
    tail = x_patch.shape[1:]
    if world == 1:
        return x_patch.copy()
    tiles = [tile_of_rank(r, px, py, gnx, gny) for r in range(world)]
    nmax = max(t[1] * t[3] for t in tiles) * nz
    buf = torch.zeros((nmax,) + tail, dtype=torch.from_numpy(x_patch).dtype, device=device)
    buf[:x_patch.shape[0]] = torch.from_numpy(x_patch).to(device)
    outs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf)
    g = _np.zeros((gnx * nz * gny,) + tail, x_patch.dtype)
    for r, (i0, nxl, j0, nyl) in enumerate(tiles):
        cells = tile_cells(gnx, nz, i0, nxl, j0, nyl)
        g[cells] = outs[r][:len(cells)].cpu().numpy()
    return g


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
