"""Worker of the multi-rank tests (spawned once per rank by test_multirank.py).

cpu mode : world_size ranks on the gloo backend exercise the host-side decomposition logic
           (wrf-partmc_b200/host/decomp.py): patch grid, neighbour table, NCCL-id broadcast, job reduction.
gpu mode : one rank per GPU; every rank owns one x-patch of a doubly periodic domain, the particle halos
           travel through libpmcgpu.so's ncclSend / ncclRecv; every rank dumps its cells for the parent,
           which compares them bit for bit with the oracle's emulation of the same decomposition.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def patch_case(case, i0, nxl, j0, nyl):
    from wrf_partmc_b200.host import decomp
    return decomp.patch_case(case, i0, nxl, j0, nyl)


def global_case(cfg):
    from wrf_partmc_b200.host import synthetic
    return synthetic.make_patch(cfg["gnx"], cfg["nz"], cfg["gny"], cfg["n_per_cell"], cfg["n_spec"], seed=cfg["seed"])


def open_boundary(case):
    """boundary conditions of the open test domain in GLOBAL cell order (wind from the south-west)"""
    n_cells, nx, nz, ny, S = case["n_cells"], case["nx"], case["nz"], case["ny"], case["n_spec"]
    vf = np.zeros((2, S))
    vf[0, 0], vf[0, 2] = 0.6, 0.4
    vf[1, 1], vf[1, 2] = 0.5, 0.5
    gi = np.arange(n_cells) % nx
    gj = np.arange(n_cells) // (nx * nz)
    throw = np.zeros(n_cells, np.int32)
    throw[(gi == 0) & (gj == 1)] = 1
    return dict(mode_type=[1, 1], weight_class=[1, 2], char_radius=[1e-8, 5.8e-8], log10_std_dev=[0.161, 0.217],
                vol_frac=vf, num_conc=np.tile(np.array([3.2e10, 2.9e10]) * 1.3, (n_cells, 1)) * case["rrho"][:, None],
                outflow=((gi == nx - 1) | (gj == ny - 1)).astype(np.int32), throw_away=throw)


def run_cpu(rank, world, port, out_dir):
    import torch.distributed as dist
    from wrf_partmc_b200.host import decomp
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    px, py = decomp.patch_grid(world)
    uid = decomp.broadcast_unique_id(lambda: bytes(range(128)), rank)
    nb = decomp.neighbours(rank, px, py)
    ext = decomp.patch_extents(rank, px, py, 5, 3)
    t, n = decomp.reduce_job([1.0 + rank, 2.0 - rank], [100.0 * (rank + 1), 7.0])
    # strong-scaling layout: one 7 x 5 domain (3 levels) cut unevenly; every rank contributes "global cell index"
    # of its tile, gather_global must return 0 .. n-1 in WRF memory order on every rank
    i0, nxl, j0, nyl = decomp.tile_of_rank(rank, px, py, 7, 5)
    mine = decomp.tile_cells(7, 3, i0, nxl, j0, nyl).astype(np.float64)
    glob = decomp.gather_global(np.stack([mine, 2.0 * mine], axis=1), 7, 3, 5, px, py)
    np.savez(os.path.join(out_dir, "cpu_%d.npz" % rank), uid=np.frombuffer(uid, np.uint8), nb=nb, px=px, py=py,
             ext=np.array([ext["global_nx"], ext["global_ny"], ext["i_offset"], ext["j_offset"]]), t=t, n=n,
             glob=glob, tile=np.array([i0, nxl, j0, nyl]))
    dist.barrier()
    dist.destroy_process_group()


def components_of_ids(ids):
    """deterministic component lists from the particle ids (both sides of the 2-rank component test use it)"""
    ids = np.asarray(ids, np.int64)
    nc = (1 + ids % 3).astype(np.int32)
    q = np.arange(20)[None, :]
    src = ((ids[:, None] + q) % 7 + 1).astype(np.int32)
    tm = ids[:, None] * 0.5 + q
    opq = (ids * 0.25)[:, None]
    return nc, src, tm, opq


def run_gpu(rank, world, port, out_dir, cfg):
    import torch.distributed as dist
    import wrf_partmc_b200 as W
    from wrf_partmc_b200.host import decomp, synthetic
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    px, py = cfg["px"], cfg["py"]
    nxl, nyl = cfg["gnx"] // px, cfg["gny"] // py
    rx, ry = decomp.patch_of_rank(rank, px, py)
    case = global_case(cfg)
    pc = patch_case(case, rx * nxl, nxl, ry * nyl, nyl)
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    is_open = bool(cfg.get("open"))
    g = W.PartMCGPU(data, nx=nxl, nz=cfg["nz"], ny=nyl, n_group=1, n_class=case["n_class"],
                    n_part_ideal=case["n_part_ideal"], cell_capacity=3 * cfg["n_per_cell"] + 64, seed=case["seed"],
                    device=rank, global_nx=cfg["gnx"], global_ny=cfg["gny"], i_offset=rx * nxl, j_offset=ry * nyl,
                    periodic_x=not is_open, periodic_y=not is_open)
    g.set_particles(pc["counts"], pc["vol"], pc["wgroup"], pc["wclass"], pc["id"], pc["n_primary"], pc["hyst"],
                    pc["lct"], pc["gct"])
    g.set_weights(pc["wtype"], pc["wmag"], pc["wexp"])
    g.set_env(temp=pc["T"], pressure=pc["P"], rel_humid=pc["RH"], inverse_density=pc["rrho"],
              cell_volume=pc["cell_volume"])
    N = int(case["counts"].sum())
    g.set_next_id(np.full(pc["n_cells"], N // case["n_cells"] + 2, np.int64))
    if cfg.get("components"):  # before comm_init: the halo record carries the cold part
        g.track_components(1)
        g.set_components(*components_of_ids(pc["id"]))
    uid = decomp.broadcast_unique_id(W.comm_unique_id, rank)
    g.comm_init(uid, world, rank, px, py)
    nc = np.load(os.path.join(out_dir, "num_conc.npy"))
    f = synthetic.transport_fields(case, nc, periodic=not is_open)
    fp = synthetic.patch_fields(f, rx * nxl, nxl, ry * nyl, nyl, periodic=not is_open)
    if is_open:
        b = open_boundary(case)
        cells = pc["cells"]
        g.set_boundary(b["mode_type"], b["weight_class"], b["char_radius"], b["log10_std_dev"], b["vol_frac"],
                       b["num_conc"][cells], b["outflow"][cells], b["throw_away"][cells])
    moved = 0
    if cfg.get("expect_overflow"):
        # the halo message of this job is too small (PMCGPU_HALO_SLAB_CAP): every rank must come back from the
        # collective exchange with PMCGPU_ERR_CAPACITY on the rank(s) that saw the overflow -- nobody may hang
        code = 0
        try:
            g.wrf_pmc_trans_driver(fp, step=1)
        except W.PmcGpuError as e:
            code = e.code
        P = g.get_particles()
        np.savez(os.path.join(out_dir, "gpu_%d.npz" % rank), code=code, unique=len(np.unique(P["id"])) == len(P["id"]),
                 max_count=int(P["count"].max()))
        dist.barrier()
        dist.destroy_process_group()
        return
    for step in range(1, cfg["n_steps"] + 1):
        g.wrf_pmc_trans_driver(fp, step=step)
        moved += int(g.trans_moved().sum())
    P = g.get_particles()
    _, m, _ = g.get_weights()
    nb, gmin, gmax = g.grid()
    extra = {}
    if cfg.get("components"):
        cn, cs, ct, cq = g.get_components()
        extra = dict(comp_n=cn, comp_src=cs, comp_time=ct, comp_opq=cq)
    np.savez(os.path.join(out_dir, "gpu_%d.npz" % rank), cells=pc["cells"], count=P["count"], vol=P["vol"], id=P["id"],
             wclass=P["wclass"], n_primary=P["n_primary"], wmag=m, nb=nb, gmin=gmin, gmax=gmax, moved=moved,
             next_id=g.get_next_id(), **extra)
    dist.barrier()
    dist.destroy_process_group()


def main(rank, mode, world, port, out_dir, cfg):
    if mode == "cpu":
        run_cpu(rank, world, port, out_dir)
    else:
        run_gpu(rank, world, port, out_dir, cfg)
