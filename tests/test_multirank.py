"""The N > 1 path.

CPU (gloo, world_size 2 and 4): the host-side decomposition logic every rank runs before it calls the
library: patch grid, neighbour table (MPI_Cart_shift order of interface/wrf_pmc_trans_aero.F90:153-159),
broadcast of the NCCL id, max-over-ranks / sum-over-ranks reduction of the benchmark.
GPU (needs >= 2 devices, skipped otherwise; run with `gpurun --gpus 2`): two ranks, one x-patch each, particle
halos through ncclSend / ncclRecv inside libpmcgpu.so; the union of the ranks' cells must equal, bit for bit,
the oracle's emulation of the same 2 x 1 decomposition (arrival order, ids, weights, bin grids)."""
import os
import socket
import tempfile

import numpy as np
import pytest

import multirank_worker as MW


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(mode, world, out_dir, cfg=None):
    import torch.multiprocessing as mp
    mp.spawn(MW.main, args=(mode, world, _free_port(), out_dir, cfg), nprocs=world, join=True)


@pytest.mark.parametrize("world", [2, 4])
def test_decomposition_host_logic_gloo(world):
    from wrf_partmc_b200.host import decomp
    with tempfile.TemporaryDirectory() as d:
        _spawn("cpu", world, d)
        r = [np.load(os.path.join(d, "cpu_%d.npz" % k)) for k in range(world)]
    px, py = int(r[0]["px"]), int(r[0]["py"])
    assert px * py == world and px >= py
    for k in range(world):
        assert np.array_equal(r[k]["uid"], np.arange(128, dtype=np.uint8))  # everybody holds rank 0's id
        nb = r[k]["nb"]
        # the neighbour relation is symmetric: my y- neighbour has me as y+, my x- neighbour has me as x+
        assert r[nb[0]]["nb"][1] == k and r[nb[1]]["nb"][0] == k
        assert r[nb[2]]["nb"][3] == k and r[nb[3]]["nb"][2] == k
        rx, ry = decomp.patch_of_rank(k, px, py)
        assert list(r[k]["ext"]) == [5 * px, 3 * py, 5 * rx, 3 * ry]
        # max over ranks of the times, sum over ranks of the particles
        assert list(r[k]["t"]) == [float(world), 2.0]
        assert list(r[k]["n"]) == [100.0 * world * (world + 1) / 2, 7.0 * world]
    # strong-scaling tiles of an uneven 7 x 5 domain cover it exactly once, and the gathered field is global order
    cover = np.zeros((5, 7), int)
    for k in range(world):
        i0, nxl, j0, nyl = r[k]["tile"]
        cover[j0:j0 + nyl, i0:i0 + nxl] += 1
        assert np.array_equal(r[k]["glob"][:, 0], np.arange(7 * 3 * 5)) and np.array_equal(r[k]["glob"][:, 1], 2.0 * np.arange(105))
    assert (cover == 1).all()
    # patches tile the global domain exactly once
    seen = set()
    for k in range(world):
        gnx, gny, i0, j0 = r[k]["ext"]
        for i in range(i0, i0 + 5):
            for j in range(j0, j0 + 3):
                assert (i, j) not in seen
                seen.add((i, j))
    assert len(seen) == 5 * px * 3 * py


def test_split_like_wrf():
    from wrf_partmc_b200.host import decomp
    assert decomp.split(150, 4) == [(0, 38), (38, 38), (76, 37), (113, 37)]
    assert sum(m for _, m in decomp.split(40, 3)) == 40
    c = decomp.tile_cells(6, 2, 2, 3, 1, 2)
    assert c[0] == 2 + 6 * (0 + 2 * 1) and len(c) == 3 * 2 * 2 and c[3] == 2 + 6 * (1 + 2 * 1)


def test_patch_slicing_matches_global_cell_order():
    """the patch a rank uploads is the global case restricted to its cells, in WRF memory order"""
    cfg = dict(gnx=4, nz=3, gny=2, n_per_cell=20, n_spec=4, seed=3)
    case = MW.global_case(cfg)
    seen = []
    for rx in range(2):
        pc = MW.patch_case(case, rx * 2, 2, 0, 2)
        assert pc["n_cells"] == 12 and pc["vol"].shape[1] == pc["counts"].sum()
        seen += list(pc["cells"])
        off = np.concatenate([[0], np.cumsum(case["counts"])])
        c0 = pc["cells"][0]
        assert np.array_equal(pc["id"][:pc["counts"][0]], case["id"][off[c0]:off[c0 + 1]])
    assert sorted(seen) == list(range(case["n_cells"]))


@pytest.mark.gpu
@pytest.mark.parametrize("n_spec,is_open,comps", [(4, False, False), (20, False, False), (4, True, False),
                                                  (4, False, True), (4, True, True)])
def test_two_rank_transport_bit_exact(n_spec, is_open, comps):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import common
    from oracle import oracle as O
    from wrf_partmc_b200.host import synthetic
    cfg = dict(gnx=8, nz=4, gny=4 if is_open else 3, px=2, py=1, n_per_cell=200, n_spec=n_spec, seed=5, n_steps=3,
               open=is_open, components=comps)
    case = MW.global_case(cfg)
    o = common.oracle_from_case(O, case)
    if comps:  # component lists and one opaque double per particle cross the rank boundary with their records
        off0 = np.concatenate([[0], np.cumsum(case["counts"])])
        for c in range(case["n_cells"]):
            o.set_components(c, *MW.components_of_ids(case["id"][off0[c]:off0[c + 1]]))
    for c in range(case["n_cells"]):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
    o.trans_setup(cfg["gnx"], cfg["nz"], cfg["gny"], px=2, py=1)
    nc = o.num_conc_by_class()
    if is_open:
        o.trans_set_boundary(True, **MW.open_boundary(case))
    f = synthetic.transport_fields(case, nc, periodic=not is_open)
    with tempfile.TemporaryDirectory() as d:
        np.save(os.path.join(d, "num_conc.npy"), nc)
        _spawn("gpu", 2, d, cfg)
        r = [dict(np.load(os.path.join(d, "gpu_%d.npz" % k))) for k in range(2)]
    moved = sum(o.wrf_pmc_trans_driver(f, step=s) for s in range(1, cfg["n_steps"] + 1))
    # (open domain: the oracle also counts what moves into boundary cells and leaves the domain; the device counts arrivals)
    assert moved > 0 and (is_open or moved == sum(int(x["moved"]) for x in r))
    n_checked = 0
    for x in r:
        off = np.concatenate([[0], np.cumsum(x["count"])])
        for lc, c in enumerate(x["cells"]):
            q = o.get_cell(int(c))
            a, b = off[lc], off[lc + 1]
            assert b - a == q["vol"].shape[1], (c, b - a, q["vol"].shape[1])
            assert np.array_equal(x["id"][a:b], q["id"]) and np.array_equal(x["wclass"][a:b], q["wclass"])
            if is_open:  # particles sampled from the background: 1e-12 (tests/test_gpu_open.py)
                np.testing.assert_allclose(x["vol"][:, a:b], q["vol"], rtol=1e-12, atol=0)
            else:
                assert np.array_equal(x["vol"][:, a:b], q["vol"])
            np.testing.assert_allclose(x["wmag"][lc].ravel(), o.get_weights(int(c))[1], rtol=1e-14)
            onb, omin, omax, _ = o.grid(int(c))
            assert x["nb"][lc] == onb
            assert x["next_id"][lc] == o.L.pmco_get_next_id(o.h, int(c))
            if comps:
                onc, osrc, otm, oopq = o.get_components(int(c), 1)
                assert np.array_equal(x["comp_n"][a:b], onc) and np.array_equal(x["comp_src"][a:b], osrc)
                assert np.array_equal(x["comp_time"][a:b], otm) and np.array_equal(x["comp_opq"][a:b], oopq)
            n_checked += 1
    assert n_checked == case["n_cells"]


@pytest.mark.gpu
def test_two_rank_halo_overflow_returns_on_every_rank():
    """one particle message too small for what crosses the patch side: the exchange still completes on both ranks
    (no rank-local early return between two collective exchanges), the overflow is a status, the state consistent"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import common
    from oracle import oracle as O
    cfg = dict(gnx=8, nz=4, gny=3, px=2, py=1, n_per_cell=200, n_spec=4, seed=5, n_steps=1, expect_overflow=True)
    case = MW.global_case(cfg)
    o = common.oracle_from_case(O, case)
    for c in range(case["n_cells"]):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
    os.environ["PMCGPU_HALO_SLAB_CAP"] = "8"
    try:
        with tempfile.TemporaryDirectory() as d:
            np.save(os.path.join(d, "num_conc.npy"), o.num_conc_by_class())
            _spawn("gpu", 2, d, cfg)   # returns only if neither rank hangs
            r = [dict(np.load(os.path.join(d, "gpu_%d.npz" % k))) for k in range(2)]
    finally:
        del os.environ["PMCGPU_HALO_SLAB_CAP"]
    assert any(int(x["code"]) == 3 for x in r)            # PMCGPU_ERR_CAPACITY where the overflow was seen
    assert all(bool(x["unique"]) for x in r)              # no record duplicated
    assert all(int(x["max_count"]) <= 3 * 200 + 64 for x in r)


@pytest.mark.gpu
def test_two_contexts_on_two_devices_in_one_process():
    """kernel attributes (shared memory above 48 KB) and occupancy figures are kept per context, and every entry
    point selects the context's device: two contexts on two devices in ONE process give the same results"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import common
    import wrf_partmc_b200 as W
    from wrf_partmc_b200.host import synthetic
    case = synthetic.make_patch(4, 3, 3, 900, 20, seed=3)
    sp = case["sp"]
    res = []
    gs = []
    for dev in (0, 1):
        data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
        g = W.PartMCGPU(data, nx=4, nz=3, ny=3, n_group=1, n_class=2, n_part_ideal=case["n_part_ideal"],
                        cell_capacity=2400, seed=case["seed"], device=dev)
        gs.append(g)
    f = None
    for g in gs:  # interleaved use of the two contexts
        g.set_particles(case["counts"], case["vol"], case["wgroup"], case["wclass"], case["id"], case["n_primary"],
                        case["hyst"], case["lct"], case["gct"])
        g.set_weights(case["wtype"], case["wmag"], case["wexp"])
        g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
                  cell_volume=case["cell_volume"])
        g.condense_equilib_particles()
    f = synthetic.patch_fields(synthetic.transport_fields(case, gs[0].partmc_to_wrf()), 0, 4, 0, 3)
    for step in range(2):
        for g in gs:
            g.partmc_timestep(temp=case["T"] - 0.1, pressure=case["P"] - 100.0, elapsed_time=step * 60.0, dt=60.0,
                              chem_dt=60.0, do_coagulation=True, coag_kernel_type=W.KERNEL_BROWN,
                              do_condensation=True, step=2 * step)
            g.wrf_pmc_trans_driver(f, step=2 * step + 1)
    P0, P1 = gs[0].get_particles(), gs[1].get_particles()
    assert np.array_equal(P0["count"], P1["count"]) and np.array_equal(P0["id"], P1["id"])
    assert np.array_equal(P0["vol"], P1["vol"])
    assert np.array_equal(gs[0].rel_humid(), gs[1].rel_humid())
