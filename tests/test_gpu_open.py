"""Open (non-periodic) domains on the GPU against the oracle (wrf_pmc_driver.F90:157-167, wrf_pmc_trans.F90:116-136,
wrf_pmc_trans_aero.F90:497-506, 689-766, 811-955, 976-1098; aero_state.F90:3529): the outermost ring of cells is the
boundary condition.  Bar: the same as for the periodic transport -- which particles move where, arrival order, ids,
weights, bin grids equal given identical draws (volumes of freshly sampled particles to 1e-12), for interior AND
boundary cells (inflow sampling, copies of the
interior neighbour where the wind blows out, resampling of the background where it blows in, throw-away cells)."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def _open_case(W, O, nx=7, nz=3, ny=6, npc=160, S=4, seed=4, bg_scale=1.0):
    from wrf_partmc_b200.host import synthetic
    case = synthetic.make_patch(nx, nz, ny, npc, S, seed=seed)
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    g = W.PartMCGPU(data, nx=nx, nz=nz, ny=ny, n_group=1, n_class=2, n_part_ideal=case["n_part_ideal"],
                    cell_capacity=4 * npc + 64, seed=case["seed"], periodic_x=False, periodic_y=False)
    g.set_particles(case["counts"], case["vol"], case["wgroup"], case["wclass"], case["id"], case["n_primary"],
                    case["hyst"], case["lct"], case["gct"])
    g.set_weights(case["wtype"], case["wmag"], case["wexp"])
    g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
              cell_volume=case["cell_volume"])
    n_cells = case["n_cells"]
    nid = np.full(n_cells, int(case["counts"].sum()) // n_cells + 2, np.int64)
    g.set_next_id(nid)
    o = common.oracle_from_case(O, case)
    for c in range(n_cells):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
    o.trans_setup(nx, nz, ny, px=1, py=1)
    # background: the two modes of the initial population, one per weight class (number per kg of air x density)
    vf = np.zeros((2, S))
    vf[0, 0], vf[0, 2] = 0.6, 0.4
    vf[1, 1], vf[1, 2] = 0.5, 0.5
    numc = np.tile(np.array([3.2e10, 2.9e10]) * bg_scale, (n_cells, 1)) * case["rrho"][:, None]
    gi = np.arange(n_cells) % nx
    gj = np.arange(n_cells) // (nx * nz)
    outflow = ((gi == nx - 1) | (gj == ny - 1)).astype(np.int32)  # wind from the south-west
    throw = np.zeros(n_cells, np.int32)
    throw[(gi == 0) & (gj == 2)] = 1
    throw[(gj == 0) & (gi == 3)] = 1
    bnd = dict(mode_type=[1, 1], weight_class=[1, 2], char_radius=[1e-8, 5.8e-8], log10_std_dev=[0.161, 0.217],
               vol_frac=vf, num_conc=numc, outflow=outflow, throw_away=throw)
    g.set_boundary(**bnd)
    o.trans_set_boundary(True, **bnd)
    f = synthetic.transport_fields(case, o.num_conc_by_class(), periodic=False)
    fp = synthetic.patch_fields(f, 0, nx, 0, ny, periodic=False)
    return case, g, o, f, fp, (gi, gj)


def _same(g, o, case, water_rtol=None):
    P = g.get_particles()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    _, m, _ = g.get_weights()
    nb, gmin, gmax = g.grid()
    iw = case["sp"]["i_water"] - 1
    for c in range(case["n_cells"]):
        q = o.get_cell(c)
        a, b = off[c], off[c + 1]
        assert b - a == q["vol"].shape[1], "cell %d: %d on device, %d in oracle" % (c, b - a, q["vol"].shape[1])
        for key in ("id", "wclass", "n_primary"):
            assert np.array_equal(P[key][a:b], q[key]), (c, key)
        # volumes of particles SAMPLED from the background carry the last-bit differences of pow / log / erf / cos of
        # the two math libraries (1e-13, like tests/test_gpu_emit.py); every record that was only moved or copied
        # is unchanged, so the bound holds for all of them
        if water_rtol is None:
            np.testing.assert_allclose(P["vol"][:, a:b], q["vol"], rtol=1e-12, atol=0, err_msg=str(c))
        else:
            dry = [s for s in range(P["vol"].shape[0]) if s != iw]
            np.testing.assert_allclose(P["vol"][dry, a:b], q["vol"][dry], rtol=1e-12, atol=0, err_msg=str(c))
            np.testing.assert_allclose(P["vol"][iw, a:b], q["vol"][iw], rtol=water_rtol, atol=1e-40)
        np.testing.assert_allclose(m[c].ravel(), o.get_weights(c)[1], rtol=1e-14)
        assert nb[c] == o.grid(c)[0], c
    assert np.array_equal(g.get_next_id(), np.array([o.L.pmco_get_next_id(o.h, c) for c in range(case["n_cells"])]))


@pytest.mark.parametrize("bg_scale", [1.0, 1.6])
def test_open_domain_transport_bit_exact(mods, bg_scale):
    """trans_aero on an open domain, 4 calls: interior sampling, boundary inflow, refresh (copies, resampling with
    removals, with additions when the background holds more particles than the cell, throw-away cells)"""
    W, O = mods
    case, g, o, f, fp, (gi, gj) = _open_case(W, O, bg_scale=bg_scale)
    ids0 = g.get_particles()
    moved = 0
    for s in range(1, 5):
        g.wrf_pmc_trans_driver(fp, step=s)
        moved += o.wrf_pmc_trans_driver(f, step=s)
        _same(g, o, case)
    assert moved > 0
    # corners never change; an outflow boundary cell is a copy of its interior neighbour with new ids
    P = g.get_particles()
    off0 = np.concatenate([[0], np.cumsum(ids0["count"])])
    off = np.concatenate([[0], np.cumsum(P["count"])])
    nx, nz = case["nx"], case["nz"]
    corner = int(np.where((gi == 0) & (gj == 0))[0][0])
    assert np.array_equal(P["id"][off[corner]:off[corner + 1]], ids0["id"][off0[corner]:off0[corner + 1]])
    c = int(np.where((gi == nx - 1) & (gj == 2))[0][0])
    assert np.array_equal(P["vol"][:, off[c]:off[c + 1]], P["vol"][:, off[c - 1]:off[c]])
    assert not np.intersect1d(P["id"][off[c]:off[c + 1]], P["id"][off[c - 1]:off[c]]).size


def test_open_domain_step_and_transport(mods):
    """partmc_timestep (Brownian mc_coag + condense_particles + rebalance: interior cells only) alternating with the
    open-domain transport: boundary cells only change through the transport"""
    W, O = mods
    from wrf_partmc_b200.host import synthetic
    case, g, o, f, fp, (gi, gj) = _open_case(W, O, seed=6)
    g.condense_equilib_particles()
    for c in range(case["n_cells"]):
        o.condense_equilib(c)
    met = synthetic.met_sequence(case, 3)
    n_cells = case["n_cells"]
    bnd = (gi == 0) | (gi == case["nx"] - 1) | (gj == 0) | (gj == case["ny"] - 1)
    for k in range(3):
        T, P_, rrho = met[k]
        before = g.get_particles()
        g.partmc_timestep(temp=T, pressure=P_, inverse_density=rrho, elapsed_time=k * 60.0, dt=60.0, chem_dt=60.0,
                          do_coagulation=True, coag_kernel_type=W.KERNEL_BROWN, do_condensation=True, step=2 * k)
        o.partmc_timestep(0, n_cells, temp=T, pressure=P_, inverse_density=rrho, elapsed_time=k * 60.0, chem_dt=60.0,
                          do_coagulation=True, do_condensation=True, step=2 * k, n_threads=4)
        after = g.get_particles()
        assert np.array_equal(before["count"][bnd], after["count"][bnd])  # no coagulation in the boundary cells
        _same(g, o, case, water_rtol=1e-9)
        g.wrf_pmc_trans_driver(fp, step=2 * k + 1)
        o.wrf_pmc_trans_driver(f, step=2 * k + 1)
        _same(g, o, case, water_rtol=1e-9)


def test_open_domain_needs_boundary_conditions(mods):
    W, O = mods
    case = common.make_cells(n_cells=27, n_per_cell=20, n_spec=1, seed=1, nx=3, nz=3, ny=3)
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    with pytest.raises(W.PmcGpuError):  # periodic in one direction only: not supported (the reference treats it as open)
        W.PartMCGPU(data, nx=3, nz=3, ny=3, periodic_x=True, periodic_y=False)
    g = W.PartMCGPU(data, nx=3, nz=3, ny=3, periodic_x=False, periodic_y=False)
    with pytest.raises(W.PmcGpuError):
        g.wrf_pmc_trans_driver(dict(dx=100.0, dy=100.0, dt=1.0), step=1)
