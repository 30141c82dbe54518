"""GPU parity of particle transport: pmcgpu_wrf_pmc_trans_driver (interface/wrf_pmc_trans.F90:37,
interface/wrf_pmc_trans_aero.F90:40) against the oracle's restatement on the same seeded periodic patch.
Bar: probabilities equal (same IEEE operations in the same order; 1e-14 relative allowed), transport
bookkeeping bit-exact given identical draws: which particles move where, the order they arrive in,
new ids, weight adjustments, rebalancing, and the bin grid the final aero_state_sort leaves behind."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def _setup(W, O, nx=6, nz=5, ny=4, n_per_cell=200, n_spec=4, seed=2, uniform=False, ragged=0.0, cap=None,
           courant=(0.3, 0.2, 0.05)):
    from wrf_partmc_b200.host import synthetic
    case = synthetic.make_patch(nx, nz, ny, n_per_cell, n_spec, seed=seed, ragged=ragged)
    if uniform:
        case["wmag"][:] = case["wmag"][0, 0, 0]
    g = common.gpu_from_case(W, case, cell_capacity=cap or (3 * n_per_cell + 64))
    o = common.oracle_from_case(O, case)
    g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
              cell_volume=case["cell_volume"])
    for c in range(case["n_cells"]):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
    o.trans_setup(nx, nz, ny, px=1, py=1)
    f = synthetic.transport_fields(case, o.num_conc_by_class(), courant_x=courant[0], courant_y=courant[1],
                                   courant_z=courant[2])
    fp = synthetic.patch_fields(f, 0, nx, 0, ny)
    return case, g, o, f, fp


def _same_everything(g, o, case):
    P = g.get_particles()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    t, m, e = g.get_weights()
    nb, gmin, gmax = g.grid()
    for c in range(case["n_cells"]):
        q = o.get_cell(c)
        a, b = off[c], off[c + 1]
        assert b - a == q["vol"].shape[1], "cell %d: %d on device, %d in oracle" % (c, b - a, q["vol"].shape[1])
        for key in ("id", "wclass", "wgroup", "n_primary", "hyst"):
            assert np.array_equal(P[key][a:b], q[key]), (c, key)
        assert np.array_equal(P["vol"][:, a:b], q["vol"]), c
        assert np.array_equal(P["lct"][a:b], q["lct"]) and np.array_equal(P["gct"][a:b], q["gct"])
        np.testing.assert_allclose(m[c].ravel(), o.get_weights(c)[1], rtol=1e-14)
        onb, omin, omax, _ = o.grid(c)
        assert nb[c] == onb
        if onb:
            np.testing.assert_allclose([gmin[c], gmax[c]], [omin, omax], rtol=1e-14)
    assert np.array_equal(g.get_next_id(), np.array([o.L.pmco_get_next_id(o.h, c) for c in range(case["n_cells"])]))


def test_probabilities_match_oracle(mods):
    W, O = mods
    case, g, o, f, fp = _setup(W, O)
    g.wrf_pmc_trans_driver(fp, step=1)
    o.trans_set_met(f)
    # the oracle's probabilities come from the state before transport: rebuild it
    case2, g2, o2, f2, fp2 = _setup(W, O)
    o2.trans_set_met(f2)
    o2.trans_probs(float(f2["dt"]))
    adv, diff, vert = g.trans_probs()
    for c in range(case["n_cells"]):
        for cl in range(case["n_class"]):
            a, d, v = o2.trans_get_probs(c, cl + 1)
            np.testing.assert_allclose(adv[c, cl], a, rtol=1e-14, atol=0)
            np.testing.assert_allclose(diff[c, cl], d, rtol=1e-14, atol=0)
            np.testing.assert_allclose(vert[c, cl], v, rtol=1e-14, atol=1e-300)


@pytest.mark.parametrize("kw", [dict(), dict(uniform=True), dict(nx=3, nz=4, ny=2, n_per_cell=500, ragged=0.5),
                                dict(nx=1, nz=3, ny=1, n_per_cell=300), dict(n_spec=20, nx=4, nz=6, ny=3),
                                # populations on both sides of half the capacity: both launches of the sampler
                                dict(nx=3, nz=4, ny=2, n_per_cell=500, ragged=0.5, cap=800)])
def test_trans_aero_bit_exact(mods, kw):
    W, O = mods
    case, g, o, f, fp = _setup(W, O, **kw)
    for step in range(1, 4):
        g.wrf_pmc_trans_driver(fp, step=step)
        moved = o.wrf_pmc_trans_driver(f, step=step)
        assert int(g.trans_moved().sum()) == moved
        _same_everything(g, o, case)


def test_preweight_scales_weights_when_a_cell_fills_up(mods):
    """a strongly convergent flow: the expected population of some cells more than doubles, so
    trans_aero_preweight calls aero_state_scale_weight (wrf_pmc_trans_aero.F90:1386-1392)"""
    W, O = mods
    case, g, o, f, fp = _setup(W, O, nx=4, nz=3, ny=4, n_per_cell=300, ragged=0.7, seed=9, cap=2000,
                               courant=(0.45, 0.4, 0.05))
    m0 = g.get_weights()[1].copy()
    g.wrf_pmc_trans_driver(fp, step=1)
    o.wrf_pmc_trans_driver(f, step=1)
    _same_everything(g, o, case)


def test_step_then_transport_sequence(mods):
    """partmc_timestep and wrf_pmc_trans_driver alternate like in module_integrate.F:405-424"""
    W, O = mods
    from wrf_partmc_b200.host import synthetic
    case, g, o, f, fp = _setup(W, O, nx=4, nz=4, ny=3, n_per_cell=300, n_spec=4)
    for c in range(case["n_cells"]):
        o.condense_equilib(c)
    g.condense_equilib_particles()
    met = synthetic.met_sequence(case, 2)
    n = case["n_cells"]
    for k, (T, P, rrho) in enumerate(met):
        g.partmc_timestep(temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * 60.0, chem_dt=60.0,
                          do_coagulation=True, do_condensation=True, step=2 * k)
        o.partmc_timestep(0, n, temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * 60.0, chem_dt=60.0,
                          do_coagulation=True, do_condensation=True, step=2 * k, n_threads=2)
        g.wrf_pmc_trans_driver(fp, step=2 * k + 1)
        o.wrf_pmc_trans_driver(f, step=2 * k + 1)
    P = g.get_particles()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    iw = case["sp"]["i_water"] - 1
    dry = [s for s in range(case["n_spec"]) if s != iw]
    for c in range(n):
        q = o.get_cell(c)
        a, b = off[c], off[c + 1]
        assert b - a == q["vol"].shape[1]
        assert np.array_equal(P["id"][a:b], q["id"])
        assert np.array_equal(P["vol"][dry, a:b], q["vol"][dry])
        np.testing.assert_allclose(P["vol"][iw, a:b], q["vol"][iw], rtol=1e-9, atol=1e-40)
