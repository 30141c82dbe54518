"""GPU parity tests for condensation: the device integrator against the CPU oracle
(1e-10 relative, north star) and against the reference's golden RH table."""
import json
import os

import numpy as np
import pytest

import common
from test_oracle_condense import _rates_inputs

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def test_condense_rates_device_vs_oracle(mods):
    """condense.F90:371-502 including the dry branch: 1e-10 relative."""
    W, O = mods
    g = W.PartMCGPU(W.AeroData([1000.0]))
    a = _rates_inputs(5000, seed=3)
    dev = g.eval_condense_rates(a)
    ref = O.condense_rates(a)[:, [0, 1, 3, 4, 5, 6]]
    scale = np.abs(ref).max(axis=0)
    rel = np.abs(dev - ref) / np.maximum(np.abs(ref), 1e-6 * scale)
    assert rel.max() < 1e-10, rel.max(axis=0)


def test_condense_equilib_device_vs_oracle(mods):
    W, O = mods
    case = common.condense_case(200, n_cells=3)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    rh = np.array([0.95, 0.8, 0.6])
    g.set_env(rel_humid=rh)
    g.condense_equilib_particles()
    for c in range(3):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=rh[c])
        o.condense_equilib(c)
    common.assert_same_population(g, o, case, exact_fp=False, rtol=1e-10)


@pytest.mark.parametrize("n_per_mode,n_cells,method", [(100, 4, 4), (200, 3, 4), (300, 2, 4), (350, 2, 4), (400, 3, 4), (540, 2, 4), (900, 2, 4), (100, 4, 5), (400, 3, 5), (900, 2, 5)])
def test_condense_particles_device_vs_oracle(mods, n_per_mode, n_cells, method):
    """condense_particles over several steps through activation: per-particle water volume, RH and
    the number of internal steps agree with the oracle (1e-10 relative on RH and on wet diameters).
    Covers the 1-, 4- (128 and 256 threads) and 8-particles-per-thread instantiations of RODAS4 (method 4, the default) and of the
    BDF integrator (method 5, the restatement of CVODE)."""
    W, O = mods
    case = common.condense_case(n_per_mode, n_cells=n_cells)
    g = common.gpu_from_case(W, case, cell_capacity=2 * n_per_mode + 64)
    g.set_condense_method(method)
    o = common.oracle_from_case(O, case)
    rh0 = np.linspace(0.95, 0.90, n_cells)
    g.set_env(rel_humid=rh0)
    g.condense_equilib_particles()
    for c in range(n_cells):
        o.set_env(c, T=290.0, p=1e5, RH=rh0[c])
        o.condense_equilib(c)
    for k in range(30):
        t1 = (k + 1) * 10.0
        T1 = np.full(n_cells, 290.0 - 10.0 * t1 / 1200.0)
        p1 = np.full(n_cells, 1e5 - 2e4 * t1 / 1200.0)
        rh, st, wre = g.condense_particles(T1, p1, 10.0)
        for c in range(n_cells):
            ost, owre = o.condense_particles(c, T1[c], p1[c], 10.0, step=k, method=method)
            # RODAS4 (the default): 1e-10 and the same accept / reject sequence.  BDF: a variable-order
            # code takes threshold decisions (Newton convergence, step and order changes) that rounding
            # flips now and then; two correct implementations then differ by a fraction of the local
            # error, so the bound is the integration tolerance, not rounding.
            tol = 1e-10 if method == 4 else 2e-7
            assert abs(rh[c] / o.env_rel_humid(c) - 1) < tol, (k, c, rh[c], o.env_rel_humid(c))
            if method == 4:
                assert (st[c, 0], st[c, 1]) == (ost[0], ost[1]), (k, c, st[c], ost)
            else:
                assert abs(int(st[c, 0]) - int(ost[0])) <= 0.2 * ost[0] + 3, (k, c, st[c], ost)
            assert abs(wre[c]) < 1e-6
    assert rh.max() > 1.0  # the case went through activation
    P = g.get_particles()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    for c in range(n_cells):
        q = o.get_cell(c)
        Dd = (P["vol"][:, off[c]:off[c + 1]].sum(axis=0)) ** (1 / 3)
        Do = (q["vol"].sum(axis=0)) ** (1 / 3)
        # RH (the integrator's cell-level output) is held to 1e-10 at every one of the 30 steps above; the wet
        # diameters of droplets that went through activation carry the rounding differences of the two
        # implementations (exp, reciprocals, summation order) amplified by the instability of activation
        # (> 1e5 for the largest droplets): 3e-10 there
        assert np.abs(Dd / Do - 1).max() < (3e-10 if method == 4 else 1e-6)
        assert np.array_equal(P["vol"][:19, off[c]:off[c + 1]], q["vol"][:19])  # solutes untouched


def test_condense_bdf_and_rodas4_agree_on_device(mods):
    """Two independent integrators (CVODE-style BDF, RODAS4) with the reference's tolerances end within a
    few 1e-8 of each other: the integration error is what the tolerance says it is."""
    W, O = mods
    res = {}
    for method in (5, 4):
        case = common.condense_case(400, n_cells=2)
        g = common.gpu_from_case(W, case)
        g.set_condense_method(method)
        g.set_env(rel_humid=np.array([0.95, 0.92]))
        g.condense_equilib_particles()
        for k in range(20):
            t1 = (k + 1) * 10.0
            rh, st, wre = g.condense_particles(np.full(2, 290.0 - 10.0 * t1 / 1200.0),
                                               np.full(2, 1e5 - 2e4 * t1 / 1200.0), 10.0)
        res[method] = (rh, g.get_particles()["vol"].sum(axis=0) ** (1 / 3), st)
    assert np.abs(res[5][0] / res[4][0] - 1).max() < 1e-7
    assert np.abs(res[5][1] / res[4][1] - 1).max() < 1e-6
    assert res[5][2][:, 2].sum() < res[4][2][:, 2].sum()  # BDF needs fewer rate evaluations


def test_condense_golden_on_device(mods):
    """K3 scenario on the GPU against ref_condense_0001_env.txt (RH 1e-4) and water mass (0.05)."""
    W, O = mods
    gold = json.load(open(os.path.join(HERE, "golden", "condense_env_golden.json")))
    case = common.condense_case(500)
    g = common.gpu_from_case(W, case)
    g.set_env(rel_humid=np.array([0.95]))
    g.condense_equilib_particles()
    _, m, _ = g.get_weights()

    def water_mass():
        P = g.get_particles()
        nc = np.where(P["wclass"] == 1, m[0, 0, 0], m[0, 1, 0])
        return float((P["vol"][19] * nc).sum() * 1000.0)
    rh, water = [0.95], [water_mass()]
    for k in range(60):
        t1 = (k + 1) * 10.0
        r, st, wre = g.condense_particles([290.0 - 10.0 * t1 / 1200.0], [1e5 - 2e4 * t1 / 1200.0], 10.0)
        rh.append(r[0])
        water.append(water_mass())
        assert abs(wre[0]) < 1e-6
    gr = np.array(gold["rel_humid"])
    assert np.linalg.norm(np.array(rh) - gr) / np.linalg.norm(gr) < 1e-4
    gw = np.array(gold["water_mass_conc"])
    assert np.linalg.norm(np.array(water) - gw) / np.linalg.norm(gw) < 0.05


@pytest.mark.parametrize("n_per_mode,steps", [(1500, 12), (5000, 6), (50000, 2)])
def test_condense_single_box_cells_device_vs_oracle(mods, n_per_mode, steps):
    """Cells above 2048 particles (single-box runs; K3 = 1e5 particles in one cell) take the grid-wide
    variant of the integrator (one cell spread over all SMs, cell sums through a grid barrier): RH and
    wet diameters against the oracle to 1e-10, same accept / reject sequence, water balance < 1e-6."""
    W, O = mods
    case = common.condense_case(n_per_mode, n_cells=2)
    case["counts"][1] = 700  # a small cell beside it: both kernel families in one call
    keep = np.concatenate([np.arange(2 * n_per_mode), 2 * n_per_mode + np.arange(700)])
    for k in ("wclass", "wgroup", "id", "n_primary", "hyst", "lct", "gct"):
        case[k] = case[k][keep]
    case["vol"] = case["vol"][:, keep]
    g = common.gpu_from_case(W, case, cell_capacity=2 * n_per_mode + 64)
    o = common.oracle_from_case(O, case)
    rh0 = np.array([0.95, 0.93])
    g.set_env(rel_humid=rh0)
    g.condense_equilib_particles()
    for c in range(2):
        o.set_env(c, T=290.0, p=1e5, RH=rh0[c])
        o.condense_equilib(c)
    for k in range(steps):
        t1 = (k + 1) * 10.0
        T1 = np.full(2, 290.0 - 10.0 * t1 / 1200.0)
        p1 = np.full(2, 1e5 - 2e4 * t1 / 1200.0)
        rh, st, wre = g.condense_particles(T1, p1, 10.0)
        for c in range(2):
            ost, owre = o.condense_particles(c, T1[c], p1[c], 10.0, step=k, method=4)
            assert abs(rh[c] / o.env_rel_humid(c) - 1) < 1e-10, (k, c, rh[c], o.env_rel_humid(c))
            assert (st[c, 0], st[c, 1]) == (ost[0], ost[1]), (k, c, st[c], ost)
            assert abs(wre[c]) < 1e-6
    P = g.get_particles()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    for c in range(2):
        q = o.get_cell(c)
        Dd = (P["vol"][:, off[c]:off[c + 1]].sum(axis=0)) ** (1 / 3)
        Do = (q["vol"].sum(axis=0)) ** (1 / 3)
        assert np.abs(Dd / Do - 1).max() < 1e-10


@pytest.mark.parametrize("n_per,n_cells", [(400, 3), (3000, 1)])
def test_condense_reweight_size_dependent_weights(mods, n_per, n_cells):
    """aero_state_reweight (aero_state.F90:597-653) after condense_equilib_particles and
    condense_particles with nummass-like weights (group 2: power -3): magnitudes rescaled, particles
    duplicated / removed by prob_round(old_nc / new_nc) backwards through the array.  The draws are
    addressed, so the device takes the same decisions as the oracle: same survivors, copies and ids in the
    same order (volumes agree to the integrator's 1e-10; a ratio within 1e-10 of a draw could flip a
    decision, which this seed does not hit)."""
    W, O = mods
    case = common.make_cells(n_cells, n_per, n_spec=3, n_group=2, weight="nummass", seed=2, ragged=False,
                             temp=285.0, num_conc=3e9, r_g=4e-8)
    case["T"][:] = 290.0
    case["P"][:] = 1e5
    g = common.gpu_from_case(W, case, cell_capacity=3 * n_per)
    o = common.oracle_from_case(O, case)
    rh0 = np.full(n_cells, 0.95)
    g.set_env(rel_humid=rh0)
    g.condense_equilib_particles(step=77)
    for c in range(n_cells):
        o.set_env(c, T=290.0, p=1e5, RH=rh0[c])
        o.condense_equilib(c, step=77)
    n_after_equilib = g.counts().copy()
    assert not np.array_equal(n_after_equilib, case["counts"]), "equilibration must change some particle counts"
    common.assert_same_population(g, o, case, exact_fp=False, rtol=1e-9)
    for k in range(6):
        t1 = (k + 1) * 10.0
        T1 = np.full(n_cells, 290.0 - 10.0 * t1 / 1200.0)
        p1 = np.full(n_cells, 1e5 - 2e4 * t1 / 1200.0)
        rh, st, wre = g.condense_particles(T1, p1, 10.0, step=k)
        for c in range(n_cells):
            o.condense_particles(c, T1[c], p1[c], 10.0, step=k, method=4)
            assert abs(rh[c] / o.env_rel_humid(c) - 1) < 1e-10
        common.assert_same_population(g, o, case, exact_fp=False, rtol=1e-9)
    assert not np.array_equal(g.counts(), n_after_equilib), "condensation must change some particle counts"
