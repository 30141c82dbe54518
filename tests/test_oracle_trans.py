"""CPU tests of the transport restatement (oracle/pmc_oracle_trans.cpp) against the invariants the
reference asserts (normalize_probs, interface/wrf_pmc_trans.F90:702-771), against conservation, and
against the reference's own sort.c (oracle/_ref) for the ordering of the move probabilities."""
import ctypes as C
import os

import numpy as np
import pytest

import common
from oracle import oracle as O
from wrf_partmc_b200.host import synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(nx=6, nz=5, ny=4, n_per_cell=200, px=1, py=1, seed=2, uniform=False, n_spec=4):
    case = synthetic.make_patch(nx, nz, ny, n_per_cell, n_spec, seed=seed)
    if uniform:
        case["wmag"][:] = case["wmag"][0, 0, 0]
    o = common.oracle_from_case(O, case)
    for c in range(case["n_cells"]):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
    o.trans_setup(nx, nz, ny, px=px, py=py)
    f = synthetic.transport_fields(case, o.num_conc_by_class())
    return case, o, f


def test_probabilities_are_valid_and_follow_the_fluxes():
    case, o, f = build()
    o.trans_set_met(f)
    o.trans_probs(float(f["dt"]))
    nx, nz, ny = case["nx"], case["nz"], case["ny"]
    for g in range(0, case["n_cells"], 7):
        k = (g // nx) % nz + 1
        for cl in (1, 2):
            adv, diff, vert = o.trans_get_probs(g, cl)
            assert (adv >= 0).all() and (diff >= 0).all() and (vert >= 0).all()
            assert (adv <= 1).all() and (diff <= 1).all() and (vert <= 1).all()
            off = np.delete(vert, k - 1).sum()
            assert adv[:4].sum() + diff.sum() + off < 1.0 + 1e-8
            # the row of B* A is a probability vector (before normalisation it sums to one)
            assert abs(vert.sum() - 1.0) < 1e-9 or adv[:4].sum() + diff.sum() + off >= 1.0 - 1e-9
            # upwind fluxes of the cell's own number concentration: p = Courant number of the outflow face
            assert adv[0] * adv[1] == 0.0 and adv[2] * adv[3] == 0.0
    # horizontal diffusion: dt K / dx^2 with K ~ 10..13 m2/s, dx = 100 m, dt = 6 s
    adv, diff, vert = o.trans_get_probs(0, 1)
    assert 0.005 < diff[0] < 0.009


def test_probability_order_matches_reference_sort_c():
    """aero_state_multisample sorts the move probabilities with real_sort -> sort.c double_sort_c
    (qsort); the oracle uses a stable ascending sort.  Same order whenever there are no ties."""
    path = os.path.join(ROOT, "oracle", "_ref", "libpmcref_sort.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    ref = C.CDLL(path)
    rng = np.random.default_rng(0)
    for n in (1, 3, 8, 23, 44):
        x = rng.random(n)
        data = x.copy()
        perm = np.zeros(n, np.int32)
        ref.double_sort_c(C.c_int(n), data.ctypes.data_as(C.c_void_p), perm.ctypes.data_as(C.c_void_p))
        mine = np.argsort(x, kind="stable")
        assert np.array_equal(perm - 1, mine)
        assert np.array_equal(data, x[mine])


def test_moves_conserve_particles_when_weights_are_uniform():
    """equal magnitudes and volumes: every transfer is a move (ratio == 1, binomial count)"""
    case, o, f = build(uniform=True)
    ids0 = np.sort(np.concatenate([o.get_cell(c)["id"] for c in range(case["n_cells"])]))
    moved = o.wrf_pmc_trans_driver(f, step=1, allow_doubling=False, allow_halving=False)
    ids1 = np.sort(np.concatenate([o.get_cell(c)["id"] for c in range(case["n_cells"])]))
    assert moved > 0.2 * len(ids0)
    assert np.array_equal(ids0, ids1)
    # mean horizontal displacement follows the Courant numbers: |cx| <= 0.3 (1 + 0.2), |cy| <= 0.2
    assert moved < 0.8 * len(ids0)


def test_number_concentration_is_conserved_in_expectation():
    case, o, f = build(nx=8, nz=4, ny=6, n_per_cell=400, seed=4)
    def total():
        return float(o.num_conc_by_class().astype(np.float64).sum())
    # equal cell volumes: sum over cells of count * magnitude * alt is a proxy of total number
    t0 = total()
    vals = []
    for step in range(1, 4):
        o.wrf_pmc_trans_driver(f, step=step)
        vals.append(total())
    assert abs(vals[-1] - t0) / t0 < 0.02, (t0, vals)


def test_decomposition_changes_only_the_arrival_order():
    """the same particles move to the same cells for any px x py (draws are addressed by the source
    cell); only the order inside the receiving cell follows the patch layout"""
    res = []
    for (px, py) in ((1, 1), (2, 2), (3, 1)):
        case, o, f = build(px=px, py=py, uniform=True)
        o.wrf_pmc_trans_driver(f, step=1, allow_doubling=False, allow_halving=False)
        res.append([np.sort(o.get_cell(c)["id"]) for c in range(case["n_cells"])])
    for other in res[1:]:
        for a, b in zip(res[0], other):
            assert np.array_equal(a, b)


def test_transport_is_deterministic_and_threads_do_not_matter():
    out = []
    for nt in (1, 4):
        case, o, f = build(px=2, py=2)
        o.wrf_pmc_trans_driver(f, step=3, n_threads=nt)
        out.append([o.get_cell(c) for c in range(case["n_cells"])])
    for a, b in zip(*out):
        for k in a:
            assert np.array_equal(a[k], b[k])
