"""Ensemble statistics (north star: "stochastic outputs must match the reference CPU PartMC's ensemble
statistics ... within 95 % confidence intervals over >= 16 seeds").

The reference draws from ONE MT19937 stream in program order (rand_gsl.c:46-58); the device (and the oracle's
Philox back end it is compared with bit for bit elsewhere) addresses every draw by (cell, step, phase, ...).
Both are samplers of the same stochastic process, so 16-member ensembles of each must agree in the mean of
every population statistic within the 95 % confidence interval of the difference (Welch's t, 30 dof).
The reference-order ensemble is the oracle with the MT19937 back end: same code path order as coagulation.F90,
one sequential stream per member.

Statistics per member after the run: number concentration, mass concentration, second and third volume
moments, number concentration in four fixed size classes."""
import numpy as np
import pytest

import common

N_MEMBERS = 16
T95 = 2.042  # Student t, 97.5 %, 30 degrees of freedom


def _population(n=2000, seed=3):
    case = common.make_cells(1, n, n_spec=3, n_class=1, seed=seed, num_conc=2e11, ragged=False)
    return case


def _stats(vol, mag, edges):
    """vol [S][n] of one member, flat weight magnitude mag"""
    v = vol.sum(axis=0)
    r = (3.0 * v / (4.0 * np.pi)) ** (1.0 / 3.0)
    out = [v.size * mag, v.sum() * mag, (v ** 2).sum() * mag, (v ** 3).sum() * mag]
    out += list(np.histogram(r, bins=edges)[0] * mag)
    return np.array(out)


def _edges(case):
    v = case["vol"].sum(axis=0)
    r = (3.0 * v / (4.0 * np.pi)) ** (1.0 / 3.0)
    q = np.quantile(r, [0.25, 0.5, 0.75])
    return np.concatenate([[0.0], q, [np.inf]])


def _oracle_member(O, case, rng, seed, steps, dt, rebalance):
    o = common.oracle_from_case(O, dict(case, seed=seed), rng=rng)
    for step in range(steps):
        o.sort(0, force_resort=True, physical=True, edge_recheck=True)
        o.mc_coag(0, O.KERNEL_BROWN, dt, step=step)
        if rebalance:
            o.rebalance(0, True, True, step=step)
    q = o.get_cell(0)
    return q["vol"], o.get_weights(0)[1][0]


def _compare(a, b, what):
    """a, b: [members][n_stat]; every mean inside the 95 % CI of the difference"""
    ma, mb = a.mean(axis=0), b.mean(axis=0)
    se = np.sqrt(a.var(axis=0, ddof=1) / a.shape[0] + b.var(axis=0, ddof=1) / b.shape[0])
    # mass is conserved exactly by both: zero variance, equal means
    tol = T95 * se + 1e-12 * np.abs(mb)
    bad = np.abs(ma - mb) > tol
    assert not bad.any(), "%s: statistics %s outside the 95 %% CI: %s vs %s (half width %s)" % (
        what, np.nonzero(bad)[0], ma[bad], mb[bad], tol[bad])
    return np.abs(ma - mb) / np.maximum(tol, 1e-300)


STEPS, DT = 6, 1200.0


def _reference_order_ensemble(O, case, edges):
    return np.array([_stats(*_oracle_member(O, case, "mt", 100 + m, STEPS, DT, False), edges) for m in range(N_MEMBERS)])


def test_addressed_draws_match_reference_order_ensemble():
    """CPU only: the oracle with addressed Philox draws (what the device reproduces bit for bit) against the
    oracle with the reference's single sequential MT19937 stream, 16 members each."""
    from oracle import oracle as O
    case = _population()
    edges = _edges(case)
    ref = _reference_order_ensemble(O, case, edges)
    phi = np.array([_stats(*_oracle_member(O, case, "philox", 200 + m, STEPS, DT, False), edges) for m in range(N_MEMBERS)])
    z = _compare(phi, ref, "philox vs mt19937")
    assert ref[:, 0].mean() < 0.8 * case["counts"][0] * case["wmag"][0, 0, 0], "the run must coagulate substantially"
    assert z.max() <= 1.0


@pytest.mark.gpu
def test_device_ensemble_matches_reference_order_ensemble():
    """16 members on the device (16 cells with the same initial population: the cell number is part of every
    draw's address) against 16 reference-order (MT19937) members of the CPU oracle."""
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    case1 = _population()
    edges = _edges(case1)
    ref = _reference_order_ensemble(O, case1, edges)
    n = int(case1["counts"][0])
    case = dict(case1, n_cells=N_MEMBERS, nx=N_MEMBERS, counts=np.full(N_MEMBERS, n, np.int32))
    for k in ("wclass", "wgroup", "n_primary", "hyst", "lct", "gct"):
        case[k] = np.tile(case1[k], N_MEMBERS)
    case["id"] = np.arange(1, N_MEMBERS * n + 1, dtype=np.int64)
    case["vol"] = np.tile(case1["vol"], (1, N_MEMBERS))
    for k in ("wtype", "wmag", "wexp"):
        case[k] = np.repeat(case1[k], N_MEMBERS, axis=0)
    case["T"] = np.full(N_MEMBERS, case1["T"][0])
    case["P"] = np.full(N_MEMBERS, case1["P"][0])
    g = common.gpu_from_case(W, case)
    for step in range(STEPS):
        g.mc_coag(W.KERNEL_BROWN, DT, step=step)
    P = g.get_particles()
    _, mag, _ = g.get_weights()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    dev = np.array([_stats(P["vol"][:, off[c]:off[c + 1]], mag[c, 0, 0], edges) for c in range(N_MEMBERS)])
    _compare(dev, ref, "device vs mt19937")
    # total mass: conserved to rounding in every member (flat weights, no rebalance)
    m0 = case1["vol"].sum() * case1["wmag"][0, 0, 0]
    assert np.abs(dev[:, 1] / m0 - 1).max() < 1e-12
