"""CPU tests of the condensation restatement: condense_rates partials against finite
differences, the integrator's order, and the reference's golden RH table."""
import json
import os

import numpy as np
import pytest

import common
from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))


def _rates_inputs(n, seed=0):
    rng = np.random.default_rng(seed)
    T = rng.uniform(270, 300, n)
    p = rng.uniform(7e4, 1e5, n)
    H = rng.uniform(0.6, 1.01, n)
    D_dry = 10 ** rng.uniform(-8, -6.3, n)
    D = D_dry * rng.uniform(1.02, 3.0, n)
    D[: n // 10] = D_dry[: n // 10] * 0.9  # dry branch (condense.F90:420-447)
    kappa = rng.uniform(0.05, 0.9, n)
    V_comp = 1.0 / 10 ** rng.uniform(5, 8, n)
    Tdot = rng.uniform(-0.02, 0.02, n)
    pdot = rng.uniform(-30, 30, n)
    return np.stack([T, Tdot, H, p, pdot, V_comp, D, D_dry, kappa], axis=1)


def test_condense_rates_partials_match_finite_differences():
    """condense.F90:371-502: dDdot/dD, dDdot/dH, dHdoti/dD, dHdoti/dH, dHdotenv/dH."""
    a = _rates_inputs(400)
    a = a[a[:, 6] > a[:, 7]]  # wet branch only: the dry branch is not differentiable at D = D_dry
    o0 = O.condense_rates(a)
    for col, jd, jh, je in ((6, 3, 5, 7), (2, 4, 6, 8)):
        ap, am = a.copy(), a.copy()
        dx = 1e-6 * a[:, col]
        ap[:, col] += dx
        am[:, col] -= dx
        op, om = O.condense_rates(ap), O.condense_rates(am)
        dD = (op[:, 0] - om[:, 0]) / (2 * dx)
        dHi = (op[:, 1] - om[:, 1]) / (2 * dx)
        dHe = (op[:, 2] - om[:, 2]) / (2 * dx)
        np.testing.assert_allclose(dD, o0[:, jd], rtol=2e-5, atol=1e-12 * np.abs(o0[:, jd]).max())
        np.testing.assert_allclose(dHi, o0[:, jh], rtol=2e-5, atol=1e-12 * np.abs(o0[:, jh]).max())
        np.testing.assert_allclose(dHe, o0[:, je], rtol=1e-6, atol=1e-14)


def test_condense_rates_equilibrium_is_kohler():
    """At Ddot = 0 the rate function reduces to kappa-Koehler: H = a_w exp(X / D)."""
    T, p = 285.0, 9e4
    D_dry, kappa = 8e-8, 0.4
    X = 4 * 18e-3 * 0.073 / (8.314472 * T * 1e3)
    for D in (1.2e-7, 2e-7, 6e-7):
        a_w = (D ** 3 - D_dry ** 3) / (D ** 3 + (kappa - 1) * D_dry ** 3)
        H = a_w * np.exp(X / D)
        out = O.condense_rates([[T, 0.0, H, p, 0.0, 1e-6, D, D_dry, kappa]])
        assert abs(out[0, 0]) < 1e-6 * 1e-9  # Ddot ~ 0 compared with typical 1e-9 m/s growth rates


def test_integrator_is_fourth_order_and_tolerance_controlled():
    case = common.condense_case(40)
    vol = case["vol"]
    D = 2 * (3 * vol.sum(axis=0) / 4 / np.pi) ** (1 / 3)
    solv = vol[:19].sum(axis=0)
    D_dry = 2 * (3 * solv / 4 / np.pi) ** (1 / 3)
    kappa = np.full(len(D), 0.3)
    nc = np.full(len(D), 1e8)
    D0 = D_dry * 1.5
    ref = O.condense_integrate(D0, 0.97, D_dry, kappa, nc, 290.0, 1e5, -8e-3, -16.0, 5.0, reltol=1e-12,
                               abstol_scale=1e-12, abstol_H=1e-14)
    assert ref[0] == 0
    errs = []
    for tol in (1e-5, 1e-7, 1e-9):
        rc, Dn, Hn, st = O.condense_integrate(D0, 0.97, D_dry, kappa, nc, 290.0, 1e5, -8e-3, -16.0, 5.0, reltol=tol,
                                              abstol_scale=tol, abstol_H=tol * 1e-2)
        assert rc == 0
        errs.append(max(np.abs(Dn / ref[1] - 1).max(), abs(Hn / ref[2] - 1)))
        assert errs[-1] < 50 * tol
    assert errs[0] > errs[1] > errs[2]


def test_condense_golden_rel_humid():
    """partmc/test/condense/test_condense_1.sh: RH(t) against ref_condense_0001_env.txt, column
    2-norm relative tolerance 1e-4 (numeric_diff --by col); test_condense_2.sh: water mass, 0.05."""
    gold = json.load(open(os.path.join(HERE, "golden", "condense_env_golden.json")))
    case = common.condense_case(300)
    o = common.oracle_from_case(O, case)
    o.set_env(0, T=290.0, p=1e5, RH=0.95)
    o.condense_equilib(0)
    rh, water = [0.95], []
    _, m, _ = o.get_weights(0)

    def water_mass():
        P = o.get_cell(0)
        nc = np.where(P["wclass"] == 1, m[0], m[1])
        return float((P["vol"][19] * nc).sum() * 1000.0)
    water.append(water_mass())
    wre_max = 0.0
    for k in range(60):
        t1 = (k + 1) * 10.0
        st, wre = o.condense_particles(0, 290.0 - 10.0 * t1 / 1200.0, 1e5 - 2e4 * t1 / 1200.0, 10.0, step=k)
        wre_max = max(wre_max, abs(wre))
        rh.append(o.env_rel_humid(0))
        water.append(water_mass())
    g = np.array(gold["rel_humid"])
    err = np.linalg.norm(np.array(rh) - g) / np.linalg.norm(g)
    assert err < 1e-4, err
    gw = np.array(gold["water_mass_conc"])
    errw = np.linalg.norm(np.array(water) - gw) / np.linalg.norm(gw)
    assert errw < 0.05, errw
    assert wre_max < 1e-6  # condense.F90:317 water-balance bound


def test_empty_cell_relative_humidity_closed_form():
    """A cell without particles: dH/dt = H (-dlnP0/dT Tdot + pdot / p) (condense.F90:446-449, Hdot_env) has the closed
    form H0 exp(int lambda dt); the integrator (with the d/dt term of the non-autonomous Rosenbrock stages) must
    reproduce it far below the tolerance, in a handful of steps."""
    from scipy.integrate import quad
    from oracle import oracle as O
    T0, p0, Td, pd = 288.0, 1e5, -0.2 / 600.0, -200.0 / 600.0

    def lam(t):
        T, p = T0 + t * Td, p0 + t * pd
        return -7.45 * np.log(10.0) * (273.15 - 38.0) / (T - 38.0) ** 2 * Td + pd / p

    exact = 0.5 * np.exp(quad(lam, 0.0, 600.0, epsabs=1e-16, epsrel=1e-13)[0])
    e = np.zeros(0)
    rc, _, H, st = O.condense_integrate(e, 0.5, e, e, e, T0, p0, Td, pd, 600.0)
    assert rc == 0 and abs(H / exact - 1.0) < 1e-12 and st[0] <= 6, (H, exact, st)
