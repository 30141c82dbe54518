"""The sedimentation coagulation kernel (partmc/src/coag_kernel_sedi.F90), pinned the way the reference pins it:
  * test/bidisperse: 10000 droplets of r = 10 um and ONE of r = 100 um, kernel sedi, del_t = 1 s, 600 s, no doubling;
    the number concentration of the small droplets and the mass concentration of the big one against the one-variable
    ODE  d n_small / dt = -k(v_small, v_big(n_small)) n_small / comp_vol  (test_bidisperse_ode.F90, RK4), column by
    column in the 2-norm at 0.3 (test_bidisperse_1.sh);
  * test/sedi: mass size distribution of the particle run (1000 particles, log-normal D_g = 20 um, log10 sigma = 0.2,
    1e9 m^-3, del_t = 5 s, doubling on) against run_sect (100 bins on D in [1e-7, 1] m, del_t = 1 s) after 300 and
    600 s at 0.7 (test_sedi_2.sh)."""
import numpy as np
import pytest

import common

R_SMALL, R_BIG, N_SMALL, DENSITY = 1e-5, 1e-4, 10000, 1000.0
NUM_CONC_SMALL = 1e9


def v_of(r):
    return 4.0 * np.pi * r ** 3 / 3.0


def bidisperse_ode(t_out, del_t=0.01):
    """test_bidisperse_ode.F90: classical RK4 on n_small (the reference steps with 1e-3 s; RK4 at 1e-2 s agrees to 1e-9)"""
    from oracle import oracle as O
    v_small, v_big_init = v_of(R_SMALL), v_of(R_BIG)
    num_conc = NUM_CONC_SMALL * (N_SMALL + 1.0) / N_SMALL
    comp_vol = (N_SMALL + 1.0) / num_conc

    def f(n):
        return -O.kernel_sedi(v_small, v_big_init + (N_SMALL - n) * v_small) / comp_vol * n

    n = float(N_SMALL)
    rows = [(0.0, n / comp_vol, v_big_init * DENSITY / comp_vol)]
    n_step = int(round(t_out[-1] / del_t))
    every = int(round((t_out[1] - t_out[0]) / del_t))
    for i in range(1, n_step + 1):
        k1 = del_t * f(n)
        k2 = del_t * f(n + k1 / 2.0)
        k3 = del_t * f(n + k2 / 2.0)
        k4 = del_t * f(n + k3)
        n = n + k1 / 6.0 + k2 / 3.0 + k3 / 3.0 + k4 / 6.0
        if i % every == 0:
            rows.append((i * del_t, n / comp_vol, (v_big_init + (N_SMALL - n) * v_small) * DENSITY / comp_vol))
    return np.array(rows), comp_vol


def bidisperse_case(seed=5):
    case = common.make_cells(1, N_SMALL + 1, n_spec=1, seed=seed, ragged=False)
    vol = np.full((1, N_SMALL + 1), v_of(R_SMALL))
    vol[0, -1] = v_of(R_BIG)
    case["vol"] = vol
    num_conc = NUM_CONC_SMALL * (N_SMALL + 1.0) / N_SMALL
    case["wmag"][:] = num_conc / (N_SMALL + 1.0)  # 1 / comp_vol
    return case


def bidisperse_columns(vol, comp_vol):
    """test_bidisperse_extract.F90: particles below / above r = 5e-5 m"""
    r = (3.0 * vol / (4.0 * np.pi)) ** (1.0 / 3.0)
    small = r < 5e-5
    return small.sum() / comp_vol, (vol[~small] * DENSITY).sum() / comp_vol


def rel_col_err(ref, got):
    return np.sqrt(((ref - got) ** 2).sum(axis=0)) / np.sqrt((ref ** 2).sum(axis=0))


def test_sedi_kernel_values():
    """limits that follow from the formula: equal droplets do not collide; symmetric; the Stokes regime of fall_g;
    the efficiency table's corner values (Hall 1980: E(300 um, ratio 1) = 1 -> capped, E(<= 6 um collector) row 1)"""
    from oracle import oracle as O
    assert O.kernel_sedi(v_of(7e-6), v_of(7e-6)) == 0.0
    a, b = O.kernel_sedi(v_of(3e-5), v_of(8e-6)), O.kernel_sedi(v_of(8e-6), v_of(3e-5))
    assert a == b and a > 0.0
    # r1 = 4 um, r2 = 2 um: both in the Stokes-Cunningham branch, ir = 1 -> E interpolated in the first row only
    r1, r2 = 4e-6, 2e-6
    eta, lam, rhow, rhoa, g = 1.818e-4, 6.62e-6, 1.0, 1.225e-3, 980.665
    w = lambda r: 2.0 * g * (rhow - rhoa) / (9.0 * eta) * ((r * 100) ** 2 + 1.257 * lam * (r * 100)) / 100.0
    k = O.kernel_sedi(v_of(r1), v_of(r2))
    ec = k / (np.pi * (r1 + r2) ** 2 * abs(w(r1) - w(r2)))
    assert ec == pytest.approx(0.04, rel=1e-9), ec  # ecoll(1, 11): collector <= 6 um, ratio 0.5
    # a 1 mm drop collecting 10 um droplets: E ~ 1 (capped), fall speed saturates above 3.5 mm radius
    k_big = O.kernel_sedi(v_of(1e-3), v_of(1e-5))
    assert 1e-6 < k_big < 1e-4
    # same radius ratio (same efficiency), both collectors above 3.5 mm: only the cross-section changes
    assert O.kernel_sedi(v_of(5e-3), v_of(1.25e-5)) / O.kernel_sedi(v_of(4e-3), v_of(1e-5)) == pytest.approx(1.25 ** 2, rel=1e-3)


def test_oracle_bidisperse_against_ode():
    from oracle import oracle as O
    t_out = np.arange(0.0, 601.0, 10.0)
    ode, comp_vol = bidisperse_ode(t_out)
    case = bidisperse_case()
    o = common.oracle_from_case(O, case)
    rows = [(0.0,) + bidisperse_columns(o.get_cell(0)["vol"][0], comp_vol)]
    for k in range(1, 601):
        o.sort(0, force_resort=True, physical=True, edge_recheck=True)
        o.mc_coag(0, O.KERNEL_SEDI, 1.0, step=k)
        if k % 10 == 0:
            rows.append((float(k),) + bidisperse_columns(o.get_cell(0)["vol"][0], comp_vol))
    part = np.array(rows)
    err = rel_col_err(ode[:, 1:], part[:, 1:])
    assert err.max() < 0.3, err  # test_bidisperse_1.sh
    assert part[-1, 1] < 0.5 * part[0, 1]  # the big drop has swept up most of the small ones
    # total volume is conserved by coagulation
    assert o.get_cell(0)["vol"].sum() == pytest.approx(case["vol"].sum(), rel=1e-12)


def nummass_oracle(O, seed, n_part=1000.0):
    """do_select_weighting no -> AERO_STATE_WEIGHT_NUMMASS (spec_file / aero_state_set_weight, aero_state.F90:440-470):
    two weight groups, flat and power -3, each sampling the whole initial distribution
    (aero_state_add_aero_dist_sample, aero_state.F90:758-831, which also sets the magnitudes)"""
    sp = common.species(1)
    data = O.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    o = O.Oracle(1, data, n_group=2, n_class=1, n_part_ideal=n_part, seed=seed)
    o.set_weights(0, np.array([[1, 2]], np.int32), np.array([[1.0, 1.0]]), np.array([[0.0, -3.0]]))
    o.set_env(0, T=288.0, p=1e5)
    n = o.add_aero_dist_sample(0, [1], [1], [1e-5], [0.2], [[1.0]], [1e9], 1.0, 1.0)
    assert 800 < n < 1200
    return o


def harmonic_num_conc(o, radius):
    wt, wm, we = [np.asarray(a).ravel() for a in o.get_weights(0)]
    inv = sum(1.0 / (wm[g] * (radius ** we[g] if wt[g] == 2 else 1.0)) for g in range(wt.size))
    return 1.0 / inv


@pytest.mark.parametrize("seed", [1, 2])
def test_oracle_sedi_against_sectional(seed):
    from oracle import oracle as O
    n_bin, d_min, d_max = 100, 1e-7, 1.0
    num_s, mass_s = O.run_sect(n_bin, d_min, d_max, 1000.0, 1e9, 2e-5, 0.2, O.KERNEL_SEDI, 288.0, 1e5, 600.0, 1.0, 300.0)
    assert mass_s.shape[0] == 3
    o = nummass_oracle(O, seed)
    edges = np.exp(np.linspace(np.log(d_min), np.log(d_max), n_bin + 1))
    w = np.log(edges[1]) - np.log(edges[0])
    num_p, mass_p = np.zeros_like(num_s), np.zeros_like(mass_s)

    def record(row):
        q = o.get_cell(0)
        radius = (3.0 * q["vol"][0] / (4.0 * np.pi)) ** (1.0 / 3.0)
        nc = harmonic_num_conc(o, radius)
        num_p[row] = np.histogram(2.0 * radius, edges, weights=nc)[0] / w
        mass_p[row] = np.histogram(2.0 * radius, edges, weights=nc * q["vol"][0] * 1000.0)[0] / w

    record(0)
    for k in range(1, 121):
        o.sort(0, force_resort=True, physical=True, edge_recheck=True)
        o.mc_coag(0, O.KERNEL_SEDI, 5.0, step=k)
        o.rebalance(0, True, True, step=k)
        if k % 60 == 0:
            record(k // 60)
    err = lambda a, b: np.sqrt(((a - b) ** 2).sum(axis=1)) / np.sqrt((a ** 2).sum(axis=1))
    assert err(num_s, num_p).max() < 0.7, err(num_s, num_p)      # test_sedi_1.sh
    assert err(mass_s, mass_p).max() < 0.7, err(mass_s, mass_p)  # test_sedi_2.sh
