"""The reference's own pin for Brownian coagulation: partmc/test/brownian compares the
particle-resolved run with the SECTIONAL solver run_sect.F90 -- 220 bins on D in [1e-10, 1e-4] m, 1e11 m^-3 log-normal
(D_g = 5e-8 m, log10 sigma = 0.24), T = 288 K, p = 1e5 Pa, 24 h, hourly output -- column by column (= output time by
output time) in the 2-norm: 0.2 for the number and 0.3 for the mass size distribution of the average of 9 repeats
with 1000 particles and del_t = 600 s (test_brownian_3.sh, test_brownian_4.sh).  The oracle restates run_sect
(oracle/pmc_oracle_sect.cpp) and its mc_coag has to pass the same comparison."""
import numpy as np
import pytest

import common

N_BIN, D_MIN, D_MAX = 220, 1e-10, 1e-4
T_MAX, T_OUT, DEL_T_PART, DEL_T_SECT = 86400.0, 3600.0, 600.0, 60.0


def sectional():
    from oracle import oracle as O
    return O.run_sect(N_BIN, D_MIN, D_MAX, 1000.0, 1e11, 5e-8, 0.24, O.KERNEL_BROWN, 288.0, 1e5, T_MAX, DEL_T_SECT, T_OUT)


def size_dists(vol, num_conc):
    """extract_aero_size --num / --mass: number and mass concentration per ln(D) bin"""
    edges = np.exp(np.linspace(np.log(D_MIN), np.log(D_MAX), N_BIN + 1))
    diam = (6.0 * vol / np.pi) ** (1.0 / 3.0)
    w = np.log(edges[1]) - np.log(edges[0])
    num = np.histogram(diam, edges, weights=num_conc)[0] / w
    mass = np.histogram(diam, edges, weights=num_conc * vol * 1000.0)[0] / w
    return num, mass


def rel_col_err(sect, part):
    """numeric_diff --by col: ||sect - part||_2 / ||sect||_2 per output time"""
    return np.sqrt(((sect - part) ** 2).sum(axis=1)) / np.sqrt((sect ** 2).sum(axis=1))


def brownian_case(n_repeat=9, n_part=1000, seed=7):
    return common.make_cells(n_repeat, n_part, n_spec=1, seed=seed, ragged=False, num_conc=1e11, r_g=2.5e-8,
                             log10_sigma=0.24, temp=288.0, pressure=1e5)


def test_sectional_conserves_mass_and_loses_number():
    num, mass = sectional()
    w = (np.log(D_MAX) - np.log(D_MIN)) / N_BIN
    m = mass.sum(axis=1) * w
    n = num.sum(axis=1) * w
    assert abs(n[0] / 1e11 - 1) < 1e-3                      # the log-normal mode integrates to its number
    assert np.abs(m / m[0] - 1).max() < 1e-6                # Bott's flux method conserves mass
    assert (np.diff(n) < 0).all() and 0.05 < n[-1] / n[0] < 0.3  # 1 / (1 + K N t / 2) = 0.19 for a monodisperse K = 1e-15


def test_oracle_mc_coag_brownian_against_sectional():
    from oracle import oracle as O
    case = brownian_case()
    case["T"][:] = 288.0
    case["P"][:] = 1e5
    case["wmag"][:] = 1e11 / 1000.0
    o = common.oracle_from_case(O, case)
    n_cells = case["n_cells"]
    num_s, mass_s = sectional()
    n_out = num_s.shape[0]
    num_p, mass_p = np.zeros_like(num_s), np.zeros_like(mass_s)

    def record(row):
        for c in range(n_cells):
            q = o.get_cell(c)
            nc = np.full(q["vol"].shape[1], o.get_weights(c)[1][0])
            a, b = size_dists(q["vol"][0], nc)
            num_p[row] += a / n_cells
            mass_p[row] += b / n_cells

    record(0)
    n_steps = int(T_MAX / DEL_T_PART)
    for k in range(1, n_steps + 1):
        for c in range(n_cells):
            o.sort(c, force_resort=True, physical=True, edge_recheck=True)
            o.mc_coag(c, O.KERNEL_BROWN, DEL_T_PART, step=k)
            o.rebalance(c, step=k)
        if k % int(T_OUT / DEL_T_PART) == 0:
            record(k // int(T_OUT / DEL_T_PART))
    e_num, e_mass = rel_col_err(num_s, num_p), rel_col_err(mass_s, mass_p)
    assert e_num.max() < 0.2, e_num       # test_brownian_3.sh
    assert e_mass.max() < 0.3, e_mass     # test_brownian_4.sh
    assert n_out == 25
