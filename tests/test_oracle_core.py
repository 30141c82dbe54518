"""CPU tests: the oracle against the reference's own known answers
and structural invariants.  No GPU needed."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle as O
from oracle import golovin

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_philox_known_answers():
    # Random123 kat_vectors for philox4x32-10
    assert [hex(x) for x in O.philox_raw([0, 0, 0, 0], [0, 0])] == ['0x6627e8d5', '0xe169c58d', '0xbc57ac4c', '0x9b00dbd8']
    assert [hex(x) for x in O.philox_raw([0xffffffff] * 4, [0xffffffff] * 2)] == \
        ['0x408f276d', '0x41c83b0e', '0xa20bc7c6', '0x6d5451fd']
    assert [hex(x) for x in O.philox_raw([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0])] \
        == ['0xd16cfe09', '0x94fdcceb', '0x5001e420', '0x24126ea1']


def test_mt19937_known_answers():
    # MT19937 reference: seed 5489 -> first output 3499211612, 10000th output 4123659995
    L = O.lib()
    assert L.pmco_mt_u32(5489, 0) == 3499211612
    assert L.pmco_mt_u32(5489, 9999) == 4123659995


@pytest.mark.parametrize("mean,tol", [(1.0, 1e-3), (4.0, 1e-3), (10.0, 3e-3), (20.0, 0.07), (30.0, 0.07)])
def test_rand_poisson_pmf(mean, tol):
    """partmc/test/rand/test_rand_{1..5}.sh: sampled pmf vs exp(-m) m^k / k!
    (column 2-norm relative error, numeric_diff.F90:239-248); means above 10 use
    the reference's normal approximation, hence the looser tolerance."""
    n = 2_000_000
    h = O.sample_poisson(mean, n, seed=7, rng="mt", hist_len=120)
    k = np.arange(len(h))
    pmf = np.array([O.lib().pmco_poisson_pmf(mean, int(i)) for i in k])
    err = np.linalg.norm(h / n - pmf) / np.linalg.norm(pmf)
    assert err < max(tol, 3e-3), err


@pytest.mark.parametrize("n,p,tol", [(10, 0.3, 3e-3), (10, 0.7, 3e-3), (30, 0.5, 0.07)])
def test_rand_binomial_pmf(n, p, tol):
    """partmc/test/rand/test_rand_{6..8}.sh"""
    ns = 2_000_000
    h = O.sample_binomial(n, p, ns, seed=5, rng="mt", hist_len=n + 1)
    pmf = np.array([O.lib().pmco_binomial_pmf(n, p, k) for k in range(n + 1)])
    err = np.linalg.norm(h / ns - pmf) / np.linalg.norm(pmf)
    assert err < tol, err


def test_philox_and_mt_samplers_agree_in_distribution():
    a = O.sample_poisson(3.0, 400_000, seed=1, rng="mt", hist_len=40) / 4e5
    b = O.sample_poisson(3.0, 400_000, seed=1, rng="philox", hist_len=40) / 4e5
    assert np.abs(a - b).max() < 4e-3


def test_bin_grid_make_and_find_edges():
    """bin_grid.F90:83-166, util.F90:532-580: exact end points, x == max -> last bin,
    below/above range -> 0 / n_bin+1."""
    nb, gmin, gmax = 40, 1e-9, 1e-5
    e, c = O.bin_grid_make(nb, gmin, gmax)
    assert e[0] == gmin and e[-1] == gmax and len(e) == nb + 1 and len(c) == nb
    assert np.all(np.diff(e) > 0) and np.all((c > e[:-1]) & (c < e[1:]))
    r = np.array([gmin, gmax, gmin * 0.5, gmax * 2.0, 1.05 * np.sqrt(gmin * gmax)])
    b = O.bin_grid_find(nb, gmin, gmax, r)
    assert list(b) == [1, nb, 0, nb + 1, nb // 2 + 1]
    # centers fall in their own bin
    assert np.array_equal(O.bin_grid_find(nb, gmin, gmax, c), np.arange(1, nb + 1))


def test_kernel_brown_against_numpy():
    """coag_kernel_brown.F90:111-197 restated independently in numpy (SURVEY 8c KAT grid)."""
    kB, NA, R, Ma, pi = 1.3806505e-23, 6.02214179e23, 8.314472, 2.89644e-2, np.pi

    def brown(v1, d1, v2, d2, T, p):
        rho = p * Ma / (R * T)
        visd = 1.8325e-05 * (416.16 / (T + 120.0)) * (T / 296.16) ** 1.5
        visk = visd / rho
        gs = np.sqrt(8 * kB * T * NA / (pi * Ma))
        gf = 2 * visk / gs

        def part(v, d):
            r = (3 * v / 4 / pi) ** (1 / 3)
            kn = gf / r
            cun = 1 + kn * (1.249 + 0.42 * np.exp(-0.87 / kn))
            dif = kB * T * cun / (6 * pi * r * visd)
            sp = 8 * kB * T / (pi * d * v)
            fp = 8 * dif / (pi * np.sqrt(sp))
            dl = ((2 * r + fp) ** 3 - (4 * r * r + fp * fp) ** 1.5) / (6 * r * fp) - 2 * r
            return r, dif, sp, dl * dl
        r1, D1, s1, l1 = part(v1, d1)
        r2, D2, s2, l2 = part(v2, d2)
        rs, Ds = r1 + r2, D1 + D2
        return 4 * pi * rs * Ds / (rs / (rs + np.sqrt(l1 + l2)) + 4 * Ds / (rs * np.sqrt(s1 + s2)))
    for r1 in (1e-9, 1e-8, 1e-7, 1e-6, 1e-5):
        for r2 in (1e-9, 3e-8, 1e-5):
            for d in (1000.0, 1800.0, 2600.0):
                for T in (250.0, 288.0, 310.0):
                    for p in (5e4, 1e5):
                        v1, v2 = 4 * pi * r1 ** 3 / 3, 4 * pi * r2 ** 3 / 3
                        a = O.kernel_brown(v1, d, v2, 1400.0, T, p)
                        b = brown(v1, d, v2, 1400.0, T, p)
                        assert abs(a - b) <= 1e-12 * abs(b), (a, b)
    # textbook magnitude: 10 nm + 1 um at 288 K is a few 1e-13 m^3/s
    k = O.kernel_brown(4 * pi * 5e-9 ** 3 / 3, 1000, 4 * pi * 5e-7 ** 3 / 3, 1000, 288, 1e5)
    assert 1e-14 < k < 1e-11


def _one_cell(n, seed, n_spec=1, n_class=1):
    import common
    return common.make_cells(1, n, n_spec=n_spec, n_class=n_class, seed=seed, ragged=False)


def test_sort_invariants_and_physical_order():
    """aero_sorted_check (aero_sorted.F90:499-560) after sort, and the (class, bin, index)
    physical order leaves every list a contiguous ascending run."""
    import common
    case = common.make_cells(3, 500, n_spec=3, n_class=2, seed=3)
    o = common.oracle_from_case(O, case)
    for c in range(3):
        o.sort(c, physical=False)  # runs aero_sorted_check inside
        nb, gmin, gmax, edges = o.grid(c)
        n = o.count(c)
        if n == 0:
            continue
        P = o.get_cell(c)
        r = (3 * P["vol"].sum(axis=0) / 4 / np.pi) ** (1 / 3)
        assert gmin <= r.min() / 3.0 * (1 + 1e-12) and gmax >= r.max() * 3.0 * (1 - 1e-12)
        assert nb == int(np.ceil((np.log10(gmax) - np.log10(gmin)) * 10))
        b = o.bins(c)
        assert np.all((edges[b - 1] <= r) & (r < edges[b]))
        ids_before = P["id"].copy()
        o.sort(c, force_resort=False, physical=True)
        Q = o.get_cell(c)
        b2 = o.bins(c)
        key = Q["wclass"].astype(np.int64) * 1000 + b2
        assert np.all(np.diff(key) >= 0)
        assert sorted(Q["id"]) == sorted(ids_before)
        # stable: inside one (class, bin) the original order (ids ascend in the generator) is kept
        for kk in np.unique(key):
            assert np.all(np.diff(Q["id"][key == kk]) > 0)


def test_golovin_additive_kernel():
    """partmc/test/additive: mc_coag with the additive kernel against the Golovin
    solution soln_additive_exp (coag_kernel_additive.F90:101-167) on 160 log bins
    1e-8..1e-3 m diameter at t = 600 s; the reference accepts 0.05 (number) and
    0.2 (mass) on 1e5 particles (test_additive_{1,2}.sh)."""
    rng = np.random.default_rng(1)
    n, N0, rmean = 100000, 1e9, 1e-5
    mean_vol = 4 * np.pi * rmean ** 3 / 3
    vol = -mean_vol * np.log(rng.random(n))  # aero_mode.F90 EXP branch
    d = O.AeroData([1000.0])
    o = O.Oracle(1, d, n_part_ideal=n, seed=3)
    o.set_cell(0, vol[None, :], ids=np.arange(1, n + 1))
    o.set_next_id(0, n + 1)
    o.set_weights(0, [1], [N0 / n], [0.0])
    o.set_env(0)
    for step in range(600):
        o.mc_coag(0, O.KERNEL_ADDITIVE, 1.0, step)
        o.rebalance(0, True, True, step)
    P = o.get_cell(0)
    _, m, _ = o.get_weights(0)
    err_n, err_v = golovin.compare(P["vol"][0], m[0], 600.0, N0, rmean)
    assert err_n < 0.05 and err_v < 0.2, (err_n, err_v)
    # mass is conserved to rounding by mc_coag; rebalance only duplicates / halves with weight change
    assert len(np.unique(P["id"])) == len(P["id"])


def test_mc_coag_conserves_mass_exactly_without_rebalance():
    case = _one_cell(3000, 5, n_spec=3)
    import common
    o = common.oracle_from_case(O, case)
    v0 = case["vol"].sum(axis=1)
    tot = 0
    for step in range(5):
        s, k, a = o.mc_coag(0, O.KERNEL_BROWN, 3600.0, step)
        tot += k
    P = o.get_cell(0)
    assert tot > 10
    assert o.count(0) == 3000 - tot  # flat single-class weights: every event removes exactly one
    np.testing.assert_allclose(P["vol"].sum(axis=1), v0, rtol=1e-13)


def test_reference_sort_c_matches_numpy():
    """oracle/_ref: the one reference source that compiles stand-alone (partmc/src/sort.c)."""
    path = os.path.join(ROOT, "oracle", "_ref", "libpmcref_sort.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built")
    L = C.CDLL(path)
    rng = np.random.default_rng(2)
    x = rng.random(57)
    data = x.copy()
    perm = np.zeros(57, np.int32)
    L.double_sort_c(C.c_int(57), data.ctypes.data_as(C.c_void_p), perm.ctypes.data_as(C.c_void_p))
    assert np.array_equal(data, np.sort(x))
    assert np.array_equal(x[perm - 1], data)


def test_wrf_to_partmc_types_follow_the_fortran():
    """wrf_to_partmc / rel_hum / get_grid_cell_volume (interface/wrf_pmc_driver.F90:576-666, 2002-2020): the
    oracle against an independent numpy statement of the same typed arithmetic (float32 sums, float32 constants
    promoted to double), the RH clamp to [0.1, 0.95] and the cell volume."""
    from oracle import oracle as O
    rng = np.random.default_rng(3)
    nx, nz, ny = 5, 4, 3
    n = nx * nz * ny
    pb = (9.0e4 * np.exp(-np.arange(nz) / 8.0))[None, :, None] * np.ones((ny, 1, nx))
    p = rng.normal(0, 300, (ny, nz, nx))
    t_2 = rng.normal(-5, 8, (ny, nz, nx))
    qv = rng.uniform(1e-4, 2.5e-2, (ny, nz, nx))
    alt = rng.uniform(0.8, 1.3, (ny, nz, nx))
    zw = np.cumsum(rng.uniform(40, 200, (ny, nz + 1, nx)), axis=1)
    msftx = rng.uniform(0.9, 1.1, (ny, nx))
    msfty = rng.uniform(0.9, 1.1, (ny, nx))
    rdx, rdy = 1.0 / 4000.0, 1.0 / 3000.0
    r = O.wrf_to_partmc(nx, nz, ny, p, pb, t_2, qv, alt, zw, rdx, rdy, msftx, msfty)
    f = np.float32
    pres = (f(p) + f(pb)).astype(np.float64).ravel()
    theta = (f(t_2) + f(300.0)).astype(np.float64).ravel()
    rcp = np.float64(f(287.0) / (f(7.0) * f(287.0) / f(2.0)))
    T = theta * (pres / 1e5) ** rcp
    rh = f(qv).astype(np.float64).ravel() / (3.8 * np.exp(np.float64(f(17.27)) * (T - 273.0) / (T - 36.0)) /
                                           (np.float64(f(0.01)) * pres))
    rh = np.clip(rh, 0.1, 0.95)
    dz = (f(zw)[:, 1:, :] - f(zw)[:, :-1, :]).astype(np.float64)
    dx = (1.0 / np.float64(f(rdx))) / f(msftx).astype(np.float64)
    dy = (1.0 / np.float64(f(rdy))) / f(msfty).astype(np.float64)
    vol = (dx * dy)[:, None, :] * dz
    np.testing.assert_allclose(r["pressure"], pres, rtol=0, atol=0)
    np.testing.assert_allclose(r["temp"], T, rtol=1e-15)
    np.testing.assert_allclose(r["rel_humid"], rh, rtol=1e-14)
    assert (r["rel_humid"] == 0.95).any() and (r["rel_humid"] == 0.1).any() and ((r["rel_humid"] > 0.1) & (r["rel_humid"] < 0.95)).any()
    np.testing.assert_allclose(r["height"], dz.ravel(), rtol=0, atol=0)
    np.testing.assert_allclose(r["cell_volume"], vol.ravel(), rtol=1e-15)
    np.testing.assert_array_equal(r["inverse_density"], f(alt).astype(np.float64).ravel())


def test_sample_without_replacement_is_uniform_over_subsets():
    """rand.F90:515-542 (selection sampling): ascending distinct indices, every element equally likely"""
    import ctypes as C
    L = O.lib()
    n_pop, k, n_rep = 12, 5, 6000
    hits = np.zeros(n_pop)
    out = (C.c_int * k)()
    for seed in range(n_rep):
        n = L.pmco_sample_without_replacement(k, n_pop, seed + 1, out)
        idx = np.array(out[:n])
        assert n == k and (np.diff(idx) > 0).all() and idx.min() >= 1 and idx.max() <= n_pop
        hits[idx - 1] += 1
    expect = n_rep * k / n_pop
    assert np.abs(hits - expect).max() < 5 * np.sqrt(expect)
    assert L.pmco_sample_without_replacement(n_pop, n_pop, 7, (C.c_int * n_pop)()) == n_pop  # the whole population


def test_coagulated_component_lists():
    """aero_particle_coagulate (aero_particle.F90:940-972): lists are concatenated (first particle first); beyond 20 the
    new list keeps prob_round(np1 / (np1 + np2) * 20) components of the first particle and the rest of the second,
    each a sample without replacement in the original order"""
    import ctypes as C
    L = O.lib()
    out = (C.c_int * 20)()
    n = L.pmco_coagulate_components(3, 3, 4, 4, 1, out)
    assert list(out[:n]) == [1000, 1001, 1002, 2000, 2001, 2002, 2003]
    n = L.pmco_coagulate_components(12, 12, 8, 8, 1, out)
    assert n == 20 and list(out[:20]) == list(range(1000, 1012)) + list(range(2000, 2008))
    firsts = []
    for seed in range(400):
        n = L.pmco_coagulate_components(15, 30, 10, 10, seed, out)   # 25 > 20: 30 / 40 * 20 = 15 exactly
        a = [x for x in out[:n] if x < 2000]
        b = [x for x in out[:n] if x >= 2000]
        assert n == 20 and len(a) == 15 and len(b) == 5 and a == sorted(a) and b == sorted(b) and out[14] < 2000 <= out[15]
        n = L.pmco_coagulate_components(18, 18, 7, 7, seed, out)     # 18 / 25 * 20 = 14.4 -> 14 or 15
        firsts.append(sum(1 for x in out[:n] if x < 2000))
        assert n == 20 and firsts[-1] in (14, 15)
    assert abs(np.mean(firsts) - 14.4) < 0.1
