"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU
oracle on the same seeded inputs.  Bar: bit-exact for bins, orders, ids, counts
and volumes (sums of IEEE doubles); 1e-10 relative for kernel / majorant values."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def test_philox_device_equals_oracle(mods):
    W, O = mods
    g = W.PartMCGPU(W.AeroData([1000.0]), seed=12345)
    rng = np.random.default_rng(0)
    n = 4096
    a = [rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32) for _ in range(5)]
    a[1] = (a[1] >> 5).astype(np.uint32)  # step < 2^27
    a[2] = (a[2] % 9).astype(np.uint32)
    j = rng.integers(0, 64, n).astype(np.uint32)
    dev = g.eval_philox(a[0], a[1], a[2], a[3], a[4], j)
    ref = np.array([O.philox_uniform(12345, int(a[0][i]), int(a[1][i]), int(a[2][i]), int(a[3][i]), int(a[4][i]),
                                     int(j[i])) for i in range(n)])
    assert np.array_equal(dev, ref)
    assert dev.min() >= 0.0 and dev.max() < 1.0


def test_kernel_brown_device_vs_oracle(mods):
    """coag_kernel_brown.F90:111-197, 1e-10 relative (north star)."""
    W, O = mods
    g = W.PartMCGPU(W.AeroData([1000.0]))
    rng = np.random.default_rng(1)
    n = 20000
    r1, r2 = 10 ** rng.uniform(-9, -5, n), 10 ** rng.uniform(-9, -5, n)
    v1, v2 = 4 * np.pi * r1 ** 3 / 3, 4 * np.pi * r2 ** 3 / 3
    d1, d2 = rng.uniform(1000, 2600, n), rng.uniform(1000, 2600, n)
    T, P = rng.uniform(230, 315, n), rng.uniform(3e4, 1.05e5, n)
    dev = g.eval_kernel_brown(v1, d1, v2, d2, T, P)
    ref = np.array([O.kernel_brown(v1[i], d1[i], v2[i], d2[i], T[i], P[i]) for i in range(n)])
    rel = np.abs(dev - ref) / np.abs(ref)
    assert rel.max() < 1e-10, rel.max()


@pytest.mark.parametrize("staged_max", [None, "0", "300"])
@pytest.mark.parametrize("n_spec,n_class,n_per,n_cells", [(1, 1, 1000, 6), (3, 2, 700, 9), (20, 3, 400, 5),
                                                            (20, 2, 1100, 3), (1, 1, 9000, 2), (1, 1, 40000, 1)])
def test_sort_bit_exact(mods, monkeypatch, n_spec, n_class, n_per, n_cells, staged_max):
    """aero_state_sort: grid (n_bin, min, max), bin of every particle, per-(class, bin)
    counts and the permuted particle order all equal the oracle's, bit for bit.
    Includes a ragged population and an empty cell.  staged_max: None = the shared-memory staged (TMA)
    variant takes every cell that fits, "0" = legacy variant only, "300" = the two variants split the cells."""
    W, O = mods
    if staged_max is None:
        monkeypatch.delenv("PMCGPU_SORT_STAGED_MAX", raising=False)
    else:
        monkeypatch.setenv("PMCGPU_SORT_STAGED_MAX", staged_max)
    case = common.make_cells(n_cells, n_per, n_spec=n_spec, n_class=n_class, seed=20 + n_spec)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    g.aero_state_sort()
    nb, gmin, gmax = g.grid()
    bins = g.bins()
    bc = g.bin_counts()
    off = np.concatenate([[0], np.cumsum(g.counts())])
    for c in range(n_cells):
        o.sort(c, force_resort=True, physical=True)
        onb, omin, omax, _ = o.grid(c)
        assert nb[c] == onb
        if onb == 0:
            assert case["counts"][c] == 0
            continue
        # grid ends are r_min/10 and r_max*10 of a libm pow(): last-bit differences between
        # CUDA's and glibc's pow are allowed ("transcendentals"); bins must still agree exactly
        np.testing.assert_allclose([gmin[c], gmax[c]], [omin, omax], rtol=1e-14)
        assert np.array_equal(bins[off[c]:off[c + 1]], o.bins(c))
        assert np.array_equal(bc[c, :, :onb], o.bin_counts(c))
        assert bc[c, :, onb:].sum() == 0
    common.assert_same_population(g, o, case)
    # idempotence: sorting a sorted state changes nothing
    before = g.get_particles()
    g.aero_state_sort()
    after = g.get_particles()
    for k in before:
        assert np.array_equal(before[k], after[k])


def test_sort_keeps_grid_until_safety_factor(mods):
    """aero_sorted.F90:338-347: the grid is remade only when a particle comes within x3 of an edge."""
    W, O = mods
    case = common.make_cells(2, 300, seed=4, ragged=False)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    g.aero_state_sort()
    for c in range(2):
        o.sort(c)
    nb0, gmin0, gmax0 = g.grid()
    # grow every particle of cell 0 by x2.5 in radius: still inside the x3 margin of a x10 over-extension
    P = g.get_particles()
    vol = P["vol"].copy()
    n0 = P["count"][0]
    vol[:, :n0] *= 2.5 ** 3
    g.set_particles(P["count"], vol, P["wgroup"], P["wclass"], P["id"], P["n_primary"], P["hyst"], P["lct"], P["gct"])
    q = o.get_cell(0)
    o.set_cell(0, q["vol"] * 2.5 ** 3, q["wgroup"], q["wclass"], q["id"], q["n_primary"], q["hyst"], q["lct"], q["gct"])
    g.aero_state_sort()
    o.sort(0)
    nb1, gmin1, gmax1 = g.grid()
    assert (nb1[0], gmin1[0], gmax1[0]) == (nb0[0], gmin0[0], gmax0[0])
    np.testing.assert_allclose([nb1[0], gmin1[0], gmax1[0]], o.grid(0)[:3], rtol=1e-14)
    # x2 more crosses the margin -> new grid, identical to the oracle's
    P = g.get_particles()
    vol = P["vol"].copy()
    vol[:, :n0] *= 2.0 ** 3
    g.set_particles(P["count"], vol, P["wgroup"], P["wclass"], P["id"], P["n_primary"], P["hyst"], P["lct"], P["gct"])
    q = o.get_cell(0)
    o.set_cell(0, q["vol"] * 2.0 ** 3, q["wgroup"], q["wclass"], q["id"], q["n_primary"], q["hyst"], q["lct"], q["gct"])
    g.aero_state_sort()
    o.sort(0)
    nb2, gmin2, gmax2 = g.grid()
    np.testing.assert_allclose([nb2[0], gmin2[0], gmax2[0]], o.grid(0)[:3], rtol=1e-14)
    assert gmax2[0] != gmax0[0]
    common.assert_same_population(g, o, case)


@pytest.mark.parametrize("kernel", ["brown", "additive"])
@pytest.mark.parametrize("n_spec", [1, 20])
def test_kernel_bounds_table(mods, kernel, n_spec):
    """est_k_minmax_binned_unweighted (coag_kernel.F90:253-341): 1e-10 relative."""
    W, O = mods
    kt = W.KERNEL_BROWN if kernel == "brown" else W.KERNEL_ADDITIVE
    case = common.make_cells(3, 300, n_spec=n_spec, seed=8, ragged=False)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    g.aero_state_sort()
    g.kernel_bounds(kt)
    for c in range(3):
        o.sort(c)
        _, kmax = o.kernel_bounds(c, kt)
        dev = g.kernel_max(c)
        assert dev.shape == kmax.shape
        rel = np.abs(dev - kmax) / np.abs(kmax)
        assert rel.max() < 1e-10, rel.max()


@pytest.mark.parametrize("n_spec,n_class,n_per,n_cells,dt", [(1, 1, 1000, 8, 3600.0), (3, 2, 600, 12, 60.0),
                                                              (20, 3, 300, 6, 60.0), (1, 1, 20000, 1, 600.0)])
def test_mc_coag_bit_exact(mods, n_spec, n_class, n_per, n_cells, dt):
    """(Multi-class cases use a short step so that the accelerated per-particle path of
    coagulation.F90:150 is not selected; test_accelerated_path_is_reported covers that.)
    mc_coag (Brownian): same sample counts, same events, same surviving particles in the
    same order with identical volumes and ids as the sequential CPU oracle, over several steps."""
    W, O = mods
    case = common.make_cells(n_cells, n_per, n_spec=n_spec, n_class=n_class, seed=31 + n_class, num_conc=3e11)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    total_events = 0
    for step in range(3):
        ns, nc = g.mc_coag(W.KERNEL_BROWN, dt, step=step)
        for c in range(n_cells):
            o.sort(c, force_resort=True, physical=True, edge_recheck=True)
            s, k, acc = o.mc_coag(c, O.KERNEL_BROWN, dt, step=step)
            assert acc == 0
            assert (s, k) == (ns[c], nc[c]), "cell %d step %d: oracle (%d,%d) device (%d,%d)" % (c, step, s, k, ns[c], nc[c])
        total_events += int(nc.sum())
        common.assert_same_population(g, o, case)
        assert np.array_equal(g.get_next_id(), [o.L.pmco_get_next_id(o.h, c) for c in range(n_cells)])
    assert total_events > 20, "the case must actually coagulate (%d events)" % total_events


def test_mc_coag_additive_with_rebalance_bit_exact(mods):
    """Additive kernel + aero_state_rebalance (doubling) every step, like test/additive."""
    W, O = mods
    rng = np.random.default_rng(1)
    n, N0, rmean = 4000, 1e9, 1e-5
    mean_vol = 4 * np.pi * rmean ** 3 / 3
    vol = -mean_vol * np.log(rng.random(n))
    case = common.make_cells(1, n, seed=9, ragged=False)
    case["vol"] = vol[None, :]
    case["wmag"][:] = N0 / n
    g = common.gpu_from_case(W, case, cell_capacity=3 * n)
    o = common.oracle_from_case(O, case)
    doubled = 0
    for step in range(120):
        ns, nc = g.mc_coag(W.KERNEL_ADDITIVE, 5.0, step=step)
        g.aero_state_rebalance(True, True, step=step)
        o.sort(0, force_resort=True, physical=True, edge_recheck=True)
        s, k, _ = o.mc_coag(0, O.KERNEL_ADDITIVE, 5.0, step=step)
        n_before = o.count(0)
        o.rebalance(0, True, True, step=step)
        doubled += o.count(0) > n_before
        assert (s, k) == (ns[0], nc[0])
    assert doubled >= 1
    common.assert_same_population(g, o, case)


def test_rebalance_halving_bit_exact(mods):
    """aero_state_halve (aero_state.F90:1726): backwards swap-with-last removal, magnitude x2."""
    W, O = mods
    case = common.make_cells(5, 900, n_spec=3, n_class=2, seed=13, n_part_ideal=300.0)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    g.aero_state_rebalance(True, True, step=4)
    for c in range(5):
        o.rebalance(c, True, True, step=4)
    assert g.counts().sum() < case["counts"].sum()
    common.assert_same_population(g, o, case)


def test_golovin_on_device(mods):
    """K1: additive kernel on the GPU against the Golovin solution, 16 seeds at once (16 cells: the cell
    number is part of every draw's address).  test/additive's tolerances 0.05 / 0.2 are for 1e5 particles;
    with 1e4 particles per member the seed-averaged distribution errors are checked against a proportionally
    wider bound, total number against the analytic decay and total mass against the initial one."""
    W, O = mods
    from oracle import golovin
    n, N0, rmean, M = 10000, 1e9, 1e-5, 16
    mean_vol = 4 * np.pi * rmean ** 3 / 3
    rng = np.random.default_rng(100)
    vol = -mean_vol * np.log(rng.random(M * n))
    case = common.make_cells(M, n, seed=1, ragged=False)
    case["vol"] = vol[None, :]
    case["wmag"][:] = N0 / n
    case["T"][:] = 288.0
    case["P"][:] = 1e5
    g = common.gpu_from_case(W, case, cell_capacity=3 * n)
    for step in range(600):
        g.mc_coag(W.KERNEL_ADDITIVE, 1.0, step=step)
        g.aero_state_rebalance(True, True, step=step)
    P = g.get_particles()
    _, m, _ = g.get_weights()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    n_expect = N0 * np.exp(-N0 * mean_vol * 1000.0 * 600.0)
    errs, num, mass = [], [], []
    for c in range(M):
        v = P["vol"][0, off[c]:off[c + 1]]
        errs.append(golovin.compare(v, m[c, 0, 0], 600.0, N0, rmean))
        num.append(v.size * m[c, 0, 0])
        mass.append(v.sum() * m[c, 0, 0] / (vol[c * n:(c + 1) * n].sum() * N0 / n))
    errs, num, mass = np.array(errs), np.array(num), np.array(mass)
    assert np.abs(num / n_expect - 1).max() < 0.1
    # ensemble mean of the number concentration inside its own 95 % confidence interval around the analytic value
    assert abs(num.mean() - n_expect) < 2.13 * num.std(ddof=1) / np.sqrt(M) + 0.01 * n_expect
    assert np.abs(mass - 1).max() < 0.08  # doubling / halving keeps mass in expectation only
    assert errs[:, 0].mean() < 0.15 and errs[:, 1].mean() < 0.45, errs.mean(axis=0)


@pytest.mark.parametrize("weight,n_spec,n_class,n_group,n_per,n_cells,dt", [
    ("flat", 3, 2, 1, 600, 12, 7200.0), ("flat", 20, 3, 1, 400, 6, 10000.0), ("nummass", 3, 1, 2, 1500, 4, 3600.0),
    ("nummass", 1, 2, 2, 800, 6, 7200.0)])
def test_accelerated_per_particle_coag_bit_exact(mods, weight, n_spec, n_class, n_group, n_per, n_cells, dt):
    """try_per_particle_coag (coagulation.F90:150-241): when determine_target_and_source / compute_n_source
    (:247-329) select the accelerated path, the device takes it like the oracle: same sample counts, same
    events, same particles (ids, groups, volumes bit for bit) in the same order, over several steps."""
    W, O = mods
    case = common.make_cells(n_cells, n_per, n_spec=n_spec, n_class=n_class, n_group=n_group, weight=weight,
                             seed=33, num_conc=3e11)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    acc = 0
    events = 0
    for step in range(3):
        ns, nc = g.mc_coag(W.KERNEL_BROWN, dt, step=step)
        for c in range(n_cells):
            o.sort(c, force_resort=True, physical=True, edge_recheck=True)
            s, k, a = o.mc_coag(c, O.KERNEL_BROWN, dt, step=step)
            acc += a
            assert (s, k) == (ns[c], nc[c]), "cell %d step %d: oracle (%d,%d) device (%d,%d)" % (c, step, s, k, ns[c], nc[c])
        events += int(nc.sum())
        common.assert_same_population(g, o, case)
        assert np.array_equal(g.get_next_id(), [o.L.pmco_get_next_id(o.h, c) for c in range(n_cells)])
    assert acc > 0, "the case must select the accelerated path"
    assert events > 20


@pytest.mark.parametrize("weight,n_group", [("flat", 1), ("nummass", 2)])
def test_mc_coag_k2_single_box_bit_exact(mods, weight, n_group):
    """BASELINE config K2 at full size: one box, 1e5 particles, Brownian kernel, flat and nummass-like
    weights (weighted sampling: coagulate_weighting's remove / create probabilities, coagulation.F90:810)."""
    W, O = mods
    case = common.make_cells(1, 100000, n_spec=1, n_group=n_group, weight=weight, seed=5, ragged=False,
                             num_conc=1e11, r_g=2.5e-8)
    g = common.gpu_from_case(W, case, cell_capacity=120000)
    o = common.oracle_from_case(O, case)
    for step in range(2):
        ns, nc = g.mc_coag(W.KERNEL_BROWN, 600.0, step=step)
        o.sort(0, force_resort=True, physical=True, edge_recheck=True)
        s, k, _ = o.mc_coag(0, O.KERNEL_BROWN, 600.0, step=step)
        assert (s, k) == (ns[0], nc[0])
        assert k > 5000
        common.assert_same_population(g, o, case)


@pytest.mark.parametrize("n_spec,n_class,n_group", [(3, 2, 1), (1, 2, 2)])
def test_scale_weight_both_branches_bit_exact(mods, n_spec, n_class, n_group):
    """aero_state_scale_weight (aero_state.F90:1832-1895): ratio > 1 (random removals with the list surgery
    of integer_rmap2_remove), ratio < 1 (backward duplication with new ids), ratio 1 and refused cases
    (allow_halving / allow_doubling false), per (group, class), against the oracle."""
    W, O = mods
    n_cells = 8
    case = common.make_cells(n_cells, 500, n_spec=n_spec, n_class=n_class, n_group=n_group, seed=21,
                             weight="flat" if n_group == 1 else "nummass")
    g = common.gpu_from_case(W, case, cell_capacity=8000)
    o = common.oracle_from_case(O, case)
    ratios = np.array([1.7, 0.4, 1.0, 3.2, 0.93, 1.01, 0.25, 2.0])
    step = 0
    for (gg, cc, ad, ah) in [(1, 1, True, True), (n_group, n_class, True, True), (1, n_class, False, True),
                             (1, 1, True, False)]:
        g.aero_state_scale_weight(gg, cc, ratios, allow_doubling=ad, allow_halving=ah, step=step)
        for c in range(n_cells):
            if ratios[c] != 1.0:
                # the device keeps no lists between calls: every call starts from a fresh aero_state_sort
                o.sort(c, force_resort=True, physical=False)
                o.scale_weight(c, gg, cc, ratios[c], ad, ah, step=step)
        common.assert_same_population(g, o, case)
        assert np.array_equal(g.get_next_id(), [o.L.pmco_get_next_id(o.h, c) for c in range(n_cells)])
        step += 1
    assert not np.array_equal(g.counts(), case["counts"])
