"""GPU parity tests for dry deposition (interface/wrf_pmc_dep.F90,
interface/wrf_pmc_dep_aero.F90): formulas to 1e-10, removal bookkeeping bit-exact given the draws."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def dep_inputs(n, seed=1):
    rng = np.random.default_rng(seed)
    a = np.zeros((n, 10))
    a[:, 0] = 10.0 ** rng.uniform(-8.3, -4.7, n)          # diameter
    a[:, 1] = rng.uniform(900.0, 2600.0, n)               # density
    a[:, 2] = rng.uniform(250.0, 310.0, n)                # temperature
    a[:, 3] = rng.uniform(0.75, 1.3, n)                   # inverse density
    a[:, 4] = 10.0 ** rng.uniform(-5.0, 0.0, n)           # ustar (below 1e-4: the max() floor)
    a[:, 5] = rng.choice([-0.1, -1e-3, -1e-7, 0.0, 5e-7, 1e-3, 0.05], n) * rng.uniform(0.5, 1.5, n)  # rmol
    a[:, 6] = 10.0 ** rng.uniform(-4.0, 0.0, n)           # znt
    a[:, 7] = rng.integers(0, 28, n)                      # ivgtyp incl. out-of-range values
    a[:, 8] = rng.choice([1.0, 6.0, 60.0], n)             # dt
    a[:, 9] = rng.uniform(10.0, 120.0, n)                 # dz
    return a


def test_deposition_formulas_device_vs_oracle(mods):
    """depvel_partmc (wrf_pmc_dep.F90:188), calculate_vs / rs / vd and the removal probability
    (wrf_pmc_dep_aero.F90:204-470) for all land-use types, stabilities and sizes: 1e-10 relative."""
    W, O = mods
    g = W.PartMCGPU(W.AeroData([1000.0]))
    a = dep_inputs(20000)
    dev = g.eval_dep(a)
    ref = O.dep_eval(a)
    assert np.isfinite(ref).all()
    rel = np.abs(dev - ref) / np.abs(ref)
    assert rel.max() < 1e-10, rel.max(axis=0)
    assert (ref[:, 2] > 0).all()


def surface_fields(nx, ny, seed=3):
    rng = np.random.default_rng(seed)
    rmol = rng.choice([-0.05, 0.0, 0.02], (ny, nx)).astype(np.float32)
    znt = (10.0 ** rng.uniform(-3, -0.3, (ny, nx))).astype(np.float32)
    ust = rng.uniform(0.05, 0.8, (ny, nx)).astype(np.float32)
    ivg = rng.integers(1, 26, (ny, nx)).astype(np.int32)
    height = rng.uniform(5.0, 20.0, (ny, nx))
    return rmol, znt, ust, ivg, height


@pytest.mark.parametrize("n_spec", [1, 20])
def test_dry_deposition_bit_exact(mods, n_spec):
    """wrf_pmc_dep_dry_driver over a 6 x 3 x 5 patch: only k = 1 cells inside the outer ring lose particles;
    the survivors, their order and everything else equal the oracle's backward swap-with-last loop."""
    W, O = mods
    nx, nz, ny = 6, 3, 5
    case = common.make_cells(nx * nz * ny, 700, n_spec=n_spec, n_class=2, seed=5, r_g=4e-7, log10_sigma=0.5,
                             nx=nx, nz=nz, ny=ny)
    g = common.gpu_from_case(W, case)
    o = common.oracle_from_case(O, case)
    rrho = np.full(case["n_cells"], 0.85)
    g.set_env(inverse_density=rrho)
    for c in range(case["n_cells"]):
        o.set_env(c, T=case["T"][c], p=case["P"][c], rrho=rrho[c])
    rmol, znt, ust, ivg, height = surface_fields(nx, ny)
    before = g.counts().copy()
    for step in range(3):
        g.wrf_pmc_dep_dry_driver(rmol, znt, ust, ivg, height, dt=60.0, step=step)
        o.dep_dry_driver(nx, nz, ny, rmol, znt, ust, ivg, height, 60.0, step=step)
        common.assert_same_population(g, o, case)
    after = g.counts().reshape(ny, nz, nx)
    b3 = before.reshape(ny, nz, nx)
    assert np.array_equal(after[:, 1:, :], b3[:, 1:, :])                       # upper layers untouched
    assert np.array_equal(after[0], b3[0]) and np.array_equal(after[-1], b3[-1])  # outer ring untouched
    assert np.array_equal(after[:, :, 0], b3[:, :, 0]) and np.array_equal(after[:, :, -1], b3[:, :, -1])
    lost = (b3[1:-1, 0, 1:-1] - after[1:-1, 0, 1:-1])
    assert lost.sum() > 50 and (lost >= 0).all()
