"""Sea salt emissions (interface/wrf_pmc_driver.F90:212-223, :1767-1924): the device call against the oracle's
restatement -- same particles added per cell, same population (ids, classes, volumes at 1e-12 like every freshly
sampled particle: erf / pow / cos differ in the last bits between libm and the device), only over water at k = 1."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


@pytest.mark.parametrize("opt,n_class", [(1, 3), (2, 5)])
def test_seasalt_emissions_match_the_oracle(mods, opt, n_class):
    W, O = mods
    nx, nz, ny = 6, 3, 4
    case = common.make_cells(nx * nz * ny, 300, n_spec=20, n_class=n_class, seed=40 + opt, nx=nx, nz=nz, ny=ny,
                             ragged=False, num_conc=1e6)
    g = common.gpu_from_case(W, case, cell_capacity=6000)
    o = common.oracle_from_case(O, case)
    rng = np.random.default_rng(3)
    u10 = (rng.random((ny, nx)) * 12.0).astype(np.float32)
    v10 = (rng.random((ny, nx)) * 6.0 - 3.0).astype(np.float32)
    xland = np.where(rng.random((ny, nx)) < 0.6, 2.0, 1.0).astype(np.float32)
    height = 40.0 + 20.0 * rng.random((ny, nx))
    n_modes = 2 if opt == 1 else 5
    wc = (np.arange(n_modes) % n_class) + 1
    i_na, i_cl = 5, 4
    total = 0
    for step, elapsed in ((1, 0.0), (2, 60.0), (3, 7200.0)):
        nd = g.seasalt_emissions(u10, v10, xland, height, opt, i_na, i_cl, wc, 60.0, elapsed, step=step)
        no = o.seasalt_emissions(nx, nz, ny, u10, v10, xland, height, opt, i_na, i_cl, wc, 60.0, elapsed, step=step)
        assert np.array_equal(nd, no), (nd[nd != no], no[nd != no])
        total += int(nd.sum())
        k = (np.arange(nx * nz * ny) // nx) % nz
        land = np.repeat(xland[:, None, :] <= 1.5, nz, axis=1).ravel()
        assert (nd[(k > 0) | land] == 0).all()
    assert total > 200
    common.assert_same_population(g, o, case, exact_fp=False, rtol=1e-12)
    P = g.get_particles()
    new = P["lct"] == 7200.0
    assert new.any()
    v = P["vol"][:, new]
    frac = v[i_na - 1] / v.sum(axis=0)
    np.testing.assert_allclose(frac, 22.990 / (22.990 + 35.450), rtol=1e-12)
    assert np.count_nonzero(v.sum(axis=1)) == 2  # Na and Cl only
