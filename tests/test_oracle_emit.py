"""CPU tests of the emission-sampling restatement (aero_state_add_aero_dist_sample, aero_state.F90:758-831):
the sampled population reproduces the mode it was drawn from."""
import numpy as np
import pytest

import common
from oracle import oracle as O


def test_lognormal_emission_reproduces_the_mode():
    """Flat weight: expected count N sample_prop / magnitude, log10 radius ~ Normal(log10 r_g, sigma) (up to the
    tail threshold of rand_normal, here 1 / (characteristic_factor n_avg) << 1), composition = vol_frac."""
    case = common.make_cells(1, 2000, n_spec=3, n_class=1, seed=3, n_part_ideal=4000.0, ragged=False)
    o = common.oracle_from_case(O, case)
    mag = case["wmag"][0, 0, 0]
    vf = np.array([[0.2, 0.5, 0.3]])
    n_target = 1500.0
    added = o.add_aero_dist_sample(0, mode_type=[1], weight_class=[1], char_radius=[5e-8], log10_std_dev=[0.2],
                                   vol_frac=vf, num_conc=[n_target * mag], sample_prop=1.0, characteristic_factor=60.0,
                                   create_time=7.0, step=1)
    assert abs(added - n_target) < 4 * np.sqrt(n_target)
    q = o.get_cell(0)
    new = q["vol"][:, 2000:]
    assert new.shape[1] == added and o.count(0) == 2000 + added
    r = (3 * new.sum(axis=0) / (4 * np.pi)) ** (1 / 3)
    lr = np.log10(r)
    assert abs(lr.mean() - np.log10(5e-8)) < 4 * 0.2 / np.sqrt(added)
    assert abs(lr.std() - 0.2) < 0.015
    np.testing.assert_allclose(new / new.sum(axis=0), np.repeat(vf.T, added, axis=1), rtol=1e-12)
    assert (q["gct"][2000:] == 7.0).all() and (q["n_primary"][2000:] == 1).all()
    assert len(set(q["id"].tolist())) == q["id"].size
    _, m1, _ = o.get_weights(0)
    assert m1[0] == mag  # 3500 particles is inside [ideal / 2, 2 ideal]: no rescaling


def test_prepare_weight_for_add_rescales_when_the_cell_would_overflow():
    case = common.make_cells(1, 900, n_spec=1, n_class=1, seed=3, n_part_ideal=500.0, ragged=False)
    o = common.oracle_from_case(O, case)
    mag = case["wmag"][0, 0, 0]
    o.add_aero_dist_sample(0, mode_type=[3], weight_class=[1], char_radius=[1e-7], log10_std_dev=[0.0],
                           vol_frac=[[1.0]], num_conc=[600.0 * mag], sample_prop=1.0, characteristic_factor=10.0, step=2)
    _, m1, _ = o.get_weights(0)
    # n_part_new = 900 + 600 = 1500 > 2 x 500: weight_ratio = 3, population thinned to ~ 300 + 200 new
    np.testing.assert_allclose(m1[0], 3.0 * mag, rtol=1e-14)
    assert 380 < o.count(0) < 620
    q = o.get_cell(0)
    r = (3 * q["vol"][0] / (4 * np.pi)) ** (1 / 3)
    n_mono = int((np.abs(r / 1e-7 - 1) < 1e-12).sum())
    assert abs(n_mono - 200) < 4 * np.sqrt(200)


def test_seasalt_mode_tables():
    """single-precision literals of the Fortran tables widened to double (real(kind=dp) :: x = [ .018, ... ])"""
    from oracle import oracle as O
    cr, ls, fl, vf = O.seasalt_modes(1, 20, 5, 4)
    assert cr[0] == float(np.float32(2.03347408e-07)) / 2 and ls[1] == np.log10(float(np.float32(2.2)))
    assert fl[0] == float(np.float32(73.7095947))
    cr, ls, fl, vf = O.seasalt_modes(2, 3, 1, 2)
    assert cr[4] == 0.5 * float(np.float32(.83)) * 1e-6
    assert fl[1] == pytest.approx(float(np.float32(.0442)) * (3.1e6 - 1e5) ** float(np.float32(1.08)), rel=1e-14)
    np.testing.assert_allclose(vf.sum(axis=1), 1.0, rtol=1e-15)


def test_seasalt_emissions_expected_number():
    """seasalt_emissions (wrf_pmc_driver.F90:1767): over water at k = 1 the expected number of new computational
    particles is sum_modes flux wind**3.41 dt / height / magnitude; nothing over land or aloft"""
    nx, nz, ny = 4, 2, 3
    case = common.make_cells(nx * nz * ny, 200, n_spec=20, n_class=2, seed=5, nx=nx, nz=nz, ny=ny, ragged=False,
                             num_conc=2e5, n_part_ideal=1000.0)
    o = common.oracle_from_case(O, case)
    u10 = np.full((ny, nx), 8.0, np.float32)
    v10 = np.full((ny, nx), 6.0, np.float32)  # wind speed 10 m/s
    xland = np.full((ny, nx), 2.0, np.float32)
    xland[0, 0] = 1.0
    height = np.full((ny, nx), 50.0)
    n_add = np.zeros(nx * nz * ny)
    n_rep = 40
    for step in range(n_rep):
        n_add += o.seasalt_emissions(nx, nz, ny, u10, v10, xland, height, 1, 5, 4, [1, 2], 60.0, 600.0, step=step,
                                     allow_doubling=False, allow_halving=False)
    n_add = n_add.reshape(ny, nz, nx)
    assert n_add[:, 1, :].sum() == 0 and n_add[0, 0, 0] == 0
    mag = case["wmag"].reshape(ny, nz, nx, 2)[:, 0, :, :]
    scale = 10.0 ** float(np.float32(3.41)) * 60.0 / 50.0
    expect = n_rep * scale * (float(np.float32(73.7095947)) / mag[..., 0] + float(np.float32(10.1706810)) / mag[..., 1])
    expect[0, 0] = 0.0
    tot, exp_tot = n_add[:, 0, :].sum(), expect.sum()
    assert abs(tot - exp_tot) < 5.0 * np.sqrt(exp_tot), (tot, exp_tot)


def test_sampled_mode_reproduces_its_table():
    """mode_type sampled (aero_mode.F90:456-491, 533-538, 581-608): expected count sum(n_k) sample_prop / magnitude,
    every bin populated in proportion to its tabulated number concentration, radii log-uniform inside a bin"""
    case = common.make_cells(1, 10, n_spec=2, n_class=1, seed=3, n_part_ideal=1e6, ragged=False)
    o = common.oracle_from_case(O, case)
    mag = case["wmag"][0, 0, 0]
    edges = np.logspace(-8, -6, 6)                      # 5 bins
    table = np.array([1.0, 4.0, 2.0, 0.0, 3.0]) * 4000.0 * mag
    n = o.add_aero_dist_sample(0, mode_type=[4], weight_class=[1], char_radius=[0.0], log10_std_dev=[0.0],
                               vol_frac=[[0.3, 0.7]], num_conc=[1.0], sample_prop=1.0, characteristic_factor=10.0,
                               step=1, allow_doubling=False, allow_halving=False, sample_radius=[edges],
                               sample_num_conc=[table])
    assert abs(n - 40000) < 5 * np.sqrt(40000)
    q = o.get_cell(0)
    r = (3 * q["vol"].sum(axis=0)[10:] / (4 * np.pi)) ** (1 / 3)
    hist = np.histogram(r, edges)[0]
    expect = table / table.sum() * n
    assert hist[3] == 0 and np.all(np.abs(hist - expect) < 5 * np.sqrt(expect + 1))
    inside = (np.log(r[(r > edges[1]) & (r < edges[2])]) - np.log(edges[1])) / np.log(edges[2] / edges[1])
    assert abs(inside.mean() - 0.5) < 0.02 and abs(inside.std() - 12 ** -0.5) < 0.02
