"""CPU tests of the dry-deposition restatement (interface/wrf_pmc_dep.F90, wrf_pmc_dep_aero.F90): an
independent numpy evaluation of the same formulas (with the Fortran's single-precision literals), physical
sanity of the deposition velocity curve, and the removal loop's bookkeeping."""
import numpy as np

import common
from oracle import oracle as O

F = np.float32


def numpy_dep(a):
    d, rho, T, rrho, ust, rmol, znt, ivg, dt, dz = [a[:, k] for k in range(10)]
    ust = ust.astype(F).astype(float)
    znt = znt.astype(F).astype(float)
    rmol = rmol.astype(F)
    rmol = np.where(np.abs(rmol) < F(1e-6), F(0), rmol)
    zr = 2.0
    c074, c001 = float(F(0.74)), float(F(0.001))
    rm = rmol.astype(float)
    with np.errstate(invalid="ignore", divide="ignore"):
        ar = ((1.0 - 9.0 * zr * rm) ** 0.25 + c001) ** 2
        ao = ((1.0 - 9.0 * znt * rm) ** 0.25 + c001) ** 2
        unstable = c074 * (np.log((ar - 1) / (ar + 1)) - np.log((ao - 1) / (ao + 1)))
    neutral = c074 * np.log(zr / znt)
    stable = neutral + (F(4.7) * rmol).astype(float) * (zr - znt)
    polint = np.where(rmol < 0, unstable, np.where(rmol == 0, neutral, stable))
    ra = polint / (0.4 * np.maximum(ust, float(F(1e-4))))
    zh = np.array([15, 7, 7, 7, 7, 10, 6, 6, 6, 6, 4, 3, 2, 1, 5, 14, 11, 10, 8, 9, 10, 9, 8, 12, 8])
    iv = ivg.astype(int)
    st = np.where((iv >= 1) & (iv <= 25), zh[np.clip(iv, 1, 25) - 1], 7)
    alpha = np.array([1.0, 0.6, 1.1, 0.8, 0.8, 1.2, 1.2, 50.0, 50.0, 1.3, 2.0, 50.0, 100.0, 100.0, 1.5], F).astype(float)[st - 1]
    A = np.array([0.002, 0.005, 0.002, 0.005, 0.005, 0.002, 0.002, 9.999, 9.999, 0.010, 0.010, 9.999, 9.999, 9.999,
                  0.010], F).astype(float)[st - 1]
    lam, visc, grav, kb = .065e-6, 1.78e-5, 9.81, 1.3806505e-23
    cc = 1.0 + (2 * lam / d) * (float(F(1.257)) + float(F(.4)) * np.exp(-(float(F(1.1)) * d) / (2 * lam)))
    vs = d ** 2 * rho * grav * cc / (18 * visc)
    diff = kb * T * cc / (3 * np.pi * visc * d)
    Sc = 1.45e-5 / diff
    St = np.where(A > 1, vs * ust ** 2 / (visc * rrho), vs * ust / (grav * A))
    EB = float(F(.2)) * Sc ** (-float(F(2.0) / F(3.0)))
    EIM = float(F(0.4)) * (St / (alpha + St)) ** 2
    EIN = .5 * (d / A) ** float(F(0.8))
    R1 = np.exp(-np.sqrt(St))
    rs = 1.0 / (3.0 * ust * (EB + EIM + EIN) * R1)
    vd = 1.0 / (ra + rs + ra * rs * vs) + vs
    return np.stack([ra, vd, dt / dz * vd], axis=1)


def test_deposition_formulas_against_numpy():
    rng = np.random.default_rng(2)
    n = 5000
    a = np.zeros((n, 10))
    a[:, 0] = 10.0 ** rng.uniform(-8.3, -4.7, n)
    a[:, 1] = rng.uniform(900.0, 2600.0, n)
    a[:, 2] = rng.uniform(250.0, 310.0, n)
    a[:, 3] = rng.uniform(0.75, 1.3, n)
    a[:, 4] = 10.0 ** rng.uniform(-5.0, 0.0, n)
    a[:, 5] = rng.choice([-0.1, -1e-3, -1e-7, 0.0, 5e-7, 1e-3, 0.05], n)
    a[:, 6] = 10.0 ** rng.uniform(-4.0, 0.0, n)
    a[:, 7] = rng.integers(0, 28, n)
    a[:, 8] = 6.0
    a[:, 9] = rng.uniform(10.0, 120.0, n)
    ref = numpy_dep(a)
    got = O.dep_eval(a)
    assert np.abs(got / ref - 1).max() < 1e-12


def test_deposition_velocity_curve_is_physical():
    """v_d(D) has its minimum in the accumulation mode and tends to the settling velocity for coarse particles."""
    d = 10.0 ** np.linspace(-8, -4.3, 60)
    a = np.zeros((d.size, 10))
    a[:, 0], a[:, 1], a[:, 2], a[:, 3], a[:, 4] = d, 1500.0, 288.0, 0.85, 0.3
    a[:, 5], a[:, 6], a[:, 7], a[:, 8], a[:, 9] = 0.0, 0.1, 7, 6.0, 50.0
    out = O.dep_eval(a)
    vd = out[:, 1]
    k = vd.argmin()
    assert 1e-7 < d[k] < 3e-6
    assert vd[0] > 3 * vd[k] and vd[-1] > 30 * vd[k]
    assert ((out[:, 2] > 0) & (out[:, 2] < 1)).all()


def test_dry_deposition_only_touches_interior_surface_cells():
    nx, nz, ny = 5, 2, 4
    case = common.make_cells(nx * nz * ny, 300, n_spec=3, seed=5, r_g=5e-7, log10_sigma=0.5, nx=nx, nz=nz, ny=ny)
    o = common.oracle_from_case(O, case)
    before = np.array([o.count(c) for c in range(case["n_cells"])]).reshape(ny, nz, nx)
    f32 = lambda v: np.full((ny, nx), v, np.float32)
    o.dep_dry_driver(nx, nz, ny, f32(0.0), f32(0.1), f32(0.4), np.full((ny, nx), 7, np.int32), np.full((ny, nx), 10.0),
                     600.0, step=1)
    after = np.array([o.count(c) for c in range(case["n_cells"])]).reshape(ny, nz, nx)
    diff = before - after
    assert (diff[:, 1] == 0).all() and (diff[0] == 0).all() and (diff[-1] == 0).all()
    assert (diff[:, :, 0] == 0).all() and (diff[:, :, -1] == 0).all()
    assert diff[1:-1, 0, 1:-1].sum() > 10
    for c in range(case["n_cells"]):
        q = o.get_cell(c)
        assert len(set(q["id"].tolist())) == q["id"].size
