"""Shared builders for the parity tests: seeded synthetic populations handed,
unchanged, to both the CUDA library (through the C ABI) and the CPU oracle."""
import numpy as np

# partmc/test/condense/aero_data.dat (20 species); H2O last
SPEC20 = dict(
    names=["SO4", "NO3", "Cl", "NH4", "MSA", "ARO1", "ARO2", "ALK1", "OLE1", "API1", "API2", "LIM1", "LIM2", "CO3",
           "Na", "Ca", "OIN", "OC", "BC", "H2O"],
    density=[1800, 1800, 2200, 1800, 1800, 1400, 1400, 1400, 1400, 1400, 1400, 1400, 1400, 2600, 2200, 2600, 2600,
             1400, 1800, 1000],
    num_ions=[1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 0, 0, 0, 0],
    molec_weight=[96e-3, 62e-3, 35.5e-3, 18e-3, 95e-3, 150e-3, 150e-3, 140e-3, 140e-3, 184e-3, 184e-3, 200e-3,
                  200e-3, 60e-3, 23e-3, 40e-3, 1e-3, 1e-3, 1e-3, 18e-3],
    kappa=[0, 0, 0, 0, 0.53, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0.1, 0, 0, 0, 0.1, 0.1, 0, 0],
    i_water=20)


def species(n_spec):
    """A species table with n_spec entries (the first n_spec-1 solutes of the 20-species table + water)."""
    if n_spec == 1:
        return dict(density=[1000.0], num_ions=[0], molec_weight=[18e-3], kappa=[0.0], i_water=1)
    idx = list(range(n_spec - 1)) + [19]
    if n_spec == 3:
        idx = [0, 17, 19]  # SO4, OC, H2O
    return dict(density=[SPEC20["density"][i] for i in idx], num_ions=[SPEC20["num_ions"][i] for i in idx],
                molec_weight=[SPEC20["molec_weight"][i] for i in idx], kappa=[SPEC20["kappa"][i] for i in idx],
                i_water=n_spec)


def make_cells(n_cells, n_per_cell, n_spec=1, n_class=1, n_group=1, seed=1, num_conc=1e11, r_g=2.5e-8,
               log10_sigma=0.24, ragged=True, weight="flat", temp=288.0, pressure=1e5, n_part_ideal=None,
               nx=None, nz=1, ny=1):
    """Log-normal populations (aero_mode.F90:549-628 sampling formula) in n_cells cells."""
    rng = np.random.default_rng(seed)
    sp = species(n_spec)
    counts = np.full(n_cells, n_per_cell, np.int32)
    if ragged and n_cells > 1:
        counts = rng.integers(max(1, n_per_cell // 2), n_per_cell + 1, n_cells).astype(np.int32)
        counts[0] = n_per_cell
        if n_cells > 3:
            counts[3] = 0  # an empty cell
    N = int(counts.sum())
    radius = r_g * 10.0 ** (log10_sigma * rng.standard_normal(N))
    vtot = 4.0 * np.pi * radius ** 3 / 3.0
    frac = rng.random((n_spec, N)) + 0.05
    if n_spec > 1:
        frac[-1] *= 0.2  # little water
    frac /= frac.sum(axis=0)
    vol = frac * vtot
    wclass = rng.integers(1, n_class + 1, N).astype(np.int32)
    wgroup = rng.integers(1, n_group + 1, N).astype(np.int32)
    ids = np.arange(1, N + 1, dtype=np.int64)
    n_primary = np.ones(N, np.int32)
    hyst = rng.integers(0, 2, N).astype(np.int32)
    lct = rng.random(N) * 10.0
    gct = lct + rng.random(N)
    shp = (n_cells, n_class, n_group)
    if weight == "flat":
        wtype = np.ones(shp, np.int32)
        wexp = np.zeros(shp)
        wmag = num_conc / np.maximum(counts, 1)[:, None, None] * (1.0 + 0.5 * rng.random(shp))
    else:  # nummass-like: group 1 flat, group 2 power -3
        assert n_group == 2
        wtype = np.ones(shp, np.int32)
        wtype[:, :, 1] = 2
        wexp = np.zeros(shp)
        wexp[:, :, 1] = -3.0
        wmag = np.zeros(shp)
        wmag[:, :, 0] = num_conc / np.maximum(counts, 1)[:, None]
        wmag[:, :, 1] = num_conc / np.maximum(counts, 1)[:, None] * (r_g ** 3)
    T = np.full(n_cells, temp) + rng.random(n_cells) * 10.0
    P = np.full(n_cells, pressure) - rng.random(n_cells) * 2e4
    if nx is None:
        nx = n_cells
    assert nx * nz * ny == n_cells
    return dict(n_cells=n_cells, nx=nx, nz=nz, ny=ny, n_spec=n_spec, n_class=n_class, n_group=n_group, sp=sp,
                counts=counts, vol=vol, wclass=wclass, wgroup=wgroup, id=ids, n_primary=n_primary, hyst=hyst, lct=lct,
                gct=gct, wtype=wtype, wmag=wmag, wexp=wexp, T=T, P=P, seed=seed,
                n_part_ideal=float(n_part_ideal if n_part_ideal is not None else n_per_cell))


def gpu_from_case(W, case, cell_capacity=None, max_bins=64):
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    g = W.PartMCGPU(data, nx=case["nx"], nz=case["nz"], ny=case["ny"], n_group=case["n_group"],
                    n_class=case["n_class"], n_part_ideal=case["n_part_ideal"], seed=case["seed"],
                    cell_capacity=cell_capacity or int(3 * case["counts"].max()) + 64, max_bins=max_bins)
    g.set_particles(case["counts"], case["vol"], case["wgroup"], case["wclass"], case["id"], case["n_primary"],
                    case["hyst"], case["lct"], case["gct"])
    g.set_weights(case["wtype"], case["wmag"], case["wexp"])
    g.set_env(temp=case["T"], pressure=case["P"])
    N = int(case["counts"].sum())
    # ids 1..N were handed out by "cell 0 .. n-1" generators; continue after them
    nid = np.full(case["n_cells"], N // max(case["n_cells"], 1) + 2, np.int64)
    g.set_next_id(nid)
    return g


def oracle_from_case(O, case, rng="philox"):
    sp = case["sp"]
    data = O.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    o = O.Oracle(case["n_cells"], data, n_group=case["n_group"], n_class=case["n_class"],
                 n_part_ideal=case["n_part_ideal"], rng=rng, seed=case["seed"])
    off = np.concatenate([[0], np.cumsum(case["counts"])])
    N = int(case["counts"].sum())
    for c in range(case["n_cells"]):
        a, b = off[c], off[c + 1]
        o.set_cell(c, case["vol"][:, a:b], case["wgroup"][a:b], case["wclass"][a:b], case["id"][a:b],
                   case["n_primary"][a:b], case["hyst"][a:b], case["lct"][a:b], case["gct"][a:b])
        o.set_weights(c, case["wtype"][c], case["wmag"][c], case["wexp"][c])
        o.set_env(c, T=case["T"][c], p=case["P"][c])
        o.set_next_id(c, N // max(case["n_cells"], 1) + 2)
    return o


def assert_same_population(g, o, case, exact_fp=True, rtol=0.0):
    """Device population == oracle population, cell by cell, in particle order."""
    P = g.get_particles()
    off = np.concatenate([[0], np.cumsum(P["count"])])
    t, m, e = g.get_weights()
    for c in range(case["n_cells"]):
        q = o.get_cell(c)
        a, b = off[c], off[c + 1]
        assert b - a == q["vol"].shape[1], "cell %d: %d particles on device, %d in oracle" % (c, b - a, q["vol"].shape[1])
        for k in ("id", "wgroup", "wclass", "n_primary", "hyst"):
            assert np.array_equal(P[k][a:b], q[k]), "cell %d field %s differs" % (c, k)
        if exact_fp:
            assert np.array_equal(P["vol"][:, a:b], q["vol"]), "cell %d volumes differ" % c
            assert np.array_equal(P["gct"][a:b], q["gct"]) and np.array_equal(P["lct"][a:b], q["lct"])
        else:
            np.testing.assert_allclose(P["vol"][:, a:b], q["vol"], rtol=rtol, atol=0)
        ot, om, oe = o.get_weights(c)
        np.testing.assert_allclose(m[c].ravel(), om, rtol=1e-13)


# ---------------------------------------------------------------- condensation scenario
def condense_case(n_per_mode=500, n_cells=1, seed=1):
    """The partmc/test/condense scenario as a deterministic population: both log-normal modes of
    aero_init_dist.dat represented by n_per_mode particles at the quantiles (i + 1/2) / n of their
    diameter distribution, one weight class per mode with flat weight N_mode / n_per_mode, composition
    OC : SO4 : NH4 = 1.375 : 1 : 0.375 by mass (aero_init_comp.dat), 20-species table (aero_data.dat)."""
    from scipy.stats import norm
    sp = SPEC20
    modes = [(3.2e9, 2e-8, 0.161), (2.9e9, 1.16e-7, 0.217)]
    mass_frac = {"OC": 1.375, "SO4": 1.0, "NH4": 0.375}
    vf = np.zeros(20)
    for k, val in mass_frac.items():
        i = sp["names"].index(k)
        vf[i] = val / sp["density"][i]
    vf /= vf.sum()
    vols, cls = [], []
    for m, (N, dg, ls) in enumerate(modes):
        q = (np.arange(n_per_mode) + 0.5) / n_per_mode
        diam = dg * 10 ** (ls * norm.ppf(q))
        v = np.pi / 6 * diam ** 3
        vols.append(vf[:, None] * v[None, :])
        cls += [m + 1] * n_per_mode
    vol1 = np.concatenate(vols, axis=1)
    cls1 = np.array(cls, np.int32)
    n1 = 2 * n_per_mode
    rng = np.random.default_rng(seed)
    vol = np.concatenate([vol1 * (1.0 + 0.05 * c) for c in range(n_cells)], axis=1)
    N = n1 * n_cells
    shp = (n_cells, 2, 1)
    wmag = np.zeros(shp)
    wmag[:, 0, 0] = modes[0][0] / n_per_mode
    wmag[:, 1, 0] = modes[1][0] / n_per_mode
    case = dict(n_cells=n_cells, nx=n_cells, nz=1, ny=1, n_spec=20, n_class=2, n_group=1,
                sp=dict(density=sp["density"], num_ions=sp["num_ions"], molec_weight=sp["molec_weight"],
                        kappa=sp["kappa"], i_water=20),
                counts=np.full(n_cells, n1, np.int32), vol=vol, wclass=np.tile(cls1, n_cells),
                wgroup=np.ones(N, np.int32), id=np.arange(1, N + 1, dtype=np.int64), n_primary=np.ones(N, np.int32),
                hyst=np.zeros(N, np.int32), lct=np.zeros(N), gct=np.zeros(N), wtype=np.ones(shp, np.int32), wmag=wmag,
                wexp=np.zeros(shp), T=np.full(n_cells, 290.0), P=np.full(n_cells, 1e5), seed=seed,
                n_part_ideal=float(n1))
    return case
