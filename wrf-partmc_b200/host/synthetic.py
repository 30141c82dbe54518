"""Seeded synthetic inputs of the shapes BASELINE.json names: particle
populations drawn with the reference's sampling formulas and WRF-like meteorology / fluxes for an
idealised patch.  Host-side helper of the benchmark and the tests; nothing here runs the aerosol
step.

References: log-normal radius sampling aero_mode.F90:549-628 (r = 10**(normal(log10 r_g, log10 sigma)));
composition by mass fractions aero_mode.F90:633; 20-species table partmc/test/condense/aero_data.dat;
the two modes of partmc/test/condense/aero_init_dist.dat; idealised WRF column wrf_pmc_init.F90:1505.
"""
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

# partmc/test/condense/aero_init_dist.dat: (num_conc m^-3, geometric mean diameter m, log10 sigma)
MODES = [(3.2e9, 2e-8, 0.161), (2.9e9, 1.16e-7, 0.217)]
# partmc/test/condense/aero_init_comp.dat, mass fractions
MASS_FRAC = {"OC": 1.375, "SO4": 1.0, "NH4": 0.375}


def species_table(n_spec=20):
    """The species the condensation scenario needs (SO4, NH4, OC, H2O) inside an n_spec table."""
    if n_spec == 20:
        idx = list(range(20))
    elif n_spec == 4:
        idx = [0, 3, 17, 19]
    else:
        raise ValueError("n_spec must be 4 or 20")
    sp = {k: [SPEC20[k][i] for i in idx] for k in ("names", "density", "num_ions", "molec_weight", "kappa")}
    sp["i_water"] = len(idx)
    return sp


def make_patch(nx, nz, ny, n_per_cell, n_spec=20, seed=1, num_conc_scale=10.0, ragged=0.0):
    """An idealised LES-style patch: every cell holds n_per_cell particles, half from each mode of the
    condensation scenario, one weight class per mode (flat weights, aero_state.F90:241 +
    wrf_pmc_init.F90:251-254), dry composition OC:SO4:NH4 = 1.375:1:0.375 by mass.  Cell order is WRF
    memory order (i fastest, then k, then j).  Returns a dict of plain numpy arrays."""
    rng = np.random.default_rng(seed)
    sp = species_table(n_spec)
    S = n_spec
    n_cells = nx * nz * ny
    n_class = 2
    per_mode = n_per_cell // 2
    counts = np.full(n_cells, 2 * per_mode, np.int32)
    if ragged > 0:
        counts = (counts * (1.0 - ragged * rng.random(n_cells))).astype(np.int32)
    N = int(counts.sum())
    cell_of = np.repeat(np.arange(n_cells), counts)
    # class 1 for the first half of each cell's particles, class 2 for the rest
    first = np.concatenate([[0], np.cumsum(counts)[:-1]])
    local = np.arange(N) - first[cell_of]
    wclass = np.where(local < (counts[cell_of] + 1) // 2, 1, 2).astype(np.int32)
    dg = np.where(wclass == 1, MODES[0][1], MODES[1][1])
    ls = np.where(wclass == 1, MODES[0][2], MODES[1][2])
    diam = dg * 10.0 ** (ls * rng.standard_normal(N))
    vdry = np.pi / 6.0 * diam ** 3
    vf = np.zeros(S)
    for k, val in MASS_FRAC.items():
        i = sp["names"].index(k)
        vf[i] = val / sp["density"][i]
    vf /= vf.sum()
    vol = np.zeros((S, N))
    for s in range(S):
        if vf[s] != 0.0:
            vol[s] = vf[s] * vdry
    shp = (n_cells, n_class, 1)
    n1 = (counts + 1) // 2
    n2 = counts - n1
    wmag = np.zeros(shp)
    wmag[:, 0, 0] = num_conc_scale * MODES[0][0] / np.maximum(n1, 1)
    wmag[:, 1, 0] = num_conc_scale * MODES[1][0] / np.maximum(n2, 1)
    # column meteorology: dry-adiabatic-like T(z), hydrostatic p, ideal-gas density
    k_of = (np.arange(n_cells) // nx) % nz
    z = (k_of + 0.5) * 100.0
    T = 290.0 - 0.0065 * z + 0.2 * rng.random(n_cells)
    P = 1e5 * np.exp(-z / 8400.0)
    rrho = 287.04 * T / P  # inverse density (grid%alt)
    RH = np.full(n_cells, 0.95)
    return dict(n_cells=n_cells, nx=nx, nz=nz, ny=ny, n_spec=S, n_class=n_class, n_group=1, sp=sp, counts=counts,
                vol=vol, wclass=wclass, wgroup=np.ones(N, np.int32), id=np.arange(1, N + 1, dtype=np.int64),
                n_primary=np.ones(N, np.int32), hyst=np.zeros(N, np.int32), lct=np.zeros(N), gct=np.zeros(N),
                wtype=np.ones(shp, np.int32), wmag=wmag, wexp=np.zeros(shp), T=T, P=P, RH=RH, rrho=rrho,
                cell_volume=np.full(n_cells, 100.0 ** 3), seed=seed, n_part_ideal=float(2 * per_mode))


def met_sequence(case, n_steps, dT=-0.1, dp=-100.0):
    """Per-step (T, p, inverse density) fields: a slow uniform cooling / lifting (the condense test drives
    290 -> 280 K and 1e5 -> 8e4 Pa over 1200 s, partmc/test/condense/temp.dat, pressure.dat)."""
    out = []
    for k in range(1, n_steps + 1):
        T = case["T"] + dT * k
        P = case["P"] + dp * k
        out.append((T, P, 287.04 * T / P))
    return out
