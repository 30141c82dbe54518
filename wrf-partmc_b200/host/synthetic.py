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


def met_sequence(case, n_steps, dT=-0.4, dp=-400.0, period=8):
    """Per-step (T, p, inverse density) fields: every cell is lifted and lowered adiabatically like a
    parcel in a boundary-layer eddy, T = T0 + dT (1 - cos(2 pi k / period)) / 2 and the same for p (the
    condense test drives 290 -> 280 K and 1e5 -> 8e4 Pa over 1200 s, i.e. -0.5 K and -1000 Pa per 60 s:
    partmc/test/condense/temp.dat, pressure.dat).  With RH0 = 0.95 the cells stay sub-saturated
    (RH 0.95 .. 0.98), so the cost of a step does not drift over the run."""
    out = []
    for k in range(1, n_steps + 1):
        f = 0.5 * (1.0 - np.cos(2.0 * np.pi * k / period))
        T = case["T"] + dT * f
        P = case["P"] + dp * f
        out.append((T, P, 287.04 * T / P))
    return out


def transport_fields(case, num_conc, courant_x=0.3, courant_y=0.2, courant_z=0.05, xkhh=10.0, dx=100.0, dy=100.0,
                     dt=6.0, seed=3, periodic=True):
    """WRF-side inputs of wrf_pmc_trans_driver for the doubly periodic patch of `case`, as global arrays in WRF
    memory order (i fastest, then k, then j; class slowest), all float32 like WRF's `real`:
      u_flux [(nx+1), nz, ny, C], v_flux [nx, nz, (ny+1), C], w_flux [nx, (nz+1), ny, C]  per-class advective
          fluxes of the number mixing ratio (WRFV3/dyn_em/module_advect_em.F:7806-7876), here first-order upwind
          fluxes of `num_conc` ([n_cells][C], what partmc_to_wrf hands to WRF) for a wind field with the given
          Courant numbers (a horizontally sheared flow that reverses with height, weak cellular vertical motion);
      xkhh [nx, nz, ny] horizontal eddy diffusivity, alt [nx, nz, ny] inverse density, z [nx, nz, ny] and
      z_at_w [nx, (nz+1), ny] heights, exch_h [nx, (nz+1), ny] vertical exchange coefficient."""
    nx, nz, ny, C = case["nx"], case["nz"], case["ny"], case["n_class"]
    rng = np.random.default_rng(seed)
    q = np.asarray(num_conc, np.float32).reshape(ny, nz, nx, C)  # [j][k][i][c]
    kk = np.arange(nz)[None, :, None]
    jj = np.arange(ny)[:, None, None]
    ii = np.arange(nx + 1)[None, None, :]
    # Courant numbers on the faces
    cx = courant_x * np.cos(np.pi * kk / max(nz - 1, 1)) * (1.0 + 0.2 * np.sin(2 * np.pi * jj / ny)) + 0.0 * ii
    cy = courant_y * np.sin(2 * np.pi * np.arange(nx)[None, None, :] / nx) * np.ones((ny + 1, nz, 1))
    cz = courant_z * np.sin(2 * np.pi * np.arange(nx)[None, None, :] / nx) * \
        np.sin(np.pi * np.arange(nz + 1)[None, :, None] / nz) * np.ones((ny, 1, 1))
    cz[:, 0, :] = 0.0
    cz[:, nz, :] = 0.0
    u_flux = np.zeros((C, ny, nz, nx + 1), np.float32)
    v_flux = np.zeros((C, ny + 1, nz, nx), np.float32)
    w_flux = np.zeros((C, ny, nz + 1, nx), np.float32)
    for c in range(C):
        qc = q[:, :, :, c]
        if periodic:
            left = np.concatenate([qc[:, :, -1:], qc], axis=2)   # cell i-1 of face i (periodic)
            right = np.concatenate([qc, qc[:, :, :1]], axis=2)   # cell i of face i
            south = np.concatenate([qc[-1:], qc], axis=0)
            north = np.concatenate([qc, qc[:1]], axis=0)
        else:  # open domain: the outermost cells stand for what lies beyond the domain faces
            left = np.concatenate([qc[:, :, :1], qc], axis=2)
            right = np.concatenate([qc, qc[:, :, -1:]], axis=2)
            south = np.concatenate([qc[:1], qc], axis=0)
            north = np.concatenate([qc, qc[-1:]], axis=0)
        u_flux[c] = np.where(cx > 0, cx * left, cx * right)
        v_flux[c] = np.where(cy > 0, cy * south, cy * north)
        below = np.concatenate([qc[:, :1], qc], axis=1)
        above = np.concatenate([qc, qc[:, -1:]], axis=1)
        w_flux[c] = np.where(cz > 0, cz * below, cz * above)
    alt = np.asarray(case["rrho"], np.float32).reshape(ny, nz, nx)
    kh = (xkhh * (1.0 + 0.3 * rng.random((ny, nz, nx)))).astype(np.float32)
    z_w1 = (np.arange(nz + 1) * 100.0).astype(np.float32)
    z_at_w = np.broadcast_to(z_w1[None, :, None], (ny, nz + 1, nx)).astype(np.float32).copy()
    z = (0.5 * (z_at_w[:, 1:, :] + z_at_w[:, :-1, :])).astype(np.float32)
    prof = 1.0 + 49.0 * np.sin(np.pi * np.arange(nz + 1) / nz) ** 2
    exch_h = np.broadcast_to(prof[None, :, None], (ny, nz + 1, nx)).astype(np.float32).copy()
    return dict(u_flux=u_flux, v_flux=v_flux, w_flux=w_flux, xkhh=kh, alt=alt, z=z, z_at_w=z_at_w, exch_h=exch_h,
                dx=np.float32(dx), dy=np.float32(dy), dt=np.float32(dt))


def patch_fields(f, i0, nxl, j0, nyl, periodic=True):
    """The slabs of the global transport fields a rank owning cells [i0, i0+nxl) x [j0, j0+nyl) hands to the
    library: staggered flux arrays of the patch, xkhh / alt with a one-cell halo in x and y (periodic wrap, or
    the nearest cell's value beyond the faces of an open domain)."""
    ny, nz, nx = f["alt"].shape
    ii = np.arange(i0, i0 + nxl)
    jj = np.arange(j0, j0 + nyl)
    if periodic:
        iih = np.arange(i0 - 1, i0 + nxl + 1) % nx
        jjh = np.arange(j0 - 1, j0 + nyl + 1) % ny
    else:
        iih = np.clip(np.arange(i0 - 1, i0 + nxl + 1), 0, nx - 1)
        jjh = np.clip(np.arange(j0 - 1, j0 + nyl + 1), 0, ny - 1)
    out = dict(dx=f["dx"], dy=f["dy"], dt=f["dt"])
    out["u_flux"] = np.ascontiguousarray(f["u_flux"][:, jj][:, :, :, np.arange(i0, i0 + nxl + 1)])
    out["v_flux"] = np.ascontiguousarray(f["v_flux"][:, np.arange(j0, j0 + nyl + 1)][:, :, :, ii])
    out["w_flux"] = np.ascontiguousarray(f["w_flux"][:, jj][:, :, :, ii])
    out["xkhh"] = np.ascontiguousarray(f["xkhh"][jjh][:, :, iih])
    out["alt"] = np.ascontiguousarray(f["alt"][jjh][:, :, iih])
    out["z"] = np.ascontiguousarray(f["z"][jj][:, :, ii])
    out["z_at_w"] = np.ascontiguousarray(f["z_at_w"][jj][:, :, ii])
    out["exch_h"] = np.ascontiguousarray(f["exch_h"][jj][:, :, ii])
    return out
