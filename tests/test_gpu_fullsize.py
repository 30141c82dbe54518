"""BASELINE.json's K4 configuration at full size (40 x 40 x 20 cells x 1000 particles, 20 species: 3.2e7
particles) through size-independent properties -- the oracle finishes cells, not 3.2e7 particles, in seconds:
  aero_state_sort     a permutation inside every cell (order-independent bit checksums of ids and of every
                      species row unchanged), (class, bin) keys non-decreasing, bins inside the cell's grid,
                      bin counts add up, idempotent;
  partmc_timestep     ids stay unique, volumes finite and non-negative, dry mass concentration conserved by
                      coagulation + condensation to the statistical bound, water present in every particle;
  trans driver        on the doubly periodic patch with uniform weights transport only moves particles: the
                      multiset of ids and the species bit checksums are unchanged.
Sub-cell parity against the oracle on the same kernels is in test_gpu_parity / test_gpu_condense / test_gpu_trans."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NX, NZ, NY, NPER, S = 40, 20, 40, 1000, 20


def _checksums(P):
    ids = P["id"].astype(np.uint64)
    out = {"n": int(P["count"].sum()), "id_sum": int(ids.sum(dtype=np.uint64)),
           "id_xor": int(np.bitwise_xor.reduce(ids))}
    bits = P["vol"].view(np.uint64)
    out["vol_xor"] = [int(np.bitwise_xor.reduce(bits[s])) for s in range(bits.shape[0])]
    out["vol_sum"] = [int(bits[s].sum(dtype=np.uint64)) for s in range(bits.shape[0])]
    return out


def test_k4_full_size_properties():
    import wrf_partmc_b200 as W
    from wrf_partmc_b200.host import synthetic
    case = synthetic.make_patch(NX, NZ, NY, NPER, S, seed=1)
    # uniform weights: transport then only moves particles (no preweighting removals or copies)
    case["wmag"][:] = case["wmag"][0, 0, 0]
    n_cells = case["n_cells"]
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    g = W.PartMCGPU(data, nx=NX, nz=NZ, ny=NY, n_group=1, n_class=case["n_class"], n_part_ideal=case["n_part_ideal"],
                    cell_capacity=int(2.5 * NPER) + 64, max_bins=64, seed=7)
    g.set_particles(case["counts"], case["vol"], case["wgroup"], case["wclass"], case["id"], case["n_primary"],
                    case["hyst"], case["lct"], case["gct"])
    g.set_weights(case["wtype"], case["wmag"], case["wexp"])
    g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
              cell_volume=case["cell_volume"])
    counts0 = case["counts"].copy()
    iw = sp["i_water"] - 1
    dens = np.array(sp["density"])
    c0 = _checksums(dict(count=counts0, id=case["id"], vol=case["vol"]))
    assert c0["n"] == int(counts0.sum()) > 3.0e7
    del case["vol"]

    # ---- aero_state_sort
    g.aero_state_sort()
    P = g.get_particles()
    assert np.array_equal(P["count"], counts0)
    c1 = _checksums(P)
    assert c1 == c0, "sort is not a permutation of the records"
    off = np.concatenate([[0], np.cumsum(P["count"])]).astype(np.int64)
    cell_of = np.repeat(np.arange(n_cells), P["count"])
    # every cell is a permutation of itself: per-cell id sums
    id_cell0 = np.add.reduceat(case["id"].astype(np.uint64), off[:-1][counts0 > 0])
    id_cell1 = np.add.reduceat(P["id"].astype(np.uint64), off[:-1][counts0 > 0])
    assert np.array_equal(id_cell0, id_cell1)
    bins = g.bins()
    nb, gmin, gmax = g.grid()
    assert bins.min() >= 1 and np.all(bins <= nb[cell_of])
    key = (P["wclass"].astype(np.int64) - 1) * 256 + bins
    d = np.diff(key)
    inner = np.ones(len(key) - 1, bool)
    inner[off[1:-1][(off[1:-1] > 0) & (off[1:-1] < len(key))] - 1] = False  # last particle of a cell -> first of the next
    assert np.all(d[inner] >= 0), "records are not in (class, bin) order"
    bc = g.bin_counts()
    assert np.array_equal(bc.reshape(n_cells, -1).sum(axis=1), counts0)
    # radius inside the bin the record carries (edges to 1e-12: the grid is built from libm logs)
    r = (3.0 * P["vol"].sum(axis=0) / (4.0 * np.pi)) ** (1.0 / 3.0)
    lo = gmin[cell_of] * (gmax[cell_of] / gmin[cell_of]) ** ((bins - 1) / nb[cell_of])
    hi = gmin[cell_of] * (gmax[cell_of] / gmin[cell_of]) ** (bins / nb[cell_of])
    assert np.all(r >= lo * (1 - 1e-12)) and np.all(r <= hi * (1 + 1e-12))
    ids_sorted = P["id"].copy()
    del r, lo, hi, key, d, inner
    g.aero_state_sort()
    assert np.array_equal(g.get_particles()["id"], ids_sorted), "sort is not idempotent"
    del ids_sorted, P

    # ---- transport on the periodic patch with uniform weights: particles only move
    f = synthetic.transport_fields(case, g.partmc_to_wrf(), dt=6.0)
    fp = synthetic.patch_fields(f, 0, NX, 0, NY)
    g.wrf_pmc_trans_driver(fp, step=1, allow_doubling=False, allow_halving=False)
    P = g.get_particles()
    assert int(g.trans_moved().sum()) > 0.2 * c0["n"]
    c2 = _checksums(P)
    assert c2 == c0, "transport with uniform weights must only move records"
    del P

    # ---- one chemistry step: equilibrate, coagulate, condense
    g.condense_equilib_particles()

    def dry_mass_conc(P):
        t, m, e = g.get_weights()
        cell = np.repeat(np.arange(n_cells), P["count"])
        nc = m[cell, P["wclass"] - 1, 0]
        mass = P["vol"] * dens[:, None]
        mass[iw] = 0.0
        return float((mass.sum(axis=0) * nc).sum())

    P = g.get_particles()
    m0 = dry_mass_conc(P)
    del P
    g.partmc_timestep(elapsed_time=0.0, dt=60.0, chem_dt=60.0, do_coagulation=True, coag_kernel_type=W.KERNEL_BROWN,
                      do_condensation=True, step=2)
    g.check_status()
    P = g.get_particles()
    assert np.all(np.isfinite(P["vol"])) and P["vol"].min() >= 0.0
    assert P["vol"][iw].min() > 0.0
    assert len(np.unique(P["id"])) == int(P["count"].sum()), "duplicate particle ids"
    assert int(P["count"].sum()) < c0["n"], "coagulation removed nothing"
    m1 = dry_mass_conc(P)
    assert abs(m1 - m0) / m0 < 1e-3, (m0, m1)
