"""pmcgpu_partmc_process_static: the allocation-free variant of pmcgpu_partmc_process (no counterpart in the reference;
its result table lives in a static device array of PMCGPU_PROCESS_STATIC_MAX_DOUBLES doubles).  Called through the
C ABI and compared with the existing entry point at the tolerance of the existing diagnostics test (1e-12: the
histogram bins are summed by shared-memory atomics whose order is not fixed), on two calls in a row with different
states and table layouts, so that stale contents of the shared array would show.  Collected last on purpose."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

KEYS = ("tot_num_conc", "tot_wet_num_conc", "tot_mass_conc", "num_dist", "mass_dist", "num_conc_by_class",
        "species_mass_conc")


def _same(a, b):
    for k in KEYS:
        assert a[k].shape == b[k].shape
        np.testing.assert_allclose(a[k], b[k], rtol=1e-12, atol=0, err_msg=k)


def test_partmc_process_static_matches_the_allocating_entry_point():
    import wrf_partmc_b200 as W
    # first call: 5 cells (one empty), 4 species, 2 classes, 100 bins
    case1 = common.make_cells(n_cells=5, n_per_cell=400, n_spec=4, n_class=2, seed=12)
    g1 = common.gpu_from_case(W, case1)
    ref1 = g1.partmc_process(n_bin=100, d_min=1e-9, d_max=1e-3)
    got1 = g1.partmc_process(n_bin=100, d_min=1e-9, d_max=1e-3, static=True)
    _same(got1, ref1)
    assert ref1["tot_num_conc"][3] == 0.0 and got1["tot_num_conc"][3] == 0.0 and ref1["tot_num_conc"].sum() > 0
    # second call in a row: another context, fewer cells and bins, so every offset in the shared table differs and
    # most of the first call's contents are still in the array behind it
    case2 = common.make_cells(n_cells=3, n_per_cell=250, n_spec=20, n_class=3, seed=7, ragged=False)
    g2 = common.gpu_from_case(W, case2)
    ref2 = g2.partmc_process(n_bin=40, d_min=1e-9, d_max=1e-4)
    got2 = g2.partmc_process(n_bin=40, d_min=1e-9, d_max=1e-4, static=True)
    _same(got2, ref2)
    # and the first context again, after coagulation changed its state
    g1.mc_coag(W.KERNEL_BROWN, 600.0, step=1)
    _same(g1.partmc_process(static=True), g1.partmc_process())


def test_partmc_process_static_rejects_tables_beyond_the_static_array():
    import wrf_partmc_b200 as W
    case = common.make_cells(n_cells=300, n_per_cell=4, n_spec=1, n_class=1, seed=3, ragged=False)
    g = common.gpu_from_case(W, case)
    # 300 cells x (3 + 2 x 2000 + 1 + 1) doubles = 1 201 500 > 2^20
    with pytest.raises(W.PmcGpuError) as ei:
        g.partmc_process(n_bin=2000, static=True)
    assert ei.value.code == 3 and "static scratch" in str(ei.value)
    ok = g.partmc_process(n_bin=1000, static=True)  # 300 x 2005 = 601 500 fits
    _same(ok, g.partmc_process(n_bin=1000))
