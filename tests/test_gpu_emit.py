"""GPU parity test for emission / inflow sampling:
aero_state_add_aero_dist_sample (aero_state.F90:758-831) with aero_state_prepare_weight_for_add (:703-757)."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def modes(n_spec, n_class):
    rng = np.random.default_rng(4)
    vf = rng.random((3, n_spec)) + 0.1
    vf[0, : max(1, n_spec // 2)] = 0.0  # species absent from a mode
    vf /= vf.sum(axis=1, keepdims=True)
    vstd = np.zeros_like(vf)
    vstd[2] = 0.2 * vf[2]
    return dict(mode_type=[1, 3, 1], weight_class=[1, n_class, n_class], char_radius=[3e-8, 2e-7, 9e-8],
                log10_std_dev=[0.25, 0.0, 0.18], vol_frac=vf, vol_frac_std=vstd)


@pytest.mark.parametrize("weight,n_spec,n_class,n_group", [("flat", 3, 2, 1), ("nummass", 4, 1, 2)])
def test_add_aero_dist_sample_matches_oracle(mods, weight, n_spec, n_class, n_group):
    """Same number of new particles per cell, same ids / classes / groups / order, same weight rescaling
    (both branches of aero_state_scale_weight are hit: cells that fill up and cells that start nearly
    empty); volumes to 1e-13 (the radii go through pow / log / cos / erf of two different math libraries)."""
    W, O = mods
    n_cells = 6
    case = common.make_cells(n_cells, 400, n_spec=n_spec, n_class=n_class, n_group=n_group, weight=weight, seed=8,
                             n_part_ideal=400.0)
    case["counts"][1] = 10  # nearly empty cell: the new particles dominate, weights go down
    # rebuild the flat arrays for the shortened cell 1
    off = np.concatenate([[0], np.cumsum(common.make_cells(n_cells, 400, n_spec=n_spec, n_class=n_class,
                                                            n_group=n_group, weight=weight, seed=8)["counts"])])
    idx = np.concatenate([np.arange(off[c], off[c] + case["counts"][c]) for c in range(n_cells)])
    for k in ("wclass", "wgroup", "id", "n_primary", "hyst", "lct", "gct"):
        case[k] = case[k][idx]
    case["vol"] = case["vol"][:, idx]
    g = common.gpu_from_case(W, case, cell_capacity=4000)
    o = common.oracle_from_case(O, case)
    m = modes(n_spec, n_class)
    mag = case["wmag"][:, 0, 0]
    total = 0
    for step in range(3):
        # number concentrations chosen so that cells receive from ~0 to ~3 x ideal new particles
        per_cell = np.array([0.0, 600.0, 40.0, 1300.0, 150.0, 5.0]) * (1.0 + 0.5 * step)
        num_conc = np.outer(per_cell * mag, [0.5, 0.2, 0.3])
        sample_prop = np.full(n_cells, 1.0)
        n_dev = g.aero_state_add_aero_dist_sample(num_conc=num_conc, sample_prop=sample_prop,
                                                  characteristic_factor=60.0, create_time=100.0 * step, step=step, **m)
        n_ora = [o.add_aero_dist_sample(c, num_conc=num_conc[c], sample_prop=sample_prop[c], characteristic_factor=60.0,
                                        create_time=100.0 * step, step=step, **m) for c in range(n_cells)]
        assert np.array_equal(n_dev, n_ora), (step, n_dev, n_ora)
        total += int(n_dev.sum())
        common.assert_same_population(g, o, case, exact_fp=False, rtol=1e-13)
        assert np.array_equal(g.get_next_id(), [o.L.pmco_get_next_id(o.h, c) for c in range(n_cells)])
    assert total > 1000
    t0, m0, e0 = case["wtype"], case["wmag"], case["wexp"]
    _, m1, _ = g.get_weights()
    assert (m1 != m0).any(), "some weights must have been rescaled"


@pytest.mark.parametrize("weight,n_group", [("flat", 1), ("nummass", 2)])
def test_sampled_mode_matches_oracle(mods, weight, n_group):
    """mode_type 4 (tabulated number concentrations per radius bin, aero_mode.F90:456-491, 533-538, 581-608) next to a
    log-normal mode: counts, ids, order equal to the oracle, volumes 1e-12; flat and power weights (the power weight
    samples the radius uniformly in 1 / num_conc and changes the expected count per bin)"""
    W, O = mods
    n_cells, S = 5, 3
    case = common.make_cells(n_cells, 300, n_spec=S, n_class=2, n_group=n_group, weight=weight, seed=14,
                             n_part_ideal=600.0)
    g = common.gpu_from_case(W, case, cell_capacity=4000)
    o = common.oracle_from_case(O, case)
    nb = 7
    edges = np.array([np.logspace(-8.3, -6.5, nb + 1), np.logspace(-8, -6, nb + 1)])
    mag = case["wmag"][:, 0, 0]
    table = np.array([[0.0] * nb, [1.0, 3.0, 5.0, 2.0, 0.0, 1.0, 0.5]]) * 40.0 * mag[0]
    vf = np.array([[0.5, 0.5, 0.0], [0.2, 0.0, 0.8]])
    m = dict(mode_type=[1, 4], weight_class=[1, 2], char_radius=[3e-8, 0.0], log10_std_dev=[0.2, 0.0], vol_frac=vf,
             sample_radius=edges, sample_num_conc=table)
    total = 0
    for step in range(3):
        num_conc = np.stack([mag * 80.0, np.full(n_cells, 1.0 + 0.5 * step)], axis=1)
        sp = np.linspace(0.5, 1.5, n_cells)
        nd = g.aero_state_add_aero_dist_sample(num_conc=num_conc, sample_prop=sp, characteristic_factor=30.0,
                                               create_time=60.0 * step, step=step, **m)
        no = [o.add_aero_dist_sample(c, num_conc=num_conc[c], sample_prop=sp[c], characteristic_factor=30.0,
                                     create_time=60.0 * step, step=step, **m) for c in range(n_cells)]
        assert np.array_equal(nd, no), (step, nd, no)
        total += int(nd.sum())
        common.assert_same_population(g, o, case, exact_fp=False, rtol=1e-12)
    assert total > 1500
    with pytest.raises(W.PmcGpuError):  # a sampled mode without its table
        g.aero_state_add_aero_dist_sample(mode_type=[4], weight_class=[1], char_radius=[0.0], log10_std_dev=[0.0],
                                          vol_frac=[[1.0, 0.0, 0.0]], num_conc=np.ones((n_cells, 1)),
                                          sample_prop=np.ones(n_cells), characteristic_factor=30.0)
