"""The cold part of the particle record: aero_particle_t%component(:) lists and the
opaque optics blob, carried on the device after pmcgpu_track_components.  Everything is compared with the oracle,
which carries the same lists in its Particle structs: bit for bit (sources, creation times, order of the lists).
  * coagulation concatenates / subsamples to 20 (aero_particle_coagulate, aero_particle.F90:955-972),
  * the accelerated path accumulates, resamples and merges into the target (coagulation.F90:440, 472-508, 565),
  * every record move keeps the list with its particle: sort, rebalance (doubling / halving), transport, emissions,
    dry deposition, checkpoints."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import wrf_partmc_b200 as W
    from oracle import oracle as O
    return W, O


def random_components(case, seed, n_opq, max_comp=3):
    """per particle: 1..max_comp components with random sources and creation times, n_primary = n_comp, opaque data"""
    rng = np.random.default_rng(seed)
    N = int(case["counts"].sum())
    nc = rng.integers(1, max_comp + 1, N).astype(np.int32)
    src = rng.integers(1, 9, (N, 20)).astype(np.int32)
    tm = np.round(rng.random((N, 20)) * 1000.0, 3)
    opq = rng.random((N, n_opq))
    case["n_primary"] = nc.copy()
    return nc, src, tm, opq


def upload(W, O, case, comps, n_opq, cap=None):
    nc, src, tm, opq = comps
    g = common.gpu_from_case(W, case, cell_capacity=cap)
    g.track_components(n_opq)
    g.set_components(nc, src, tm, opq if n_opq else None)
    o = common.oracle_from_case(O, case)
    off = np.concatenate([[0], np.cumsum(case["counts"])])
    for c in range(case["n_cells"]):
        a, b = off[c], off[c + 1]
        o.set_components(c, nc[a:b], src[a:b], tm[a:b], opq[a:b] if n_opq else None)
    return g, o


def assert_same_components(g, o, case, n_opq):
    nc, src, tm, opq = g.get_components()
    cnt = g.counts()
    off = np.concatenate([[0], np.cumsum(cnt)])
    longest = 0
    for c in range(case["n_cells"]):
        a, b = off[c], off[c + 1]
        onc, osrc, otm, oopq = o.get_components(c, n_opq)
        assert b - a == onc.size, c
        assert np.array_equal(nc[a:b], onc), "cell %d: component counts differ" % c
        assert np.array_equal(src[a:b], osrc), "cell %d: sources differ" % c
        assert np.array_equal(tm[a:b], otm), "cell %d: creation times differ" % c
        if n_opq:
            assert np.array_equal(opq[a:b], oopq), "cell %d: opaque optics differ" % c
        longest = max(longest, int(onc.max()) if onc.size else 0)
    return longest


def test_default_components_and_round_trip(mods):
    W, O = mods
    case = common.make_cells(4, 120, n_spec=3, n_class=2, seed=3)
    g = common.gpu_from_case(W, case)
    with pytest.raises(W.PmcGpuError):
        g.get_components()  # not tracked yet
    g.track_components(2)
    nc, src, tm, opq = g.get_components()
    assert (nc == 1).all() and np.array_equal(src[:, 0], case["wclass"]) and np.array_equal(tm[:, 0], case["lct"])
    assert (src[:, 1:] == 0).all() and (opq == 0).all()
    comps = random_components(case, 5, 2)
    g.set_components(*comps)
    nc2, src2, tm2, opq2 = g.get_components()
    assert np.array_equal(nc2, comps[0]) and np.array_equal(opq2, comps[3])
    for i in range(nc2.size):
        assert np.array_equal(src2[i, :nc2[i]], comps[1][i, :nc2[i]]) and (src2[i, nc2[i]:] == 0).all()
        assert np.array_equal(tm2[i, :nc2[i]], comps[2][i, :nc2[i]])
    # a sort permutes the lists with their particles
    P0 = g.get_particles()
    g.aero_state_sort()
    P1 = g.get_particles()
    nc3, src3, tm3, opq3 = g.get_components()
    pos = {int(i): k for k, i in enumerate(P0["id"])}
    perm = np.array([pos[int(i)] for i in P1["id"]])
    assert not np.array_equal(perm, np.arange(perm.size))
    assert np.array_equal(nc3, nc2[perm]) and np.array_equal(src3, src2[perm]) and np.array_equal(opq3, opq2[perm])


def test_num_concs_by_source(mods):
    """aero_state_num_concs_by_source (aero_state.F90:1310-1347), integer quotient n_primary_parts / n_components kept"""
    W, O = mods
    case = common.make_cells(5, 150, n_spec=2, n_class=2, seed=4)
    comps = random_components(case, 11, 0, max_comp=5)
    case["n_primary"] = (comps[0] * 2 + 1).astype(np.int32)  # quotient 2 except for one component (3)
    g, _ = upload(W, O, case, comps, 0)
    got = g.num_concs_by_source(8)
    nc, src, tm, _ = comps
    off = np.concatenate([[0], np.cumsum(case["counts"])])
    want = np.zeros((5, 8))
    for c in range(5):
        for p in range(off[c], off[c + 1]):
            num_conc = case["wmag"][c, case["wclass"][p] - 1, 0]
            for q in range(nc[p]):
                want[c, src[p, q] - 1] += num_conc * (case["n_primary"][p] // nc[p])
    np.testing.assert_allclose(got, want, rtol=1e-14)
    assert want[3].sum() == 0 and got[3].sum() == 0 and want.sum() > 0


@pytest.mark.parametrize("weight,n_class,n_group,dt,accel", [("flat", 1, 1, 3600.0, False), ("flat", 2, 1, 7200.0, True),
                                                              ("nummass", 1, 2, 3600.0, True)])
def test_coagulation_merges_component_lists(mods, weight, n_class, n_group, dt, accel):
    """several long coagulation steps with rebalancing: lists grow past 20 and are subsampled; per-set and
    accelerated path; ids / volumes stay bit-exact as without tracking"""
    W, O = mods
    n_opq = 3
    case = common.make_cells(6, 700, n_spec=3, n_class=n_class, n_group=n_group, weight=weight, seed=33, num_conc=3e11)
    comps = random_components(case, 9, n_opq, max_comp=8)
    g, o = upload(W, O, case, comps, n_opq)
    acc = events = 0
    for step in range(6):
        ns, nc = g.mc_coag(W.KERNEL_BROWN, dt, step=step)
        g.aero_state_rebalance(True, True, step=step)
        for c in range(case["n_cells"]):
            o.sort(c, force_resort=True, physical=True, edge_recheck=True)
            s, k, a = o.mc_coag(c, O.KERNEL_BROWN, dt, step=step)
            o.rebalance(c, True, True, step=step)
            acc += a
            assert (s, k) == (ns[c], nc[c]), (c, step)
        events += int(nc.sum())
        common.assert_same_population(g, o, case)
        longest = assert_same_components(g, o, case, n_opq)
    assert events > 300
    assert (acc > 0) == accel
    assert longest == 20, "the case must reach the 20-component limit (longest list %d)" % longest
    # coagulated particles carry zero optics, untouched ones their own
    nc_, src_, tm_, opq_ = g.get_components()
    P = g.get_particles()
    assert (opq_[P["n_primary"] > 8] == 0).all() and (opq_ != 0).any()


def test_components_through_the_whole_step_and_transport(mods, tmp_path):
    """partmc_timestep (coagulation, water uptake, rebalance) + emissions + dry deposition + transport on a periodic
    patch: lists equal to the oracle's after every step; checkpoint / restart keeps them"""
    W, O = mods
    from wrf_partmc_b200.host import synthetic
    nx, nz, ny, npc, S = 4, 3, 3, 250, 4
    case = synthetic.make_patch(nx, nz, ny, npc, S, seed=6, ragged=0.3)
    n_opq = 1
    comps = random_components(case, 2, n_opq)
    g, o = upload(W, O, case, comps, n_opq, cap=3 * npc + 64)
    n = case["n_cells"]
    g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
              cell_volume=case["cell_volume"])
    for c in range(n):
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
    o.trans_setup(nx, nz, ny, px=1, py=1)
    f = synthetic.transport_fields(case, o.num_conc_by_class())
    fp = synthetic.patch_fields(f, 0, nx, 0, ny)
    for c in range(n):
        o.condense_equilib(c)
    g.condense_equilib_particles()
    met = synthetic.met_sequence(case, 3)
    rng = np.random.default_rng(8)
    vf = np.zeros((2, S))
    vf[0, 0] = 1.0
    vf[1, 1] = 0.5
    vf[1, 2] = 0.5
    modes = dict(mode_type=[1, 3], weight_class=[1, case["n_class"]], char_radius=[4e-8, 1e-7], log10_std_dev=[0.2, 0.0],
                 vol_frac=vf)
    mag = case["wmag"][:, 0, 0]
    dep = dict(rmol=np.full((ny, nx), 0.01, np.float32), znt=np.full((ny, nx), 0.1, np.float32),
               ust=np.full((ny, nx), 0.4, np.float32), ivgtyp=np.full((ny, nx), 7, np.int32),
               height=np.full((ny, nx), 40.0))
    for k, (T, P, rrho) in enumerate(met):
        g.partmc_timestep(temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * 600.0, chem_dt=600.0,
                          do_coagulation=True, do_condensation=True, step=4 * k)
        o.partmc_timestep(0, n, temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * 600.0, chem_dt=600.0,
                          do_coagulation=True, do_condensation=True, step=4 * k, n_threads=2)
        num_conc = np.outer(mag * 30.0, [0.6, 0.4])
        g.aero_state_add_aero_dist_sample(num_conc=num_conc, sample_prop=np.ones(n), characteristic_factor=60.0,
                                          create_time=600.0 * (k + 1), step=4 * k + 1, **modes)
        for c in range(n):
            o.add_aero_dist_sample(c, num_conc=num_conc[c], sample_prop=1.0, characteristic_factor=60.0,
                                   create_time=600.0 * (k + 1), step=4 * k + 1, **modes)
        g.wrf_pmc_dep_dry_driver(dt=600.0, step=4 * k + 2, **dep)
        o.dep_dry_driver(nx, nz, ny, dep["rmol"], dep["znt"], dep["ust"], dep["ivgtyp"], dep["height"], 600.0,
                         step=4 * k + 2)
        g.wrf_pmc_trans_driver(fp, step=4 * k + 3)
        o.wrf_pmc_trans_driver(f, step=4 * k + 3)
        P_ = g.get_particles()
        off = np.concatenate([[0], np.cumsum(P_["count"])])
        for c in range(n):
            assert np.array_equal(P_["id"][off[c]:off[c + 1]], o.get_cell(c)["id"]), (k, c)
        assert_same_components(g, o, case, n_opq)
    nc, src, tm, opq = g.get_components()
    assert (nc > 1).any() and (tm[:, 0] >= 600.0).any(), "coagulated and emitted particles must be present"
    # checkpoint / restart
    path = str(tmp_path / "ck.bin")
    g.checkpoint_write(path)
    g2 = common.gpu_from_case(W, case, cell_capacity=3 * npc + 64)
    g2.checkpoint_read(path)
    g2.n_opaque = n_opq
    for a, b in zip(g.get_components(), g2.get_components()):
        assert np.array_equal(a, b)
