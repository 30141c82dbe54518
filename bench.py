#!/usr/bin/env python
"""bench.py -- particle-timesteps/sec of the PartMC per-grid-cell aerosol step on B200.

One "step" = one pass of the hot path over the patch of this rank:
  partmc_timestep   (sort -> mc_coag(Brownian) -> density rescale -> condense_particles -> rebalance ->
                     partmc_to_wrf; interface/wrf_pmc_driver.F90:46-290)
  wrf_pmc_trans_driver (particle transport; interface/wrf_pmc_trans.F90:37).
Workload: BASELINE.json config K4, ONE em_les-style 40 x 40 x 20 domain with 10^3 particles per cell (3.2e7
particles, 20-species table, 2 weight classes).  At N > 1 that fixed domain is cut into px x py x-y patches, one
per GPU (WRFV3/main/module_wrf_top.F:505-510), particles crossing patch sides travel as NCCL send/recv: strong
scaling (`--scaling weak` gives every rank a whole K4 patch instead).  Before the timed region every N > 1 run
checks a small decomposed domain against the CPU oracle's emulation of the same decomposition
("multirank_parity"), and the N = 8 line carries BASELINE.json's K5 (150 x 150 x 40 x 500) as an extra key.

  value  : whole-job particle-timesteps/s with all inputs resident in HBM, CUDA events on the library's stream;
           particle-timesteps are counted on the device (particles at the start of every partmc_timestep)
  e2e    : the same through the C ABI with the per-step WRF fields in pinned HOST buffers (H2D of env fields
           and D2H of partmc_to_wrf's number concentrations inside the timed region, wall clock)
  roofline / cpu_baseline : see DESIGN.md ("Measurement").
`--impl reference` times the CPU restatement of the reference (oracle/, kind "port": the Fortran reference
cannot be built here) with the reference's integrator (the restatement of CVODE's BDF method,
condense_solver.c:131-166) on a bounded sample of the same workload with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "particle-timesteps/sec (coag+cond+transport)"
UNIT = "particle-timesteps/s"
CHEM_DT = 60.0
TRANS_DT = 6.0  # Courant numbers 0.3 / 0.2 / 0.05 at dx = dy = dz = 100 m
FK = ("u_flux", "v_flux", "w_flux", "xkhh", "alt", "z", "z_at_w", "exch_h")
PHASES = ["sort", "mc_coag(brownian)", "condense_particles", "rebalance", "partmc_to_wrf"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="N > 1: strong = the one K4 domain cut into patches; weak = one K4 patch per GPU")
    ap.add_argument("--nx", type=int, default=40)
    ap.add_argument("--ny", type=int, default=40)
    ap.add_argument("--nz", type=int, default=20)
    ap.add_argument("--n-per-cell", type=int, default=1000)
    ap.add_argument("--n-spec", type=int, default=20)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU work of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-transport", action="store_true")
    ap.add_argument("--courant-scale", type=float, default=1.0,
                    help="scales the Courant numbers 0.3 / 0.2 / 0.05 and the diffusivity of the transport call "
                         "(1/6 = em_les' own 1 s step instead of the 6 s of the benchmark step)")
    ap.add_argument("--no-parity", action="store_true", help="skip the multi-rank parity check (N > 1)")
    ap.add_argument("--k5", default="auto", choices=["auto", "on", "off"],
                    help="K5 (150x150x40 cells x 500 particles) as an extra key; auto = at N = 8")
    ap.add_argument("--condense-method", type=int, default=4, choices=[4, 5],
                    help="device integrator: 4 RODAS4 (default), 5 the BDF restatement")
    return ap.parse_args()


def workload_name(a):
    return "K4 em_les-style domain %dx%dx%d cells x %d particles/cell, S=%d species, 2 weight classes" % (
        a.nx, a.ny, a.nz, a.n_per_cell, a.n_spec)


# ----------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline: the oracle on a bounded sample of cells, all host threads
# ----------------------------------------------------------------------------------------------
def oracle_sample(case, n_sample_cells):
    from oracle import oracle as O
    sp = case["sp"]
    data = O.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    o = O.Oracle(n_sample_cells, data, n_group=1, n_class=case["n_class"], n_part_ideal=case["n_part_ideal"],
                 rng="philox", seed=case["seed"], n_cells_global=case["n_cells"])
    off = np.concatenate([[0], np.cumsum(case["counts"])])
    for c in range(n_sample_cells):
        a, b = off[c], off[c + 1]
        o.set_cell(c, case["vol"][:, a:b], case["wgroup"][a:b], case["wclass"][a:b], case["id"][a:b],
                   case["n_primary"][a:b], case["hyst"][a:b], case["lct"][a:b], case["gct"][a:b])
        o.set_weights(c, case["wtype"][c], case["wmag"][c], case["wexp"][c])
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
        o.condense_equilib(c)
    return o


def time_oracle(a, nx_s, ny_s, steps, warmup, n_threads, transport, method):
    """The CPU restatement on its own small doubly periodic domain of nx_s x nz x ny_s cells with the
    per-cell workload of the benchmark, decomposed into n_threads patches along x (one thread per patch,
    like one MPI rank per core).  method 5 = the restatement of CVODE's BDF (what the reference runs),
    4 = RODAS4 (what the device runs).  Returns (particle-timesteps/s, seconds per step, description)."""
    from wrf_partmc_b200.host import synthetic
    case = synthetic.make_patch(nx_s, a.nz, ny_s, a.n_per_cell, a.n_spec, seed=1)
    n_cells = case["n_cells"]
    o = oracle_sample(case, n_cells)
    o.set_condense_method(method)
    met = synthetic.met_sequence(case, warmup + steps)
    if transport:
        px = n_threads
        while nx_s % px:
            px -= 1
        o.trans_setup(nx_s, a.nz, ny_s, px=px, py=1)
        f = synthetic.transport_fields(case, o.num_conc_by_class(), dt=TRANS_DT, **courant_kw(a.courant_scale))
    t_total, pts = 0.0, 0
    for k in range(warmup + steps):
        T, P, rrho = met[k]
        n_now = sum(o.count(c) for c in range(n_cells))
        t0 = time.perf_counter()
        o.partmc_timestep(0, n_cells, temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * CHEM_DT,
                          chem_dt=CHEM_DT, do_coagulation=True, do_condensation=True, step=2 * k,
                          n_threads=n_threads)
        if transport:
            o.wrf_pmc_trans_driver(f, step=2 * k + 1, n_threads=n_threads)
        t1 = time.perf_counter()
        if k >= warmup:
            t_total += t1 - t0
            pts += n_now
    desc = "periodic sample domain %dx%dx%d cells (%d particles at the start) with the per-cell workload of the benchmark" % (
        nx_s, a.nz, ny_s, int(case["counts"].sum()))
    return pts / t_total, t_total / steps, desc


def sized_oracle_run(a, steps, warmup, seconds, transport, method=5, n_threads=None):
    """probe one small domain, then size ny_s so that the whole run takes about `seconds`"""
    ncores = n_threads or os.cpu_count() or 1
    nx_s = max(2, min(ncores, 16))
    v0, s0, _ = time_oracle(a, nx_s, 2, 1, 1, ncores, transport, method)
    want = seconds / max(steps + warmup, 1)
    ny_s = int(max(2, min(64, round(2 * want / max(s0, 1e-6)))))
    return time_oracle(a, nx_s, ny_s, steps, warmup, ncores, transport, method) + (ncores,)


def cpu_baseline_object(a, steps, warmup, seconds, transport):
    """the reference's algorithm on the host cores (BDF, all threads), with the RODAS4 time and the single-core
    figure beside it"""
    v, sps, desc, ncores = sized_oracle_run(a, steps, warmup, 0.6 * seconds, transport, method=5)
    out = {"value": v, "unit": UNIT, "cores": ncores, "kind": "port", "integrator": "bdf",
           "sample": desc + ", %d timed steps after %d warm-up; CPU restatement of the reference (oracle/) with the "
                            "restatement of CVODE's BDF method (condense_solver.c:131-166), Philox draws, one patch "
                            "per thread" % (steps, warmup)}
    try:
        v4, _, _, _ = sized_oracle_run(a, 1, 1, 0.2 * seconds, transport, method=4)
        out["rodas4_value"] = v4
        v1, _, d1, _ = sized_oracle_run(a, 1, 1, 0.2 * seconds, transport, method=5, n_threads=1)
        out["single_core_value"] = v1
        out["single_core_sample"] = d1
    except Exception as e:  # the extra figures must not cost the line
        out["extra_error"] = str(e)[:200]
    out["seconds_per_sample_step"] = sps
    return out


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    transport = not a.no_transport
    cb = cpu_baseline_object(a, a.steps, a.warmup, max(a.cpu_seconds, 30.0), transport)
    v = cb["value"]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": cb["seconds_per_sample_step"] * 1e3, "higher_is_better": True,
            "scaling": a.scaling if a.gpus > 1 else "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "integrator": "bdf",
            "config": {"workload": workload_name(a), "phases": PHASES + (["trans_aero"] if transport else []),
                       "transport": transport, "chem_dt_s": CHEM_DT, "sample": cb["sample"]},
            "cpu_baseline": cb,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ----------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None
        self.mark_line = 0

    def _lines(self):
        try:
            with open(self.f.name) as fh:
                return fh.read().splitlines()
        except OSError:
            return []

    def mark(self):
        """start of the timed region: the sampler was started before the warm-up steps (nvidia-smi needs a few
        hundred ms to deliver its first line; a strongly scaled timed region is shorter than that)"""
        if self.p is None:
            return
        t0 = time.time()
        while not self._lines() and time.time() - t0 < 3.0:
            time.sleep(0.02)
        self.mark_line = len(self._lines())

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, power = [], [], set(), []
        lines = self.f.read().splitlines()
        window = "timed region"
        if len(lines) > self.mark_line:
            lines = lines[self.mark_line:]
        else:  # a timed region shorter than the sampling period: the last warm-up samples (same load)
            lines = lines[-3:]
            window = "last warm-up steps (the timed region is shorter than the 50 ms sampling period)"
        for line in lines:
            q = [x.strip() for x in line.split(",")]
            if len(q) < 9:
                continue
            try:
                sm.append(float(q[1]))
                mx.append(float(q[2]))
                power.append(float(q[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), q[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "power_w_max": max(power), "samples": len(sm), "window": window}
        return out


# ----------------------------------------------------------------------------------------------
# the B200 arm
# ----------------------------------------------------------------------------------------------
def courant_kw(scale):
    return dict(courant_x=0.3 * scale, courant_y=0.2 * scale, courant_z=0.05 * scale, xkhh=10.0 * scale)


def algorithmic_bytes(kernel, n_part, n_cells, S, C, nbcap, n_moved=0):
    """ALGORITHMIC bytes one launch of `kernel` has to move ("Kernels")."""
    P = 8 * S + 44  # record: vol[S] f64, id i64, (group, class) i32, n_primary i32, hyst i32, 2 create times f64, bin i32
    if kernel == "k_sort_cells":
        return n_part * 2 * P                      # read record + write record (species rows staged once in smem)
    if kernel == "k_mc_coag":
        return n_part * (4 + 4 + 8)                # read bin + class, write list index; + 16S per sample (small)
    if kernel == "k_condense":
        return n_part * (8 * S + 4 + 8)            # read vol[S] + class, write water volume
    if kernel == "k_rebalance":
        return n_part * 4                          # read (group, class)
    if kernel == "k_num_conc_by_class":
        return n_part * 4 + n_cells * C * 4
    if kernel == "k_kernel_bounds":
        return n_cells * 8 * nbcap * nbcap / 8     # tables only for cells with a new grid (rare)
    if kernel in ("k_tr_gather", "k_tr_stage"):
        # survivors stay in place: the sampled records are staged once (k_tr_stage) and appended once behind the
        # survivors (k_tr_gather; the holes the removals left are filled from the tail: up to as many moves again)
        return n_moved * 2 * (P - 4)
    return 0


class Patch:
    """One rank's tile of a doubly periodic domain on its GPU, with the WRF-side fields of the transport driver."""

    def __init__(self, W, synthetic, decomp, torch, dist, rank, world, local, gnx, gny, nz, n_per_cell, S, scaling,
                 has_transport, seed0=1, cap_factor=2.5, condense_method=4, comm=True, courant_scale=1.0):
        self.W, self.torch, self.dist, self.decomp = W, torch, dist, decomp
        self.rank, self.world, self.local = rank, world, local
        self.dev = torch.device("cuda", local)
        px, py = decomp.patch_grid(world)
        if scaling == "weak":  # every rank one gnx x gny patch; the global domain grows with N
            rx, ry = decomp.patch_of_rank(rank, px, py)
            i0, nxl, j0, nyl = rx * gnx, gnx, ry * gny, gny
            gnx, gny = gnx * px, gny * py
        else:
            i0, nxl, j0, nyl = decomp.tile_of_rank(rank, px, py, gnx, gny)
        self.px, self.py, self.gnx, self.gny, self.nz = px, py, gnx, gny, nz
        self.tile = (i0, nxl, j0, nyl)
        case = synthetic.make_patch(nxl, nz, nyl, n_per_cell, S, seed=seed0 + rank)
        self.case = case
        self.n_cells, self.S, self.C = case["n_cells"], S, case["n_class"]
        sp = case["sp"]
        data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
        cap = int(cap_factor * n_per_cell) + 64
        g = W.PartMCGPU(data, nx=nxl, nz=nz, ny=nyl, n_group=1, n_class=self.C, n_part_ideal=case["n_part_ideal"],
                        cell_capacity=cap, max_bins=64, seed=7, device=local, global_nx=gnx, global_ny=gny,
                        i_offset=i0, j_offset=j0)
        self.g = g
        g.set_particles(case["counts"], case["vol"], case["wgroup"], case["wclass"], case["id"], case["n_primary"],
                        case["hyst"], case["lct"], case["gct"])
        case["vol"] = None  # the host copy of the particles is not needed any more
        g.set_weights(case["wtype"], case["wmag"], case["wexp"])
        g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
                  cell_volume=case["cell_volume"])
        if condense_method != 4:
            g.set_condense_method(condense_method)
        g.condense_equilib_particles()
        self.has_transport = has_transport
        self.f_patch = None
        if has_transport:
            # WRF's side of the loop: per-class advective fluxes of the number mixing ratio partmc_to_wrf reports,
            # computed on the GLOBAL domain (every rank contributes its tile), then cut back into patches
            nc_glob = decomp.gather_global(g.partmc_to_wrf().reshape(self.n_cells, self.C), gnx, nz, gny, px, py,
                                           device=self.dev)
            rrho_glob = decomp.gather_global(case["rrho"], gnx, nz, gny, px, py, device=self.dev)
            gcase = dict(nx=gnx, nz=nz, ny=gny, n_class=self.C, rrho=rrho_glob)
            f_glob = synthetic.transport_fields(gcase, nc_glob, dt=TRANS_DT, **courant_kw(courant_scale))
            self.f_patch = synthetic.patch_fields(f_glob, i0, nxl, j0, nyl)
            if world > 1 and comm:
                uid = decomp.broadcast_unique_id(W.comm_unique_id, rank, device=self.dev)
                g.comm_init(uid, world, rank, px, py)
        self.stream = torch.cuda.ExternalStream(g.stream(), device=self.dev)

    def barrier(self):
        self.g.synchronize()
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def prepare(self, synthetic, n_seq, e2e_steps):
        torch = self.torch
        met = synthetic.met_sequence(self.case, n_seq + e2e_steps)
        self.met_dev = [tuple(torch.from_numpy(x).to(self.dev) for x in m) for m in met[:n_seq]]
        self.met_pin = [tuple(torch.from_numpy(x).pin_memory() for x in m) for m in met[n_seq:]]
        self.numconc_dev = torch.zeros(self.n_cells * self.C, dtype=torch.float32, device=self.dev)
        self.numconc_pin = torch.zeros(self.n_cells * self.C, dtype=torch.float32).pin_memory()
        self.f_dev = self.f_pin = None
        if self.has_transport:
            self.f_dev = {k: torch.from_numpy(self.f_patch[k]).to(self.dev) for k in FK}
            self.f_pin = {k: torch.from_numpy(self.f_patch[k]).pin_memory() for k in FK}
            for d_ in (self.f_dev, self.f_pin):
                d_.update(dx=self.f_patch["dx"], dy=self.f_patch["dy"], dt=self.f_patch["dt"])

    def one_step(self, k, m, out, fields):
        W = self.W
        self.g.partmc_timestep(temp=m[0], pressure=m[1], inverse_density=m[2], elapsed_time=k * CHEM_DT, dt=CHEM_DT,
                               chem_dt=CHEM_DT, do_coagulation=True, coag_kernel_type=W.KERNEL_BROWN,
                               do_condensation=True, step=2 * k, num_conc_out=out, no_wait=True)
        if self.has_transport:
            self.g.wrf_pmc_trans_driver(fields, step=2 * k + 1, no_wait=True)

    def measure(self, synthetic, steps, warmup, sample_clocks=True):
        """warm-up, the device-timed region (inputs resident in HBM), then the end-to-end region (host fields)"""
        torch, g = self.torch, self.g
        n_seq = warmup + steps
        self.prepare(synthetic, n_seq, steps)
        clocks = ClockSampler(self.local) if sample_clocks else None
        for k in range(warmup):
            self.one_step(k, self.met_dev[k], self.numconc_dev, self.f_dev)
        g.check_status()
        g.work(reset=True)
        g.profile_read(reset=True)
        g.profile_enable(True)
        launches0 = g.launch_count()
        if clocks:
            clocks.mark()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for k in range(warmup, n_seq):
            self.one_step(k, self.met_dev[k], self.numconc_dev, self.f_dev)
        e1.record(self.stream)
        self.barrier()
        ms = e0.elapsed_time(e1)
        clk = clocks.stop() if clocks else None
        g.check_status()
        out = dict(ms=ms, clocks=clk, launches=g.launch_count() - launches0, prof=g.profile_read(reset=True),
                   work=g.work(reset=True), counts_end=g.counts(), cond_stats=g.condense_stats(),
                   coag_stats=g.coag_stats(), n_moved=int(g.trans_moved().sum()) if self.has_transport else 0)
        g.profile_enable(False)
        # ---- end to end through the C ABI with host buffers (wall clock)
        self.barrier()
        t0 = time.perf_counter()
        for k in range(steps):
            self.one_step(n_seq + k, self.met_pin[k % len(self.met_pin)], self.numconc_pin, self.f_pin)
        g.check_status()
        self.barrier()
        out["e2e_s"] = time.perf_counter() - t0
        out["e2e_work"] = g.work(reset=True)
        h2d = 3 * self.n_cells * 8
        if self.has_transport:
            h2d += sum(self.f_patch[k].nbytes for k in FK)
        out["h2d"], out["d2h"] = h2d, self.n_cells * self.C * 4 + 4
        return out


def multirank_parity(W, synthetic, decomp, torch, dist, rank, world, local, n_steps=3):
    """A small doubly periodic domain cut into the px x py patches of this job: `n_steps` alternating
    partmc_timestep (Brownian mc_coag + condense_particles + rebalance) and wrf_pmc_trans_driver calls on the
    GPUs, particle halos through NCCL, compared on every rank with the CPU oracle's emulation of the same
    decomposition (the oracle is the checker here, outside every timed region): counts, ids, classes, weight
    magnitudes (1e-13), bin grids and all dry species volumes bit for bit; water volumes, which come out of the
    condensation integrator, to 1e-9 relative."""
    from oracle import oracle as O
    px, py = decomp.patch_grid(world)
    gnx, gny, nz, npc, S = 4 * px, 3 * py, 4, 200, 4
    dev = torch.device("cuda", local)
    case = synthetic.make_patch(gnx, nz, gny, npc, S, seed=5)
    n_glob = case["n_cells"]
    i0, nxl, j0, nyl = decomp.tile_of_rank(rank, px, py, gnx, gny)
    pc = decomp.patch_case(case, i0, nxl, j0, nyl)
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    g = W.PartMCGPU(data, nx=nxl, nz=nz, ny=nyl, n_group=1, n_class=case["n_class"], n_part_ideal=case["n_part_ideal"],
                    cell_capacity=3 * npc + 64, seed=case["seed"], device=local, global_nx=gnx, global_ny=gny,
                    i_offset=i0, j_offset=j0)
    g.set_particles(pc["counts"], pc["vol"], pc["wgroup"], pc["wclass"], pc["id"], pc["n_primary"], pc["hyst"],
                    pc["lct"], pc["gct"])
    g.set_weights(pc["wtype"], pc["wmag"], pc["wexp"])
    g.set_env(temp=pc["T"], pressure=pc["P"], rel_humid=pc["RH"], inverse_density=pc["rrho"],
              cell_volume=pc["cell_volume"])
    nid = int(case["counts"].sum()) // n_glob + 2
    g.set_next_id(np.full(pc["n_cells"], nid, np.int64))
    g.condense_equilib_particles()
    uid = decomp.broadcast_unique_id(W.comm_unique_id, rank, device=dev)
    g.comm_init(uid, world, rank, px, py)
    # the oracle: the whole small domain on every rank (deterministic, so all ranks hold the same answer)
    odata = O.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    o = O.Oracle(n_glob, odata, n_group=1, n_class=case["n_class"], n_part_ideal=case["n_part_ideal"], rng="philox",
                 seed=case["seed"])
    off = np.concatenate([[0], np.cumsum(case["counts"])])
    for c in range(n_glob):
        a, b = off[c], off[c + 1]
        o.set_cell(c, case["vol"][:, a:b], case["wgroup"][a:b], case["wclass"][a:b], case["id"][a:b],
                   case["n_primary"][a:b], case["hyst"][a:b], case["lct"][a:b], case["gct"][a:b])
        o.set_weights(c, case["wtype"][c], case["wmag"][c], case["wexp"][c])
        o.set_env(c, T=case["T"][c], p=case["P"][c], RH=case["RH"][c], rrho=case["rrho"][c],
                  cell_volume=case["cell_volume"][c])
        o.set_next_id(c, nid)
        o.condense_equilib(c)
    o.trans_setup(gnx, nz, gny, px=px, py=py)
    f = synthetic.transport_fields(case, o.num_conc_by_class())
    fp = synthetic.patch_fields(f, i0, nxl, j0, nyl)
    met = synthetic.met_sequence(case, n_steps)
    cells = pc["cells"]
    moved = 0
    for k in range(n_steps):
        T, P, rrho = met[k]
        g.partmc_timestep(temp=T[cells], pressure=P[cells], inverse_density=rrho[cells], elapsed_time=k * CHEM_DT,
                          dt=CHEM_DT, chem_dt=CHEM_DT, do_coagulation=True, coag_kernel_type=W.KERNEL_BROWN,
                          do_condensation=True, step=2 * k)
        g.wrf_pmc_trans_driver(fp, step=2 * k + 1)
        moved += int(g.trans_moved().sum())
        o.partmc_timestep(0, n_glob, temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * CHEM_DT,
                          chem_dt=CHEM_DT, do_coagulation=True, do_condensation=True, step=2 * k, n_threads=4)
        o.wrf_pmc_trans_driver(f, step=2 * k + 1, n_threads=4)
    Pg = g.get_particles()
    _, wm, _ = g.get_weights()
    nb, _, _ = g.grid()
    iw = sp["i_water"] - 1
    dry = [s for s in range(S) if s != iw]
    goff = np.concatenate([[0], np.cumsum(Pg["count"])])
    bad, n_checked, max_water = [], 0, 0.0
    for lc, c in enumerate(cells):
        q = o.get_cell(int(c))
        a, b = goff[lc], goff[lc + 1]
        ok = (b - a == q["vol"].shape[1])
        if ok:
            ok = (np.array_equal(Pg["id"][a:b], q["id"]) and np.array_equal(Pg["wclass"][a:b], q["wclass"]) and
                  np.array_equal(Pg["vol"][dry, a:b], q["vol"][dry]) and int(nb[lc]) == int(o.grid(int(c))[0]) and
                  np.allclose(wm[lc].ravel(), o.get_weights(int(c))[1], rtol=1e-13, atol=0))
        if ok and b > a:
            w0, w1 = Pg["vol"][iw, a:b], q["vol"][iw]
            rel = float(np.max(np.abs(w0 - w1) / np.maximum(np.abs(w1), 1e-300)))
            max_water = max(max_water, rel)
            ok = rel <= 1e-9
        if not ok:
            bad.append(int(c))
        n_checked += int(b - a)
    g.comm_destroy()
    g.close()
    t = torch.tensor([len(bad), n_checked, moved], dtype=torch.float64, device=dev)
    mw = torch.tensor([max_water], dtype=torch.float64, device=dev)
    dist.all_reduce(t)
    dist.all_reduce(mw, op=dist.ReduceOp.MAX)
    n_bad, n_chk, n_moved = (int(x) for x in t.cpu())
    return {"ranks": world, "patches": "%dx%d" % (px, py), "cells": n_glob, "steps": n_steps,
            "particles_checked": n_chk, "particles_moved": n_moved, "cells_mismatched": n_bad,
            "bit_exact": n_bad == 0 and n_moved > 0,
            "what": "ids, classes, dry species volumes, bin grids bit for bit, weights 1e-13, water volume 1e-9 "
                    "against the CPU oracle's emulated decomposition",
            "water_volume_max_rel_diff": float(mw.cpu()[0]), "first_bad_cells_rank0": bad[:8]}


def run_b200(a):
    import torch
    import torch.distributed as dist
    import wrf_partmc_b200 as W
    from wrf_partmc_b200.host import decomp, synthetic

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the aerosol step has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    has_transport = not a.no_transport
    S = a.n_spec

    parity = None
    if world > 1 and not a.no_parity:
        parity = multirank_parity(W, synthetic, decomp, torch, dist, rank, world, local)
        if not parity["bit_exact"]:
            if rank == 0:
                sys.stderr.write("bench.py: multi-rank parity FAILED: %s\n" % json.dumps(parity))
            dist.barrier()
            dist.destroy_process_group()
            raise SystemExit(3)

    scaling = a.scaling
    pt = Patch(W, synthetic, decomp, torch, dist, rank, world, local, a.nx, a.ny, a.nz, a.n_per_cell, S, scaling,
               has_transport, condense_method=a.condense_method, courant_scale=a.courant_scale)
    g = pt.g
    n_cells, n_class = pt.n_cells, pt.C
    m = pt.measure(synthetic, a.steps, a.warmup)
    ms = m["ms"]
    # ---- max over ranks of the times, sum over ranks of the particle-timesteps (counted on the device)
    t, work = decomp.reduce_job([ms * 1e-3, m["e2e_s"]],
                                [m["work"]["particle_timesteps"], m["e2e_work"]["particle_timesteps"],
                                 m["work"]["cond_rate_evals"], m["work"]["cond_particles"]], device=dev)
    t_dev, t_e2e = t
    value = work[0] / t_dev
    e2e = work[1] / t_e2e
    pts_rank = m["work"]["particle_timesteps"]          # this rank
    n_mean = pts_rank / a.steps                           # mean population of this rank's patch at step start
    prof = m["prof"]
    cnt_end = m["counts_end"]
    cst = m["cond_stats"]
    csamp, ccoag = m["coag_stats"]

    k5 = None
    want_k5 = (a.k5 == "on") or (a.k5 == "auto" and world == 8)
    fp64_peak = g.measure_fp64_peak() if rank == 0 else None
    if want_k5:
        g.comm_destroy()
        g.close()
        pt = None
        torch.cuda.empty_cache()
        try:
            p5 = Patch(W, synthetic, decomp, torch, dist, rank, world, local, 150, 150, 40, 500, 20, "strong",
                       has_transport, seed0=101)
            m5 = p5.measure(synthetic, 3, 2, sample_clocks=False)
            t5, w5 = decomp.reduce_job([m5["ms"] * 1e-3, m5["e2e_s"]],
                                       [m5["work"]["particle_timesteps"], m5["e2e_work"]["particle_timesteps"]],
                                       device=dev)
            k5 = {"workload": "K5 regional domain 150x150x40 cells x 500 particles/cell, S=20, doubly periodic, cut into "
                              "%dx%d patches" % (p5.px, p5.py),
                  "value": w5[0] / t5[0], "unit": UNIT, "ms_per_step": t5[0] * 1e3 / 3, "steps": 3, "warmup": 2,
                  "particles": w5[0] / 3, "e2e_value": w5[1] / t5[1]}
            p5.g.comm_destroy()
            p5.g.close()
        except Exception as e:  # K5 must not cost the K4 line
            k5 = {"error": str(e)[:300]}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (
            6650.0, "fallback (B200_PROFILING.md)")
        kern = []
        for name, (kms, nl) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
            per_launch_s = kms * 1e-3 / max(nl, 1)
            by = algorithmic_bytes(name, n_mean, n_cells, S, n_class, 64, n_moved=m.get("n_moved", 0))
            kern.append({"kernel": name, "ms_per_step": kms / a.steps, "share": kms / ms, "launches": nl,
                         "alg_bytes_per_launch": by, "hbm_gbs": by / per_launch_s / 1e9 if per_launch_s > 0 else None})
        # FP64 work of the condensation integrator, counted on the device: sum over cells and steps of
        # (rate evaluations of the cell x particles of the cell) on THIS rank
        evals_per_step = m["work"]["cond_rate_evals"] / a.steps
        cond_particles_per_step = m["work"]["cond_particles"] / a.steps
        # counted by ncu on this kernel (profiles/): FP64 thread instructions per rate evaluation and particle,
        # integrator arithmetic included
        FLOP_PER_RATE_EVAL = 187.3   # DFMA = 2 flop
        FP64_INST_PER_RATE_EVAL = 128.8
        cond_ms = prof.get("k_condense", (0.0, 1))[0] / a.steps
        cond_tflops = evals_per_step * FLOP_PER_RATE_EVAL / (cond_ms * 1e-3) / 1e12 if cond_ms > 0 else None
        cond_pipe = (evals_per_step * FP64_INST_PER_RATE_EVAL / (cond_ms * 1e-3) / (fp64_peak * 1e12 / 2.0)
                     if cond_ms > 0 and fp64_peak else None)
        dom = kern[0] if kern else None
        if dom and dom["kernel"] == "k_condense":
            roofline = {"kernel": "k_condense", "bound": "fp64", "achieved": cond_tflops, "peak": fp64_peak,
                        "unit": "TFLOP/s", "frac": cond_tflops / fp64_peak if fp64_peak else None,
                        "traffic": 173.7 * cond_particles_per_step,  # ncu r02_c4: 1.272 GB read + 63 MB written for 7.686e6 particles
                        "traffic_unit": "bytes per launch (ncu capture, scaled by the particle count)",
                        "peak_source": "FP64 FMA microbenchmark in this run (pmcgpu_measure_fp64_peak); "
                                       "MEASURED_PEAKS.json holds no FP64 figure",
                        "fp64_pipe_slots_used": cond_pipe,
                        "rate_evals_x_particles_per_step": evals_per_step,
                        "note": "the dominant kernel is bounded by the FP64 pipe, not by HBM or the tensor cores "
                                "(north star: no dense contraction on the path); achieved = (rate evaluations x "
                                "particles, counted on the device in the timed region) x %.1f flop (ncu-counted, "
                                "DFMA = 2) / kernel time; fp64_pipe_slots_used counts DFMA, DMUL and DADD alike "
                                "(%.1f per evaluation) against peak / 2 instructions/s" % (
                                    FLOP_PER_RATE_EVAL, FP64_INST_PER_RATE_EVAL)}
        elif dom:
            roofline = {"kernel": dom["kernel"], "bound": "hbm", "achieved": dom["hbm_gbs"], "peak": hbm_peak,
                        "unit": "GB/s", "frac": dom["hbm_gbs"] / hbm_peak, "traffic": None, "peak_source": hbm_src}
        else:
            roofline = None
        sortk = next((k for k in kern if k["kernel"] == "k_sort_cells"), None)
        roofline_hbm = None
        if sortk:
            roofline_hbm = {"kernel": "k_sort_cells", "bound": "hbm", "achieved": sortk["hbm_gbs"], "peak": hbm_peak,
                            "unit": "GB/s", "frac": sortk["hbm_gbs"] / hbm_peak,
                            "traffic": 397.0 * n_mean, "traffic_unit": "bytes per launch (ncu capture, scaled)",
                            "peak_source": hbm_src}
        wl = workload_name(a)
        if world > 1:
            wl += (", cut into %dx%d x-y patches (one per GPU)" % decomp.patch_grid(world) if scaling == "strong"
                   else ", one such patch per GPU")
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": t_dev * 1e3 / a.steps, "higher_is_better": True,
                "scaling": scaling if world > 1 else "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": wl, "phases": PHASES + (["trans_aero"] if has_transport else []),
                           "transport": bool(has_transport), "chem_dt_s": CHEM_DT,
                           "mean_particles_per_cell": work[0] / a.steps / (
                               a.nx * a.ny * a.nz * (world if scaling == "weak" else 1)),
                           "particles_per_step_all_gpus": work[0] / a.steps,
                           "particles_per_gpu": n_mean,
                           "cell_population_end": {"max": int(cnt_end.max()), "over_1024": float((cnt_end > 1024).mean())},
                           "l2": "inputs (%.1f GB of particle state per GPU) larger than L2" % (
                               n_mean * (8 * S + 44) / 1e9),
                           "particle_count": "particles at the start of every partmc_timestep, counted on the device",
                           "coag_samples_per_step": int(csamp.sum()), "coag_events_per_step": int(ccoag.sum()),
                           "condense_integrator": "rodas4" if a.condense_method == 4 else "bdf",
                           "condense_steps_per_cell": float(cst[:, 0].mean()),
                           "condense_rate_evals_per_particle": m["work"]["cond_rate_evals"] / max(m["work"]["cond_particles"], 1)},
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                        "ms_per_step": t_e2e * 1e3 / a.steps,
                        "what": "pmcgpu_partmc_timestep + pmcgpu_wrf_pmc_trans_driver through the C ABI: T, p, 1/rho "
                                "and the transport fields from pinned host memory, partmc_to_wrf number "
                                "concentrations back to host, status read; particles stay resident (DESIGN.md "
                                "'ownership'); byte counts per GPU"},
                "gpu_launches": int(m["launches"]), "roofline": roofline, "roofline_hbm": roofline_hbm, "kernels": kern,
                "fp64_peak_tflops": fp64_peak, "clocks": m["clocks"]}
        if parity is not None:
            line["multirank_parity"] = parity
        if k5 is not None:
            line["k5"] = k5
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_object(a, 2, 1, a.cpu_seconds, has_transport)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """the ONE JSON line, on the process's real stdout"""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    a = parse()
    # libraries (NCCL's version banner, torch warnings) write to fd 1: send everything but the JSON line to stderr
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
