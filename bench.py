#!/usr/bin/env python
"""bench.py -- particle-timesteps/sec of the PartMC per-grid-cell aerosol step on B200.

One "step" = one pass of the hot path over the patch of this rank:
  partmc_timestep   (sort -> mc_coag(Brownian) -> density rescale -> condense_particles -> rebalance ->
                     partmc_to_wrf; interface/wrf_pmc_driver.F90:46-290)
  wrf_pmc_trans_driver (particle transport; interface/wrf_pmc_trans.F90:37) when the library exports it.
Workload at N = 1: BASELINE.json config K4, an em_les-style 40 x 40 x 20 patch with 10^3 particles per cell
(3.2e7 particles, 20-species table, 2 weight classes); at N > 1 every rank holds one such patch (weak scaling).

  value  : whole-job particle-timesteps/s with all inputs resident in HBM, CUDA events on the library's stream
  e2e    : the same through the C ABI with the per-step WRF fields in pinned HOST buffers (H2D of env fields
           and D2H of partmc_to_wrf's number concentrations inside the timed region, wall clock)
  roofline / cpu_baseline : see DESIGN.md ("Measurement").
`--impl reference` times the CPU restatement of the reference (oracle/, kind "port": the Fortran reference
cannot be built here) on a bounded sample of the same workload with all host threads.
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nx", type=int, default=40)
    ap.add_argument("--ny", type=int, default=40)
    ap.add_argument("--nz", type=int, default=20)
    ap.add_argument("--n-per-cell", type=int, default=1000)
    ap.add_argument("--n-spec", type=int, default=20)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU work of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-transport", action="store_true")
    return ap.parse_args()


def workload_name(a, n_gpus):
    base = "K4 em_les-style patch %dx%dx%d cells x %d particles/cell, S=%d species, 2 weight classes" % (
        a.nx, a.ny, a.nz, a.n_per_cell, a.n_spec)
    return base + (" per GPU" if n_gpus > 1 else "")


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


def time_oracle(a, nx_s, ny_s, steps, warmup, n_threads, transport):
    """The CPU restatement on its own small doubly periodic domain of nx_s x nz x ny_s cells with the
    per-cell workload of the benchmark, decomposed into n_threads patches along x (one thread per patch,
    like one MPI rank per core).  Returns (particle-timesteps/s, seconds per step, description)."""
    from wrf_partmc_b200.host import synthetic
    case = synthetic.make_patch(nx_s, a.nz, ny_s, a.n_per_cell, a.n_spec, seed=1)
    n_cells = case["n_cells"]
    o = oracle_sample(case, n_cells)
    met = synthetic.met_sequence(case, warmup + steps)
    if transport:
        px = n_threads
        while nx_s % px:
            px -= 1
        o.trans_setup(nx_s, a.nz, ny_s, px=px, py=1)
        f = synthetic.transport_fields(case, o.num_conc_by_class(), dt=TRANS_DT)
    n_part = int(case["counts"].sum())
    t_total = 0.0
    for k in range(warmup + steps):
        T, P, rrho = met[k]
        t0 = time.perf_counter()
        o.partmc_timestep(0, n_cells, temp=T, pressure=P, inverse_density=rrho, elapsed_time=k * CHEM_DT,
                          chem_dt=CHEM_DT, do_coagulation=True, do_condensation=True, step=2 * k,
                          n_threads=n_threads)
        if transport:
            o.wrf_pmc_trans_driver(f, step=2 * k + 1, n_threads=n_threads)
        t1 = time.perf_counter()
        if k >= warmup:
            t_total += t1 - t0
    desc = "periodic sample domain %dx%dx%d cells (%d particles) with the per-cell workload of the benchmark" % (
        nx_s, a.nz, ny_s, n_part)
    return n_part * steps / t_total, t_total / steps, desc


def sized_oracle_run(a, steps, warmup, seconds, transport):
    """probe one small domain, then size ny_s so that the whole run takes about `seconds`"""
    ncores = os.cpu_count() or 1
    nx_s = max(2, min(ncores, 16))
    v0, s0, _ = time_oracle(a, nx_s, 2, 1, 1, ncores, transport)
    want = seconds / max(steps + warmup, 1)
    ny_s = int(max(2, min(64, round(2 * want / max(s0, 1e-6)))))
    return time_oracle(a, nx_s, ny_s, steps, warmup, ncores, transport) + (ncores,)


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    transport = not a.no_transport
    v, sps, desc, ncores = sized_oracle_run(a, a.steps, a.warmup, a.cpu_seconds, transport)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a, a.gpus), "sample": desc, "transport": transport},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": ncores, "kind": "port",
                             "sample": desc + "; CPU restatement of the reference (oracle/), Philox draws, "
                                              "one patch per thread"},
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
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

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
        for line in self.f.read().splitlines():
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
                   "power_w_max": max(power), "samples": len(sm)}
        return out


# ----------------------------------------------------------------------------------------------
# the B200 arm
# ----------------------------------------------------------------------------------------------
def algorithmic_bytes(kernel, n_part, n_cells, S, C, nbcap):
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
    return 0


def run_b200(a):
    import torch
    import torch.distributed as dist
    import wrf_partmc_b200 as W
    from wrf_partmc_b200.host import synthetic

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the aerosol step has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    S = a.n_spec
    case = synthetic.make_patch(a.nx, a.nz, a.ny, a.n_per_cell, S, seed=1 + rank)
    n_cells = case["n_cells"]
    sp = case["sp"]
    data = W.AeroData(sp["density"], sp["kappa"], sp["molec_weight"], sp["num_ions"], sp["i_water"])
    cap = int(2.5 * a.n_per_cell) + 64
    # weak scaling: one K4 patch per GPU, patches arranged px x py like WRF's processor grid
    from wrf_partmc_b200.host import decomp
    px, py = decomp.patch_grid(world)
    ext = decomp.patch_extents(rank, px, py, a.nx, a.ny)
    g = W.PartMCGPU(data, nx=a.nx, nz=a.nz, ny=a.ny, n_group=1, n_class=case["n_class"],
                    n_part_ideal=case["n_part_ideal"], cell_capacity=cap, max_bins=64, seed=7, device=local, **ext)
    g.set_particles(case["counts"], case["vol"], case["wgroup"], case["wclass"], case["id"], case["n_primary"],
                    case["hyst"], case["lct"], case["gct"])
    g.set_weights(case["wtype"], case["wmag"], case["wexp"])
    g.set_env(temp=case["T"], pressure=case["P"], rel_humid=case["RH"], inverse_density=case["rrho"],
              cell_volume=case["cell_volume"])
    g.condense_equilib_particles()
    has_transport = not a.no_transport
    if has_transport:
        # WRF's side of the loop: per-class advective fluxes of the number mixing ratio partmc_to_wrf reports
        f_glob = synthetic.transport_fields(case, g.partmc_to_wrf(), dt=TRANS_DT)
        f_patch = synthetic.patch_fields(f_glob, 0, a.nx, 0, a.ny)
        if world > 1:
            uid = decomp.broadcast_unique_id(W.comm_unique_id, rank, device=torch.device("cuda", local))
            g.comm_init(uid, world, rank, px, py)

    n_seq = a.warmup + a.steps
    met = synthetic.met_sequence(case, 2 * n_seq)
    dev = torch.device("cuda", local)
    met_dev = [tuple(torch.from_numpy(x).to(dev) for x in m) for m in met[:n_seq]]
    met_pin = [tuple(torch.from_numpy(x).pin_memory() for x in m) for m in met[n_seq:]]
    numconc_dev = torch.zeros(n_cells * case["n_class"], dtype=torch.float32, device=dev)
    numconc_pin = torch.zeros(n_cells * case["n_class"], dtype=torch.float32).pin_memory()
    stream = torch.cuda.ExternalStream(g.stream(), device=dev)
    FK = ("u_flux", "v_flux", "w_flux", "xkhh", "alt", "z", "z_at_w", "exch_h")
    if has_transport:
        f_dev = {k: torch.from_numpy(f_patch[k]).to(dev) for k in FK}
        f_pin = {k: torch.from_numpy(f_patch[k]).pin_memory() for k in FK}
        for d_ in (f_dev, f_pin):
            d_.update(dx=f_patch["dx"], dy=f_patch["dy"], dt=f_patch["dt"])

    def barrier():
        g.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def one_step(k, m, out, fields):
        g.partmc_timestep(temp=m[0], pressure=m[1], inverse_density=m[2], elapsed_time=k * CHEM_DT, dt=CHEM_DT,
                          chem_dt=CHEM_DT, do_coagulation=True, coag_kernel_type=W.KERNEL_BROWN,
                          do_condensation=True, step=2 * k, num_conc_out=out, no_wait=True)
        if has_transport:
            g.wrf_pmc_trans_driver(fields, step=2 * k + 1, no_wait=True)

    # ---- warm-up, then the timed region (device-resident inputs)
    for k in range(a.warmup):
        one_step(k, met_dev[k], numconc_dev, f_dev if has_transport else None)
    g.check_status()
    n_start = int(g.counts().sum())
    g.profile_read(reset=True)
    g.profile_enable(True)
    launches0 = g.launch_count()
    barrier()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(a.warmup, n_seq):
        one_step(k, met_dev[k], numconc_dev, f_dev if has_transport else None)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    g.check_status()
    launches = g.launch_count() - launches0
    prof = g.profile_read(reset=True)
    g.profile_enable(False)
    cnt_end = g.counts()
    n_end = int(cnt_end.sum())
    cst = g.condense_stats()
    csamp, ccoag = g.coag_stats()
    n_avg = 0.5 * (n_start + n_end)

    # ---- end to end through the C ABI with host buffers (wall clock)
    barrier()
    t0 = time.perf_counter()
    for k in range(a.steps):
        kk = n_seq + k
        one_step(kk, met_pin[k % len(met_pin)], numconc_pin, f_pin if has_transport else None)
    g.check_status()
    barrier()
    e2e_s = time.perf_counter() - t0
    n_end2 = int(g.counts().sum())
    h2d = 3 * n_cells * 8
    if has_transport:
        h2d += sum(f_patch[k].nbytes for k in FK)
    d2h = n_cells * case["n_class"] * 4 + 4

    # ---- max over ranks, sum of particles
    t, npart = decomp.reduce_job([ms * 1e-3, e2e_s], [n_avg, 0.5 * (n_end + n_end2)], device=dev)
    t_dev, t_e2e = t
    value = npart[0] * a.steps / t_dev
    e2e = npart[1] * a.steps / t_e2e

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (
            6650.0, "fallback (B200_PROFILING.md)")
        fp64_peak = g.measure_fp64_peak()
        kern = []
        for name, (kms, nl) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
            per_launch_s = kms * 1e-3 / max(nl, 1)
            by = algorithmic_bytes(name, n_avg, n_cells, S, case["n_class"], 64)
            kern.append({"kernel": name, "ms_per_step": kms / a.steps, "share": kms / ms, "launches": nl,
                         "alg_bytes_per_launch": by, "hbm_gbs": by / per_launch_s / 1e9 if per_launch_s > 0 else None})
        # FP64 work of the condensation integrator: rate evaluations x counted FP64 operations per evaluation
        evals_per_step = float((cst[:, 2].astype(np.float64) * case["counts"]).sum())
        # counted by ncu on this kernel (profiles/r01_s15_condense_ncu.md): per rate evaluation and particle,
        # integrator arithmetic included, 58.5 DFMA + 49.5 DMUL + 20.8 DADD thread instructions
        FLOP_PER_RATE_EVAL = 187.3   # DFMA = 2 flop
        FP64_INST_PER_RATE_EVAL = 128.8
        cond_ms = prof.get("k_condense", (0.0, 1))[0] / a.steps
        cond_tflops = evals_per_step * FLOP_PER_RATE_EVAL / (cond_ms * 1e-3) / 1e12 if cond_ms > 0 else None
        # share of the FP64 pipe's issue slots in use (a DMUL or DADD takes the slot of a DFMA but counts one flop)
        cond_pipe = (evals_per_step * FP64_INST_PER_RATE_EVAL / (cond_ms * 1e-3) / (fp64_peak * 1e12 / 2.0)
                     if cond_ms > 0 and fp64_peak else None)
        dom = kern[0] if kern else None
        if dom and dom["kernel"] == "k_condense":
            roofline = {"kernel": "k_condense", "bound": "fp64", "achieved": cond_tflops, "peak": fp64_peak,
                        "unit": "TFLOP/s", "frac": cond_tflops / fp64_peak if fp64_peak else None,
                        # dram__bytes_read + write of one ncu --set full capture of this kernel, scaled by the
                        # particle count: 1.323 GB + 64 MB for 7.64e6 particles (profiles/r01_s15_condense_ncu.md)
                        "traffic": 181.6 * n_avg, "traffic_unit": "bytes per launch (ncu capture, scaled)",
                        "peak_source": "FP64 FMA microbenchmark in this run (pmcgpu_measure_fp64_peak); "
                                       "MEASURED_PEAKS.json holds no FP64 figure",
                        "fp64_pipe_slots_used": cond_pipe,
                        "note": "the dominant kernel is bounded by the FP64 pipe, not by HBM or the tensor cores "
                                "(north star: no dense contraction on the path); achieved = rate evaluations x "
                                "187.3 flop (ncu-counted, DFMA = 2) / kernel time; fp64_pipe_slots_used counts "
                                "DFMA, DMUL and DADD alike (128.8 per evaluation) against peak / 2 instructions/s"}
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
                            # 1.536 GB + 1.497 GB for 7.64e6 records (profiles/r01_s15_sort_ncu.md)
                            "traffic": 397.0 * n_avg, "traffic_unit": "bytes per launch (ncu capture, scaled)",
                            "peak_source": hbm_src}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": t_dev * 1e3 / a.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(a, world), "particles_per_gpu": n_avg,
                           "phases": ["sort", "mc_coag(brownian)", "condense_particles", "rebalance", "partmc_to_wrf"] +
                                     (["trans_aero"] if has_transport else []),
                           "transport": bool(has_transport), "chem_dt_s": CHEM_DT,
                           "cell_population_end": {"max": int(cnt_end.max()), "over_1024": float((cnt_end > 1024).mean())},
                           "l2": "inputs (%.1f GB of particle state per GPU) larger than L2" % (
                               n_avg * (8 * S + 44) / 1e9),
                           "particle_count": "mean of the population at the start and the end of the timed region",
                           "coag_samples_per_step": int(csamp.sum()), "coag_events_per_step": int(ccoag.sum()),
                           "condense_steps_per_cell": float(cst[:, 0].mean()),
                           "condense_rate_evals_per_particle": float(cst[:, 2].mean())},
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": t_e2e * 1e3 / a.steps,
                        "what": "pmcgpu_partmc_timestep through the C ABI: T, p, 1/rho fields from pinned host memory, "
                                "partmc_to_wrf number concentrations back to host, status read; particles stay "
                                "resident (DESIGN.md 'ownership')"},
                "gpu_launches": int(launches), "roofline": roofline, "roofline_hbm": roofline_hbm, "kernels": kern,
                "fp64_peak_tflops": fp64_peak, "clocks": clk}
        if world == 1 and not a.no_cpu_baseline:
            v, sps, desc, ncores = sized_oracle_run(a, 2, 1, a.cpu_seconds, has_transport)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": ncores, "kind": "port",
                                    "sample": desc + ", 2 timed steps after 1 warm-up; CPU restatement of the "
                                                     "reference (oracle/), one patch per thread"}
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
