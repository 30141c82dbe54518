"""Timing of the single-box configurations K2 (Brownian mc_coag, 1e5 particles in one cell) and K3
(condense_particles, 1e5 particles in one cell, test/condense scenario) on one GPU, next to the CPU oracle.
Run on a GPU box from the repository root: python profiles/k2_k3_single_box_timing.py"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402
import wrf_partmc_b200 as W  # noqa: E402
from oracle import oracle as O  # noqa: E402

out = {}
# ---- K2
case = common.make_cells(1, 100000, n_spec=1, seed=5, ragged=False, num_conc=1e11, r_g=2.5e-8)
g = common.gpu_from_case(W, case, cell_capacity=120000)
o = common.oracle_from_case(O, case)
g.mc_coag(W.KERNEL_BROWN, 60.0, step=0)
g.synchronize()
t0 = time.perf_counter()
n_ev = 0
for step in range(1, 6):
    ns, nc = g.mc_coag(W.KERNEL_BROWN, 60.0, step=step)
    n_ev += int(nc[0])
g.synchronize()
t_dev = (time.perf_counter() - t0) / 5
o.sort(0, force_resort=True, physical=True, edge_recheck=True)
o.mc_coag(0, O.KERNEL_BROWN, 60.0, step=0)
t0 = time.perf_counter()
for step in range(1, 6):
    o.sort(0, force_resort=True, physical=True, edge_recheck=True)
    o.mc_coag(0, O.KERNEL_BROWN, 60.0, step=step)
t_cpu = (time.perf_counter() - t0) / 5
out["K2"] = {"particles": 100000, "dt_s": 60.0, "events_per_step": n_ev / 5, "gpu_ms_per_step": t_dev * 1e3,
             "cpu_oracle_ms_per_step_1core": t_cpu * 1e3}
# ---- K2 as a 16-member ensemble: 16 boxes side by side (the cell number is part of every draw's address)
case16 = common.make_cells(16, 100000, n_spec=1, seed=5, ragged=False, num_conc=1e11, r_g=2.5e-8)
g16 = common.gpu_from_case(W, case16, cell_capacity=120000)
g16.mc_coag(W.KERNEL_BROWN, 60.0, step=0)
g16.synchronize()
t0 = time.perf_counter()
for step in range(1, 6):
    g16.mc_coag(W.KERNEL_BROWN, 60.0, step=step)
g16.synchronize()
out["K2_ensemble16"] = {"particles": 1600000, "gpu_ms_per_step": (time.perf_counter() - t0) / 5 * 1e3,
                        "cpu_oracle_ms_per_step_1core": 16 * t_cpu * 1e3}
del g16
# ---- K3
case = common.condense_case(50000, n_cells=1)
g = common.gpu_from_case(W, case, cell_capacity=100064)
o = common.oracle_from_case(O, case)
g.set_env(rel_humid=np.array([0.95]))
g.condense_equilib_particles()
o.set_env(0, T=290.0, p=1e5, RH=0.95)
o.condense_equilib(0)
ts, tc, steps = [], [], []
for k in range(6):
    t1 = (k + 1) * 10.0
    T1, p1 = 290.0 - 10.0 * t1 / 1200.0, 1e5 - 2e4 * t1 / 1200.0
    g.synchronize()
    t0 = time.perf_counter()
    rh, st, wre = g.condense_particles([T1], [p1], 10.0)
    g.synchronize()
    ts.append(time.perf_counter() - t0)
    steps.append(int(st[0, 0]))
    if k < 2:
        t0 = time.perf_counter()
        o.condense_particles(0, T1, p1, 10.0, step=k, method=4)
        tc.append(time.perf_counter() - t0)
out["K3"] = {"particles": 100000, "dt_s": 10.0, "integrator_steps_per_call": float(np.mean(steps[1:])),
             "gpu_ms_per_step": float(np.mean(ts[1:]) * 1e3), "cpu_oracle_ms_per_step_1core": float(np.mean(tc) * 1e3)}
print(json.dumps(out))
