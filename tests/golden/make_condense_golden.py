"""Generates tests/golden/condense_env_golden.json from the reference's stored golden
table partmc/test/condense/ref_condense_0001_env.txt (columns: time, temp, rel_humid,
pressure, height; 61 rows, t = 0..600 s) and the scenario parameters of
partmc/test/condense/{run_part.spec, aero_init_dist.dat, aero_init_comp.dat, temp.dat,
pressure.dat}.  Run in the build container (the reference tree is not on the GPU box)."""
import json
import os

import numpy as np

REF = "/root/reference/partmc/test/condense"
g = np.loadtxt(os.path.join(REF, "ref_condense_0001_env.txt"))
out = dict(
    source="partmc/test/condense/ref_condense_0001_env.txt (reference golden, tolerance 1e-4 in test_condense_1.sh)",
    time=g[:, 0].tolist(), temp=g[:, 1].tolist(), rel_humid=g[:, 2].tolist(), pressure=g[:, 3].tolist(),
    scenario=dict(del_t=10.0, t_max=600.0, rel_humidity=0.95, temp=[[0, 1200], [290, 280]],
                  pressure=[[0, 1200], [1e5, 8e4]],
                  modes=[dict(num_conc=3.2e9, geom_mean_diam=2e-8, log10_geom_std_dev=0.161),
                         dict(num_conc=2.9e9, geom_mean_diam=1.16e-7, log10_geom_std_dev=0.217)],
                  mass_frac=dict(OC=1.375, SO4=1.0, NH4=0.375)))
# water mass concentration, column 23 of ref_condense_0001_aero_time.txt (tolerance 0.05, test_condense_2.sh)
a = np.loadtxt(os.path.join(REF, "ref_condense_0001_aero_time.txt"))
out["water_mass_conc"] = a[:, 22].tolist()
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "condense_env_golden.json"), "w") as f:
    json.dump(out, f)
print("rows", g.shape, "water col range", a[:, 22].min(), a[:, 22].max())
