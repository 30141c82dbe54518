"""Copies the two reference tables of the reference's fractal coagulation tests into a fixture that travels to the
GPU box (/root/reference does not):
  partmc/test/fractal/ref_free_df_2_4_dimless_time_regrid.txt  (test_fractal_2.sh, free molecular regime, df = 2.4)
  partmc/test/fractal/ref_cont_df_1_8_dimless_time_regrid.txt  (test_fractal_3.sh, continuum regime, df = 1.8)
Each row: dimensionless time, dimensionless number concentration (test_fractal_dimless_time.F90:126-147); the
reference compares its own run with them by `numeric_diff --by col --rel-tol 0.1`.
Run here: python tests/golden/make_fractal_golden.py"""
import json
import os

import numpy as np

REF = "/root/reference/partmc/test/fractal/"
out = {}
for key, name in (("free_df_2_4", "ref_free_df_2_4_dimless_time_regrid.txt"),
                  ("cont_df_1_8", "ref_cont_df_1_8_dimless_time_regrid.txt")):
    out[key] = np.loadtxt(REF + name).tolist()
out["source"] = "partmc/test/fractal/ref_{free_df_2_4,cont_df_1_8}_dimless_time_regrid.txt"
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "fractal_dimless_time.json"), "w") as f:
    json.dump(out, f, indent=0)
print({k: len(v) for k, v in out.items() if k != "source"})
