"""Timing experiments on the cluster triangular solve (k_tri_cluster_lell) with the profiling build of the library
(-DMMADMM_TRI_PROF, build/prof/libmmadmm_b200_prof.so): MMADMM_TRI_PROF_MODE = 0 normal, 1 no `out` gathers, 2 no factor loads,
3 barrier only, 4 no fences.  Results are WRONG for modes != 0: only the device time of a preconditioner application is read.
usage: MMADMM_TRI_PROF_MODE=m python scripts/m2_prof_modes.py [nx] [row_block]"""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
mm.LIB_PATH = os.path.join(ROOT, "build", "prof", "libmmadmm_b200_prof.so")

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 2
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
pcg.set_option("tri_layout", 1)
pcg.set_option("row_block", bs)
i2 = pcg.info()
ms = pcg.time_precond(5)
lv = i2["lower_levels"] + i2["upper_levels"]
print(json.dumps({"mode": int(os.environ.get("MMADMM_TRI_PROF_MODE", "0")), "nx": nx, "row_block": bs, "levels": lv,
                  "precond_ms_device": round(ms, 3), "us_per_level": round(ms * 1e3 / lv, 3)}))
