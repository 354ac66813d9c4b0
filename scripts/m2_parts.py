"""Where a Method 2 step spends its time: the ILU(0) triangular-solve pair in both tri_modes, one BiCGSTAB solve, the SpMV,
the residual evaluation.  Wall clock around synchronous C-ABI calls.  usage: python scripts/m2_parts.py [nx]"""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 256
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
t = time.perf_counter(); rowptr, col, val = solver.backward_euler_jacobian(); t_jac = time.perf_counter() - t
t = time.perf_counter(); pcg = mm.PCG(rowptr, col, val, precond="ilu0"); t_create = time.perf_counter() - t
info = pcg.info()
rng = np.random.default_rng(1)
r = rng.standard_normal(len(rowptr) - 1)


def best(f, reps=5):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); f(); ts.append(time.perf_counter() - t0)
    return round(min(ts) * 1e3, 3)


out = {"nx": nx, "rows": len(r), "nnz": len(val), "levels": [info["lower_levels"], info["upper_levels"]],
       "jacobian_build_incl_ilu_ms": round(t_jac * 1e3, 1), "pcg_create_ms": round(t_create * 1e3, 1),
       "spmv_ms": round(pcg.time_spmv(20), 4), "energy_call_ms": best(solver.getEnergy)}
for mode in (0, 1):
    pcg.set_option("tri_mode", mode)
    out[f"precond_ms_tri{mode}"] = best(lambda: pcg.precond(r))
    out[f"bicgstab_ms_tri{mode}"] = best(lambda: pcg.bicgstab(r * 1e-3, 1e-6, 10000), 3)
    out[f"bicgstab_iters_tri{mode}"] = pcg.bicgstab(r * 1e-3, 1e-6, 10000)[1]
pcg.set_option("tri_mode", 1)
for lay in (0, 1):            # 1: level-ELL copy of the factor for the cluster kernels
    pcg.set_option("tri_layout", lay)
    for bs in (1, 2):         # block rows: groups of bs consecutive rows per thread (device-timed preconditioner application)
        pcg.set_option("row_block", bs)
        i2 = pcg.info()
        out[f"layout_{lay}_row_block_{bs}"] = {"levels": [i2["lower_levels"], i2["upper_levels"]], "precond_ms_device": round(pcg.time_precond(5), 3)}
pcg.set_option("tri_mode", 2)     # dataflow solve: no barrier between levels
for bs in (1, 2):
    pcg.set_option("row_block", bs)
    i2 = pcg.info()
    out[f"dataflow_row_block_{bs}"] = {"levels": [i2["lower_levels"], i2["upper_levels"]], "precond_ms_device": round(pcg.time_precond(5), 3)}
print(json.dumps(out))
