"""Dataflow triangular solve: device time of a preconditioner application per grid size (option tri_sf_ctas_per_sm_x2: CTAs = SMs x value / 2).
(profiles/r02_al_dataflow_grid_sweep.jsonl was taken with an environment override of the CTA count, 8 ... 296, that the library no longer has.)"""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
if os.environ.get("MMADMM_USE_PROF_LIB"): mm.LIB_PATH = os.path.join(ROOT, "build", "prof", "libmmadmm_b200_prof.so")
nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
pcg.set_option("tri_mode", 2); pcg.set_option("row_block", 2)
for x2 in (1, 2, 4):
    pcg.set_option("tri_sf_ctas_per_sm_x2", x2)
    print(json.dumps({"ctas_per_sm": x2 / 2, "pmode": os.environ.get("MMADMM_TRI_PROF_MODE"), "precond_ms_device": round(pcg.time_precond(5), 3)}))
