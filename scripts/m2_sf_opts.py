"""Dataflow triangular solve (block form): device time of a preconditioner application for a few option settings."""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
pcg.set_option("tri_mode", 2); pcg.set_option("row_block", 2)
r = np.random.default_rng(3).standard_normal(len(rowptr) - 1)
pcg.set_option("tri_mode", 1); z1 = pcg.precond(r); pcg.set_option("tri_mode", 2)
for proxy in (1, 0):
    for x2 in (1, 2, 4):
        pcg.set_option("tri_proxy", proxy); pcg.set_option("tri_sf_ctas_per_sm_x2", x2)
        ok = bool(np.array_equal(pcg.precond(r), z1))
        print(json.dumps({"proxy": proxy, "ctas_per_sm": x2 / 2, "bit_identical_to_cluster_solve": ok, "precond_ms_device": round(pcg.time_precond(5), 3)}))
