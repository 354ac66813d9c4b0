"""Generic (non-block) dataflow triangular solve on the C2-size Method 2 Jacobian: device time per row_block / proxy setting."""
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
r = np.random.default_rng(3).standard_normal(len(rowptr) - 1)
z1 = pcg.precond(r)
pcg.set_option("tri_blocks", 0); pcg.set_option("tri_mode", 2)
for bs in (1, 2):
    pcg.set_option("row_block", bs)
    for proxy in (0, 1):
        pcg.set_option("tri_proxy", proxy)
        ok = bool(np.array_equal(pcg.precond(r), z1))
        print(json.dumps({"form": "generic", "row_block": bs, "proxy": proxy, "bit_identical_to_cluster_solve": ok, "precond_ms_device": round(pcg.time_precond(5), 3)}))
