"""Method 2 preconditioner application in 3D (SquareGrid nx^3): device time per tri_mode / row_block.  usage: m2_parts3d.py [nx]"""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
nx = int(sys.argv[1]) if len(sys.argv) > 1 else 32
Vp, F, mask = mm.generate_uniform_rect_mesh(3, nx, nx, nx)
mon = mm.TestMonitor(3, 3).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
out = {"nx": nx, "tets": int(F.shape[0]), "rows": len(rowptr) - 1, "nnz": len(val)}
r = np.random.default_rng(5).standard_normal(len(rowptr) - 1)
ref = None
for mode, bs, lay in ((1, 1, 0), (1, 3, 0), (1, 3, 1), (2, 1, 0), (2, 3, 0)):
    pcg.set_option("tri_mode", mode); pcg.set_option("tri_layout", lay); pcg.set_option("row_block", bs)
    z = pcg.precond(r)
    if ref is None: ref = z
    i2 = pcg.info()
    out[f"tri_mode_{mode}_row_block_{bs}_layout_{lay}"] = {"levels": [i2["lower_levels"], i2["upper_levels"]], "bit_identical": bool(np.array_equal(z, ref)),
                                                          "precond_ms_device": round(pcg.time_precond(5), 3)}
import time
t0 = time.perf_counter(); solver.backwardsEulerStep(0.025, 1e-3); t1 = time.perf_counter(); solver.backwardsEulerStep(0.025, 1e-3); t2 = time.perf_counter()
out["backward_euler_step_ms"] = [round((t1 - t0) * 1e3, 1), round((t2 - t1) * 1e3, 1)]
out["last"] = solver.last
print(json.dumps(out))
