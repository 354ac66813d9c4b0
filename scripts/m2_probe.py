"""Method 2 (backward Euler) on the device at size: wall time per MeshIntegrator::backwardsEulerStep (host-driven Newton loop,
so wall clock with the stream synchronised is the meaningful figure), Newton / BiCGSTAB iteration counts, kernel launches.
usage: python scripts/m2_probe.py [nx] [steps] [D] [tri_mode]"""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
D = int(sys.argv[3]) if len(sys.argv) > 3 else 2
tri_mode = int(sys.argv[4]) if len(sys.argv) > 4 else 1
Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
mon = mm.TestMonitor(D, 4 if D == 2 else 3).evaluate_at_vertices(Vp)   # MEx4 shock layer (C2) / MEx33D (C4)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
solver.set_option("m2_tri_mode", tri_mode)
out = {"tri_mode": tri_mode, "workload": f"SquareGrid {D}D nx={nx}", "elements": int(F.shape[0]), "unknowns": int(Vp.size), "steps": []}
for s in range(steps):
    l0 = solver.stats()["kernel_launches"]
    t = time.perf_counter()
    ih = solver.backwardsEulerStep(0.025, 1e-3)
    dt = time.perf_counter() - t
    out["steps"].append({"ms": round(dt * 1e3, 3), "ctx_launches": solver.stats()["kernel_launches"] - l0, **solver.last})
rowptr, col, val = solver.backward_euler_jacobian()
out["jacobian_nnz"] = int(len(val))
print(json.dumps(out))
