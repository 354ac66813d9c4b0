"""Debug: dataflow triangular solve vs the per-level kernels on the 3D Method 2 Jacobian; prints the rows that differ."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
D, nx = 3, 6
Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx)
mon = mm.TestMonitor(D, 3).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
n = len(rowptr) - 1
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
r = np.random.default_rng(12).standard_normal(n)
pcg.set_option("tri_mode", 0); pcg.set_option("row_block", 1)
z0 = pcg.precond(r)
lens = np.diff(rowptr)
print("n", n, "max row len", lens.max(), "grid env", os.environ.get("MMADMM_SF_GRID"))
for bs in (1, 3, 4, 6):
    pcg.set_option("tri_mode", 2); pcg.set_option("row_block", bs)
    for rep in range(2):
        z = pcg.precond(r)
        bad = np.nonzero(z != z0)[0]
        print("bs", bs, "rep", rep, "levels", pcg.info()["lower_levels"], pcg.info()["upper_levels"], "bad rows", bad[:20], "n bad", len(bad),
              "max rel", (np.abs(z - z0).max() / np.abs(z0).max()) if len(bad) else 0.0)
