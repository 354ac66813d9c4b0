"""One ILU(0) apply + one BiCGSTAB solve on the Method 2 Jacobian (SquareGrid 2D nx, MEx4) for an ncu launch list.
usage: python scripts/m2_ncu.py [nx] [tri_mode]"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
nx = int(sys.argv[1]) if len(sys.argv) > 1 else 256
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 1
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
pcg.set_option("tri_mode", mode)
r = np.random.default_rng(1).standard_normal(len(rowptr) - 1)
pcg.precond(r)
pcg.precond(r)
if mode == 1:
    print(pcg.bicgstab(r * 1e-3, 1e-6, 10000)[1:])
