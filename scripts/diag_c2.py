import importlib, sys, time, numpy as np
sys.path.insert(0, '.')
mm = importlib.import_module("mm-admm_b200")
nx = int(sys.argv[1]); nsteps = int(sys.argv[2]); rho = float(sys.argv[3]) if len(sys.argv) > 3 else 50.0
dt = float(sys.argv[4]) if len(sys.argv) > 4 else 0.025
mon_kind = int(sys.argv[5]) if len(sys.argv) > 5 else 4
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx)
mon = mm.TestMonitor(2, mon_kind).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp, F, mask, mon, rho=rho, tau=0.5)
s = mm.MeshIntegrator(dt, mesh, sync_vp=False)
print("E", mesh.E, "grid", s.grid_dims(), "E0", s.getEnergy())
for i in range(nsteps):
    try:
        t = time.time(); ih = s.step(10, 1e-3); print(i, "%.12g" % ih, s.last, "%.1f ms" % ((time.time() - t) * 1e3), flush=True)
    except Exception as e:
        print(i, "FAILED", e); break
print(s.stats()["bfgs_hist"][:12])
