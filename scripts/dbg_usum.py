import importlib, numpy as np
mm = importlib.import_module("mm-admm_b200")
for (D, nx, kind, nit, tol) in [(2, 24, 4, 10, 1e-5), (2, 24, 4, 10, 1e-3), (2, 24, 4, 1, 1e-3), (3, 6, 4, 10, 1e-5)]:
    Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
    mon = mm.TestMonitor(D, kind).evaluate_at_vertices(Vp)
    ref = None
    for period in (0, 1, 3, 64):
        mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5)
        s = mm.MeshIntegrator(0.025, mesh)
        s.set_option("usum_period", period)
        tr = []
        try:
            for k in range(8):
                I = s.step(nit, tol)
                tr.append((I, mesh.Vp.copy()))
        except Exception as e:
            print(D, nx, kind, nit, tol, "period", period, "failed at step", len(tr), str(e)[:60])
        if period == 0:
            ref = tr
        else:
            d = [float(np.abs(a[1] - b[1]).max()) for a, b in zip(ref, tr)]
            print(D, nx, kind, nit, tol, "period", period, "steps", len(tr), "max dx per step", ["%.1e" % v for v in d], "iters", s.stats()["admm_iters"])
        s.close()
