"""Debug aid: 3D FREE-boundary / comp-mesh / GradUse cases against the oracle under several kernel options."""
import importlib, sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
mm = importlib.import_module("mm-admm_b200")
import oracle_lib as O
for btype, kind in ((0, 4), (1, 3)):
    Vp, F, mask = O.uniform_rect_mesh(3, 5, 5, 5, btype=btype)
    mon = O.monitor_eval(3, kind, Vp)
    rng = np.random.default_rng(5)
    Vc = Vp + 0.003 * rng.standard_normal(Vp.shape) * (mask == 2)[:, None]
    for kwname, kw in (("plain", {}), ("comp", dict(Vc=Vc)), ("graduse", dict(grad_use=True))):
        orc = O.Oracle(Vp, F, mask, mon, 50.0, 0.5, 0.025, nthreads=4, **{("Vc" if k == "Vc" else k): v for k, v in kw.items()})
        xs = []
        for s in range(4):
            orc.step(10, 1e-3); xs.append(orc.x().copy())
        for opts in ({}, {"prox_variant": 1}, {"iso_grid": 0}, {"usum_period": 0}, {"prox_variant": 1, "usum_period": 0, "iso_grid": 0}):
            mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, grad_use=kw.get("grad_use", False), Vc=None if "Vc" not in kw else Vc.copy())
            sol = mm.MeshIntegrator(0.025, mesh)
            for k, v in opts.items():
                sol.set_option(k, v)
            errs = []
            for s in range(4):
                sol.step(10, 1e-3)
                errs.append(float(np.abs(mesh.Vp - xs[s]).max() / np.abs(xs[s]).max()))
            hist = sol.stats()["bfgs_hist"]
            print(btype, kind, kwname, opts, ["%.1e" % e for e in errs], "bfgs hist", {int(i): int(h) for i, h in enumerate(hist) if h})
            sol.close()
