"""Small hot-path run for compute-sanitizer (memcheck / racecheck / synccheck): 2D 20x20 square and 3D 5^3 cube, 4 ADMM outer steps
each (finite-difference Hessians, gradient predictor, persistent prox kernels with their mbarrier / cp.async / bulk-copy pipelines,
fused x-updates), an explicit-Euler step, checkpoint / restore and a deferred write-back.  With --ranks N (N >= 2) the same on N GPUs,
one process per GPU (NVLink mailbox exchange, fused push / finish kernels).  Usage:
    compute-sanitizer --tool racecheck python scripts/sanitize_small.py
    compute-sanitizer --tool memcheck --target-processes all python scripts/sanitize_small.py --ranks 2"""
import argparse
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(rank, world, id128, q):
    mm = importlib.import_module("mm-admm_b200")
    out = []
    for D, nx, kind in ((2, 20, 4), (3, 5, 3)):
        Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
        mon = mm.TestMonitor(D, kind).evaluate_at_vertices(Vp)
        mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5)
        s = mm.MeshIntegrator(0.025, mesh, device=rank, sync_vp="deferred", comm=(rank, world, id128) if world > 1 else None)
        ih = [s.getEnergy()]
        for k in range(4):
            ih.append(s.step(10, 1e-3))
        s.checkpoint()
        ih.append(s.step(10, 1e-3))
        s.restore()
        ih.append(s.step(10, 1e-3))
        assert ih[-1] == ih[-2]
        ih.append(s.eulerStep())
        s.done()
        out.append((ih, float(np.abs(mesh.Vp).sum())))
        s.close()
    if q is not None:
        q.put((rank, out))
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=1)
    a = ap.parse_args()
    if a.ranks == 1:
        print(run(0, 1, None, None))
    else:
        import torch.multiprocessing as mp
        mm = importlib.import_module("mm-admm_b200")
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        id128 = mm.comm_unique_id()
        ps = [ctx.Process(target=run, args=(r, a.ranks, id128, q)) for r in range(a.ranks)]
        for p in ps:
            p.start()
        res = sorted(q.get(timeout=3000) for _ in ps)
        for p in ps:
            p.join()
        assert all(r[1] == res[0][1] for r in res), "ranks disagree"
        print(res[0])
