"""Multi-GPU parity (needs >= 2 B200s on one box): the element-block-partitioned solve, one process per GPU over NCCL,
against the single-GPU solve (1e-12: same arithmetic, only the association of the shared-vertex sums differs) and against
the CPU oracle (1e-9, the north-star tolerance)."""
import importlib
import os

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu
mm = importlib.import_module("mm-admm_b200")


def _case_mesh(m, case):
    """case = (D, nx, monitor, rho, boundary type, compMesh[, "disc"]) -> Vp, F, mask, tau, dt.  A disc case is BASELINE
    config 3 at test size: nx hexagonal rings, Monitor310's tau/dt; its element blocks are annuli, so every rank has two
    neighbours whose interfaces are whole rings of different lengths."""
    D, nx, kind, rho, btype, comp = case[:6]
    if len(case) > 6 and case[6] == "disc":
        return m.generate_disc_mesh(nx, 1.0, (0.0, 0.0), btype) + (0.1, 0.05)
    return m.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0, boundary_type=btype) + (0.5, 0.025)


def _worker(rank, world, id128, case, nsteps, q, neuler=0):
    import importlib as il
    m = il.import_module("mm-admm_b200")
    D, nx, kind, rho, btype, comp = case[:6]
    Vp, F, mask, tau, dt = _case_mesh(m, case)
    mon = m.TestMonitor(D, kind).evaluate_at_vertices(Vp)
    mesh = m.Mesh(Vp.copy(), F.copy(), mask, mon, rho=rho, tau=tau, Vc=Vp.copy() if comp else None)
    s = m.MeshIntegrator(dt, mesh, device=rank, comm=(rank, world, id128))
    e0 = s.getEnergy()
    ihs, its = [], []
    for _ in range(neuler):   # Method 1 steps first (collective too: assembled gradient of the shared vertices)
        ihs.append(s.eulerStep())
        its.append(0)
    for _ in range(nsteps):
        ihs.append(s.step(10, 1e-3))
        its.append(s.last["iters"])
    e1 = s.getEnergy()
    s.done()
    q.put((rank, e0, ihs, its, e1, mesh.Vp.copy(), s.local_sizes()))
    s.close()


def run_dist(world, case, nsteps, neuler=0):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    id128 = mm.comm_unique_id()
    procs = [ctx.Process(target=_worker, args=(r, world, id128, case, nsteps, q, neuler)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=600) for _ in procs], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return res


def ngpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("case,nsteps", [((2, 48, 3, 100.0, 1, False), 6), ((2, 40, 1, 50.0, 0, True), 5), ((3, 8, 3, 50.0, 1, False), 4),
                                         ((2, 40, 5, 5.0, 0, False, "disc"), 5)])
@pytest.mark.parametrize("world", [2, 4])
def test_partitioned_solve_matches_single_gpu_and_oracle(case, nsteps, world):
    if ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    D, nx, kind, rho, btype, comp = case[:6]
    res = run_dist(world, case, nsteps)
    # every rank returns the same scalars and the same assembled mesh
    for r in res[1:]:
        assert r[1] == res[0][1] and r[2] == res[0][2] and r[3] == res[0][3] and r[4] == res[0][4]
        assert np.array_equal(r[5], res[0][5])
    # single GPU
    Vp, F, mask, tau, dt = _case_mesh(mm, case)
    assert sum(r[6][1] for r in res) == F.shape[0]
    mon = mm.TestMonitor(D, kind).evaluate_at_vertices(Vp)
    Vc = Vp.copy() if comp else None
    mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=rho, tau=tau, Vc=Vc)
    s = mm.MeshIntegrator(dt, mesh)
    orc = O.Oracle(Vp, F, mask, mon, rho, tau, dt, Vc=Vc, nthreads=4)
    assert abs(res[0][1] - s.getEnergy()) <= 1e-12 * abs(res[0][1])
    for k in range(nsteps):
        ih = s.step(10, 1e-3)
        ro = orc.step(10, 1e-3)
        assert s.last["iters"] == res[0][3][k] == ro["iters"]
        assert abs(res[0][2][k] - ih) <= 1e-12 * abs(ih)
        assert abs(res[0][2][k] - ro["Ih"]) <= 1e-9 * abs(ro["Ih"])
    s.done()
    xs, xd, xo = mesh.Vp, res[0][5], orc.x()
    assert np.abs(xd - xs).max() <= 1e-12 * np.abs(xs).max()
    assert np.abs(xd - xo).max() <= 1e-9 * np.abs(xo).max()
    assert abs(res[0][4] - orc.energy()) <= 1e-9 * abs(orc.energy())
    s.close()


@pytest.mark.parametrize("case", [(2, 48, 3, 100.0, 1, False), (3, 8, 3, 50.0, 1, False)])
def test_partitioned_explicit_euler_matches_single_gpu_and_oracle(case):
    """Method 1 on a partitioned mesh: 5 eulerStep calls, then 3 ADMM steps from that state."""
    if ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    D, nx, kind, rho, btype, comp = case[:6]
    res = run_dist(2, case, 3, neuler=5)
    assert res[0][2] == res[1][2] and np.array_equal(res[0][5], res[1][5])
    Vp, F, mask, tau, dt = _case_mesh(mm, case)
    mon = mm.TestMonitor(D, kind).evaluate_at_vertices(Vp)
    mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=rho, tau=tau)
    s = mm.MeshIntegrator(dt, mesh)
    orc = O.Oracle(Vp, F, mask, mon, rho, tau, dt, nthreads=4)
    for k in range(8):
        ih = s.eulerStep() if k < 5 else s.step(10, 1e-3)
        ro = orc.euler_step() if k < 5 else orc.step(10, 1e-3)
        assert ro["status"] == 0
        assert abs(res[0][2][k] - ih) <= 1e-12 * abs(ih)
        assert abs(res[0][2][k] - ro["Ih"]) <= 1e-9 * abs(ro["Ih"])
    s.done()
    assert np.abs(res[0][5] - mesh.Vp).max() <= 1e-12 * np.abs(mesh.Vp).max()
    assert np.abs(res[0][5] - orc.x()).max() <= 1e-9 * np.abs(orc.x()).max()
    s.close()
