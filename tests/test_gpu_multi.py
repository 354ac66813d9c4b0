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


def _worker(rank, world, id128, case, nsteps, q, neuler=0, delay=None):
    import importlib as il
    import time
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
    for k in range(nsteps):
        if delay and delay[0] == rank and delay[1] == k:
            time.sleep(delay[2])      # this rank's host stalls between two steps (file output, GC, a debugger ...)
        ihs.append(s.step(10, 1e-3))
        its.append(s.last["iters"])
    e1 = s.getEnergy()
    s.done()
    q.put((rank, e0, ihs, its, e1, mesh.Vp.copy(), s.local_sizes()))
    s.close()


def run_dist(world, case, nsteps, neuler=0, delay=None):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    id128 = mm.comm_unique_id()
    procs = [ctx.Process(target=_worker, args=(r, world, id128, case, nsteps, q, neuler, delay)) for r in range(world)]
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
@pytest.mark.parametrize("world", [2, 4, 8])
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


def test_a_stalled_rank_does_not_kill_its_peers():
    """Ranks enter mmadmm_step at independent host times.  While one rank's host is busy for 15 s between two steps, its peers
    sit in the flag waits of k_xfinish_p2p / k_scal_sync; the wait is bounded by a TIME-OUT (option "peer_timeout_s", default
    300 s), not by a poll count, and a peer that really never arrives yields MMADMM_EPEER on every rank instead of a trap that
    destroys the CUDA context."""
    if ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    case = (2, 48, 3, 100.0, 1, False)
    res = run_dist(2, case, 4, delay=(1, 2, 15.0))
    ref = run_dist(2, case, 4)
    assert res[0][2] == res[1][2] == ref[0][2]
    assert np.array_equal(res[0][5], ref[0][5])


def test_reference_driver_on_two_gpus_reproduces_committed_trace(tmp_path, monkeypatch):
    """`MMADMM_GPUS=2 mesh.exe Monitor2160 0 1` -- the reference's main.cpp, unchanged, forks one worker per extra
    GPU inside Mesh<D>'s constructor (mm-admm_b200/host/dropin/src/Mesh.h) and must write the committed Ih0.txt (162 lines, the
    DtTol stop included) and the same final mesh as the single-GPU run."""
    import json
    from test_gpu_dropin import EXE, TRACES, run_driver
    if ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    if not os.path.exists(EXE):
        pytest.skip("build/dropin/mesh.exe not built")
    name = "Monitor2160"
    spec0 = TRACES[name]["input"]
    out1 = run_driver(tmp_path / "one", name, spec0["nSteps"], extra={"DtTol": spec0["DtTol"]})
    monkeypatch.setenv("MMADMM_GPUS", "2")
    out2 = run_driver(tmp_path / "two", name, spec0["nSteps"], extra={"DtTol": spec0["DtTol"]})
    ref = TRACES[name]["Ih"]
    assert len(out2[2]) == len(ref) == 162
    for i, (g, e) in enumerate(zip(out2[2], ref)):
        assert g == ("%g" % e), f"step {i}: 2-GPU drop-in driver wrote {g}, reference trace has {e:g}"
    p1 = open(out1[1] / "points.txt").read()
    p2 = open(out2[1] / "points.txt").read()
    a = np.array([[float(w) for w in ln.split(",")] for ln in p1.split("\n") if ln.strip()])
    b = np.array([[float(w) for w in ln.split(",")] for ln in p2.split("\n") if ln.strip()])
    assert a.shape == b.shape and np.abs(a - b).max() <= 2e-6 * np.abs(a).max()
    assert open(out1[1] / "triangles.txt").read() == open(out2[1] / "triangles.txt").read()
