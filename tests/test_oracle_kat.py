"""Known-answer tests that PIN the oracle to the reference.

The only results the reference commits for the ADMM path are the per-step energy traces
Experiments/Results/<Name>/Ih0.txt (main.cpp:131-140,245-248; 6 significant digits).  The oracle must
reproduce them digit for digit from the shipped Experiments/InputFiles/<Name>.json parameters.
"""
import json
import os

import numpy as np
import pytest

import oracle_lib as O

TRACES = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "ih_traces.json")))


def fmt6(v):
    return float("%.6g" % v)


def build_from_input(name, golden_dir, nthreads=1):
    spec = TRACES[name]["input"]
    D = spec["Dim"]
    if spec["TestType"] == "SquareGrid":
        Vp, F, mask = O.uniform_rect_mesh(D, spec["nx"], spec["ny"], spec.get("nz", 0),
                                          (spec["xa"], spec["xb"], spec["ya"], spec["yb"], spec.get("za", 0), spec.get("zb", 0)),
                                          1 if spec["BoundaryType"] != 0 else 0)
    else:
        stem = os.path.basename(spec["PntsFile"]).replace("points.txt", "")
        Vp, F, mask = O.read_triangles(D, f"{golden_dir}/{stem}triangles.txt", f"{golden_dir}/{stem}points.txt",
                                       f"{golden_dir}/{stem}mask.txt")
    mon = O.monitor_eval(D, spec["MonType"], Vp)
    orc = O.Oracle(Vp, F, mask, mon, spec["rho"], spec["tau"], spec["dt"], Vc=Vp.copy() if spec["CompMesh"] else None,
                   grad_use=spec["GradUse"], nthreads=nthreads)
    return orc, spec


@pytest.mark.parametrize("name,nsteps,nthreads", [("Monitor210", 19, 1), ("Monitor220", 12, 1), ("Monitor310", 30, 1),
                                                  ("3DMonitor210", 6, 4)])
def test_energy_trace_matches_reference(name, nsteps, nthreads, golden_dir):
    orc, spec = build_from_input(name, golden_dir, nthreads)
    ref = TRACES[name]["Ih"]
    got = [orc.energy()]
    for _ in range(nsteps):
        r = orc.step(spec["AdmmIter"], 1e-3)  # tol hard-coded in main.cpp:184
        assert r["status"] == 0
        got.append(r["Ih"])
    for i, (g, e) in enumerate(zip(got, ref)):
        assert fmt6(g) == e, f"{name} step {i}: oracle {g!r} vs reference {e}"


@pytest.mark.parametrize("name,max_steps", [("Monitor2160", None), ("3DMonitor220", None), ("Monitor2320", 40)])
def test_large_committed_traces_match_reference_including_run_length(name, max_steps, golden_dir):
    """The reference's large committed runs (SquareGrid 160^2 = 102,400 triangles, 162 lines; 3D 20^3 = 96,000 tets, 101 lines;
    320^2 = 409,600 triangles, 327 lines -- the first 40 here, ALL of them on the GPU through mesh.exe in
    tests/test_gpu_parity_large.py and, for the oracle, by tests/golden/check_full_traces.py): the driver loop of
    main.cpp:180-209 with the shipped DtTol, so the step at which the run stops is pinned too."""
    orc, spec = build_from_input(name, golden_dir, nthreads=os.cpu_count() or 1)
    ref = TRACES[name]["Ih"]
    got, prev = [orc.energy()], float("inf")
    for i in range(spec["nSteps"] if max_steps is None else max_steps):
        r = orc.step(spec["AdmmIter"], 1e-3)
        assert r["status"] == 0
        got.append(r["Ih"])
        if i != 0 and abs((r["Ih"] - prev) / spec["dt"]) < spec["DtTol"]:     # main.cpp:203-206
            break
        prev = r["Ih"]
    if max_steps is None:
        assert len(got) == len(ref) == TRACES[name]["lines_committed"]
    for i, (g, e) in enumerate(zip(got, ref)):
        assert fmt6(g) == e, f"{name} step {i}: oracle {g!r} vs reference {e}"


def test_oracle_clone_is_a_deep_copy(golden_dir):
    a, spec = build_from_input("Monitor210", golden_dir, 1)
    a.step(spec["AdmmIter"])
    b = a.clone()
    b.set_threads(3)
    ra = [a.step(spec["AdmmIter"]) for _ in range(3)]
    rb = [b.step(spec["AdmmIter"]) for _ in range(3)]
    assert ra == rb and np.array_equal(a.x(), b.x()) and np.array_equal(a.hess(), b.hess())


def test_oracle_threads_do_not_change_results(golden_dir):
    a, spec = build_from_input("Monitor210", golden_dir, 1)
    b, _ = build_from_input("Monitor210", golden_dir, 4)
    for _ in range(4):
        ra = a.step(spec["AdmmIter"]); rb = b.step(spec["AdmmIter"])
        assert ra == rb
    assert np.array_equal(a.x(), b.x())


def test_lu_inverse_matches_numpy():
    rng = np.random.default_rng(0)
    for n in (6, 12):
        A = rng.standard_normal((n, n)) + n * np.eye(n)
        assert np.allclose(O.lu_inverse(A), np.linalg.inv(A), rtol=1e-12, atol=1e-13)


def test_monitor_quirks():
    # MEx0 identity, MEx3 2D == MEx33D radial form, MEx53D ignores d/dy (MEx53D.h:20-22)
    X = np.array([[0.3, 0.4], [0.5, 0.5]])
    assert np.array_equal(O.monitor_eval(2, 0, X), np.tile([1.0, 0, 0, 1.0], (2, 1)))
    m3 = O.monitor_eval(2, 3, X)
    r = np.sqrt((X[:, 0] - .5) ** 2 + (X[:, 1] - .5) ** 2)
    assert np.allclose(m3[:, 0], np.sqrt(0.01 / (2 + np.cos(8 * np.pi * r))), rtol=1e-14)
    assert np.all(m3[:, 1] == 0) and np.array_equal(m3[:, 0], m3[:, 3])


def test_explicit_euler_trace_matches_reference(golden_dir):
    """Method 1 (MeshIntegrator::eulerStep, src/MeshIntegrator.cpp:87-95): Experiments/Results/Monitor220/Ih1.txt, all 49 steps."""
    orc, spec = build_from_input("Monitor220", golden_dir)
    ref = TRACES["Monitor220"]["Ih1"]
    got = [orc.energy()]
    for _ in range(len(ref) - 1):
        r = orc.euler_step()
        assert r["status"] == 0
        got.append(r["Ih"])
    assert len(ref) == 50
    for i, (g, e) in enumerate(zip(got, ref)):
        assert fmt6(g) == e, f"euler step {i}: oracle {g!r} vs reference {e}"


def _method2_solver(orc, restated=False):
    import pcg_oracle as PO
    rows, cols = orc.jacobian_entries()
    cls = PO.RestatedJac if (restated or not PO.have_lasolver()) else PO.LASolverJac
    return cls(orc.P * orc.D, rows, cols)


@pytest.mark.parametrize("name,restated", [("Monitor210", False), ("Monitor220", False), ("Monitor310", False),
                                           ("3DMonitor210", False), ("Monitor210", True)])
def test_backward_euler_trace_matches_reference(name, restated, golden_dir):
    """Method 2 (Mesh::backwardsEulerStep, src/Mesh.cpp:1263-1341: Newton on x - xn + (dt/tau) grad I = 0 with the
    finite-difference Jacobian of FSubJac and LASolver's ILU(0) + CGSTAB): Experiments/Results/<name>/Ih2.txt, ALL
    committed steps (18 / 52 / 355 / 86), linear solves by the reference's own LASolver (oracle/_ref) or, `restated`,
    by the numpy ILU(0) + CGSTAB restatement."""
    if name == "3DMonitor210":
        import pcg_oracle as PO
        if not PO.have_lasolver():
            pytest.skip("3D trace needs oracle/_ref (the numpy ILU(0) is too slow at this size)")
    orc, spec = build_from_input(name, golden_dir)
    ref = TRACES[name]["Ih2"]
    S = _method2_solver(orc, restated)
    got = [orc.energy()]
    for _ in range(len(ref) - 1):
        r = orc.backwards_euler_step(S, 1e-3)  # tol hard-coded in main.cpp:190
        assert r["status"] == 0 and r["newton_iters"] < 1000
        got.append(r["Ih"])
    for i, (g, e) in enumerate(zip(got, ref)):
        assert fmt6(g) == e, f"{name} backward-Euler step {i}: oracle {g!r} vs reference {e}"


def test_restated_cgstab_equals_lasolver(golden_dir):
    """The numpy ILU(0) + CGSTAB restatement against the reference's LASolver on the Method 2 Jacobian of Monitor220:
    same CSR pattern from the same set_entry sequence, same iteration counts, iterates equal to rounding."""
    import pcg_oracle as PO
    if not PO.have_lasolver():
        pytest.skip("oracle/_ref not built")
    a, _ = build_from_input("Monitor220", golden_dir)
    b, _ = build_from_input("Monitor220", golden_dir)
    S, R = _method2_solver(a), _method2_solver(b, restated=True)
    assert np.array_equal(S.ia, R.ia) and np.array_equal(S.ja, R.ja)
    for _ in range(3):
        ra, rb = a.backwards_euler_step(S), b.backwards_euler_step(R)
        assert ra["inner_iters"] == rb["inner_iters"] and ra["newton_iters"] == rb["newton_iters"]
        assert abs(ra["Ih"] - rb["Ih"]) <= 1e-13 * abs(ra["Ih"])
        assert np.abs(a.x() - b.x()).max() <= 1e-13


def test_backward_euler_jacobian_is_fd_hessian(golden_dir):
    """FSubJac structure: FIXED rows carry 1 + (dt/tau) * (#incident elements in which the vertex is local vertex 0) on
    the diagonal and nothing else; INTERIOR rows are I + (dt/tau) * (forward-difference derivative of the assembled
    gradient at Vp), checked against a directional difference of eulerStepMod."""
    orc, spec = build_from_input("Monitor220", golden_dir)
    S = _method2_solver(orc, restated=True)
    a, rc = orc.build_euler_jac(S.ia, S.ja)
    assert rc == 0
    import scipy.sparse as sp
    n = orc.P * orc.D
    J = sp.csr_matrix((a, S.ja, S.ia), shape=(n, n)).toarray()
    D = orc.D
    Fo = orc.F_oriented()
    lead = np.bincount(Fo[:, 0], minlength=orc.P)
    s = spec["dt"] / spec["tau"]
    inter = np.repeat(orc.mask == 2, D)
    for v in np.nonzero(orc.mask == 1)[0]:
        for p in range(D):
            row = J[D * v + p].copy()
            assert row[D * v + p] == 1.0 + s * lead[v]
            row[D * v + p] = 0.0
            assert not row.any()
    # INTERIOR rows: row (v, p) holds d grad / d x_{v,p} (FSubJac stores the perturbed direction as the row), so for a
    # perturbation dv of the interior vertices  J^T dv = dv + (dt/tau) * (g(Vp + eps dv) - g(Vp)) / eps  on interior dofs,
    # g = the assembled INTERIOR gradient of eulerStepMod.  (J is NOT symmetric: blockGrad is not an exact gradient.)
    Ji = J[np.ix_(inter, inter)]
    x0 = orc.x().reshape(-1)
    rng = np.random.default_rng(5)
    dv = np.where(inter, rng.standard_normal(n), 0.0)
    eps = 10.0 * np.sqrt(np.finfo(float).eps)
    _, g0, _ = orc.euler_step_mod(x0)
    _, g1, _ = orc.euler_step_mod(x0 + eps * dv)
    want = dv[inter] + s * (g1 - g0)[inter] / eps
    got = Ji.T @ dv[inter]
    assert np.abs(got - want).max() < 1e-5 * np.abs(want).max()


def test_restated_cgstab_equals_lasolver_3d():
    """Same cross-check on a small 3D Jacobian (cube nx = 3, MEx33D, 3 x 3 blocks): pattern, iteration counts and iterates of
    the numpy ILU(0) + CGSTAB restatement against the reference's LASolver."""
    import pcg_oracle as PO
    if not PO.have_lasolver():
        pytest.skip("oracle/_ref not built")
    Vp, F, mask = O.uniform_rect_mesh(3, 3, 3, 3)
    mon = O.monitor_eval(3, 3, Vp)
    a = O.Oracle(Vp, F, mask, mon, 50.0, 0.5, 0.025)
    b = O.Oracle(Vp, F, mask, mon, 50.0, 0.5, 0.025)
    S, R = _method2_solver(a), _method2_solver(b, restated=True)
    assert np.array_equal(S.ia, R.ia) and np.array_equal(S.ja, R.ja)
    for _ in range(3):
        ra, rb = a.backwards_euler_step(S), b.backwards_euler_step(R)
        assert ra["status"] == 0 and rb["status"] == 0
        assert ra["inner_iters"] == rb["inner_iters"] and ra["newton_iters"] == rb["newton_iters"] >= 1
        assert abs(ra["Ih"] - rb["Ih"]) <= 1e-13 * abs(ra["Ih"])
        assert np.abs(a.x() - b.x()).max() <= 1e-13
