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
