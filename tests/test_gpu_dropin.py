"""The reference's own driver (main.cpp, unchanged) built against the drop-in headers in
mm-admm_b200/host/dropin and linked with libmmadmm_b200.so (build/dropin/mesh.exe, built by
__graft_entry__.build() in the container where /root/reference is mounted) must reproduce the energy traces the
reference commits (Experiments/Results/<Name>/Ih0.txt, 6 significant digits) from the shipped input files.

This is the drop-in boundary test of SURVEY.md 8(b): same CLI, same JSON inputs, same result files.
Monitor110 additionally goes through the Shoulder set-up (main.cpp:403-630: removeRow, srand(69) + Vector::Random
perturbations, GradUse = true) -- a fifth committed trace, pinned end to end through the reference's own driver.
Nothing here reads /root/reference at run time: the inputs are regenerated from tests/golden/.
"""
import json
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "build", "dropin", "mesh.exe")
GOLD = os.path.join(ROOT, "tests", "golden")
TRACES = json.load(open(os.path.join(GOLD, "ih_traces.json")))


def run_driver(tmp_path, name, nsteps, extra=None, method=0):
    spec = dict(TRACES[name]["input"])
    spec["nSteps"] = nsteps
    spec["DtTol"] = 0          # never stop early: compare a fixed number of steps
    spec.update(extra or {})
    wd = tmp_path / name
    (wd / "Experiments" / "InputFiles").mkdir(parents=True)
    if spec["TestType"] == "FromFile":
        base = wd / "Experiments" / "Results" / ("BaseCircle3D" if spec["Dim"] == 3 else "BaseCircle")
        base.mkdir(parents=True)
        for key in ("MaskFile", "PntsFile", "TrianglesFile"):
            fn = os.path.basename(spec[key])
            shutil.copy(os.path.join(GOLD, fn), base / fn)
            spec[key] = "./" + os.path.relpath(base / fn, wd)
    json.dump(spec, open(wd / "Experiments" / "InputFiles" / f"{name}.json", "w"))
    r = subprocess.run([EXE, name, str(method), "1"], cwd=wd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    out = wd / "Experiments" / "Results" / name
    rows = [ln.split(",") for ln in open(out / f"Ih{method}.txt").read().split("\n") if ln.strip()]
    ih = [row[1].strip() for row in rows]
    return spec, out, ih, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("name,nsteps", [("Monitor310", 30), ("Monitor210", 19), ("Monitor220", 12), ("3DMonitor210", 6),
                                         ("Monitor110", 19)])
def test_reference_driver_reproduces_committed_trace(tmp_path, name, nsteps):
    if not os.path.exists(EXE):
        pytest.skip("build/dropin/mesh.exe not built (needs the reference's main.cpp at build time)")
    spec, out, ih, stdout = run_driver(tmp_path, name, nsteps)
    ref = TRACES[name]["Ih"]
    assert len(ih) == nsteps + 1
    for i, (g, e) in enumerate(zip(ih, ref)):
        assert g == ("%g" % e), f"{name} step {i}: drop-in driver wrote {g}, reference trace has {e:g}"
    # result files the reference writes (main.cpp:225-248)
    pts = [ln for ln in open(out / "points.txt").read().split("\n") if ln.strip()]
    tris = [ln for ln in open(out / "triangles.txt").read().split("\n") if ln.strip()]
    assert len(pts[0].split(",")) == spec["Dim"] and len(tris[0].split(",")) == spec["Dim"] + 1
    assert f"Number of simplices = {len(tris)}" in stdout and f"Number of points = {len(pts)}" in stdout
    assert os.path.exists(out / "IhPara1.txt")


@pytest.mark.gpu
def test_reference_driver_explicit_euler_reproduces_committed_trace(tmp_path):
    """`mesh.exe Monitor220 1 1`: Method 1 = MeshIntegrator::eulerStep (main.cpp:186-188) -> mmadmm_euler_step; the reference
    commits Experiments/Results/Monitor220/Ih1.txt for the shipped input (49 steps, 6 digits)."""
    if not os.path.exists(EXE):
        pytest.skip("build/dropin/mesh.exe not built")
    ref = TRACES["Monitor220"]["Ih1"]
    spec, out, ih, stdout = run_driver(tmp_path, "Monitor220", len(ref) - 1, method=1)
    assert len(ih) == len(ref) == 50
    for i, (g, e) in enumerate(zip(ih, ref)):
        assert g == ("%g" % e), f"euler step {i}: drop-in driver wrote {g}, reference trace has {e:g}"


@pytest.mark.gpu
def test_reference_driver_level_set_mesh_aborts_like_the_reference(tmp_path):
    """LevelSet set-up (main.cpp:257-400 -> utils::meshFromLevelSetFun).  The reference ships no LevelSet input, and on
    the disc it cuts (thin boundary elements) its own algorithm inverts an element in the very first step -- the
    oracle reports status 1 there for every parameter set tried (tests/test_dropin_host.py pins the generator
    itself).  The reference then dies on assert(Edet > 0) (src/AdaptationFunctional.cpp:174); the drop-in driver must
    generate the same mesh sizes and abort with EINVERTED instead of continuing on garbage."""
    if not os.path.exists(EXE):
        pytest.skip("build/dropin/mesh.exe not built")
    import numpy as np
    import oracle_lib as O
    from test_dropin_host import reference_level_set_2d
    spec = {"TestType": "LevelSet", "Dim": 2, "MonType": 1, "Method": 0, "CompMesh": False, "BoundaryType": 1,
            "GradUse": False, "nSteps": 8, "AdmmIter": 10, "DtTol": 0, "dt": 0.025, "tau": 0.5, "rho": 50, "w": 3.5355,
            "nx": 24, "ny": 24, "xa": 0, "xb": 1, "ya": 0, "yb": 1}
    V, F, mask = reference_level_set_2d(24, 24, 1)
    orc = O.Oracle(V, F.astype(np.int32), mask[: len(V)].astype(np.int32), O.monitor_eval(2, 1, V), 50.0, 0.5, 0.025)
    assert orc.step(10, 1e-3)["status"] == 1          # the reference algorithm itself inverts here
    wd = tmp_path / "ls"
    (wd / "Experiments" / "InputFiles").mkdir(parents=True)
    (wd / "Experiments" / "Results" / "ls").mkdir(parents=True)    # phi.txt is written before runAlgo's mkdir
    json.dump(spec, open(wd / "Experiments" / "InputFiles" / "ls.json", "w"))
    r = subprocess.run([EXE, "ls", "0", "1"], cwd=wd, capture_output=True, text=True, timeout=300)
    assert f"size of Vp {len(V)}, 2" in r.stdout and f"size of F {len(F)}, 3" in r.stdout
    assert f"size of mask {len(mask)}" in r.stdout                 # the un-compacted mask of the reference
    assert r.returncode != 0 and "EINVERTED" in r.stderr
    phi = np.loadtxt(wd / "Experiments" / "Results" / "ls" / "phi.txt", delimiter=",")
    assert phi.shape == (25 * 25, 3)


@pytest.mark.gpu
@pytest.mark.parametrize("name,nsteps,wb", [("Monitor210", 12, "deferred"), ("Monitor210", 12, "strict"), ("3DMonitor210", 5, "deferred")])
def test_reference_driver_final_points_match_oracle(tmp_path, name, nsteps, wb, monkeypatch):
    """points.txt written by the reference's driver after done() (main.cpp:225-236: Mesh::outputPoints of the caller's Vp, which
    the library refreshed through its column-major write-back) equals the oracle's final node positions at the 6 printed digits,
    with the per-step write-back overlapped (default) or strict (MMADMM_WRITEBACK=strict)."""
    if not os.path.exists(EXE):
        pytest.skip("build/dropin/mesh.exe not built")
    import numpy as np
    import oracle_lib as O
    from test_oracle_kat import build_from_input
    monkeypatch.setenv("MMADMM_WRITEBACK", wb)
    spec, out, ih, stdout = run_driver(tmp_path, name, nsteps)
    orc, _ = build_from_input(name, GOLD, nthreads=4)
    for _ in range(nsteps):
        assert orc.step(spec["AdmmIter"], 1e-3)["status"] == 0
    xo = orc.x()
    pts = np.array([[float(w) for w in ln.split(",")] for ln in open(out / "points.txt").read().split("\n") if ln.strip()])
    assert pts.shape == xo.shape
    want = np.array([[float("%g" % v) for v in row] for row in xo])
    assert np.abs(pts - want).max() <= 1e-6 * np.abs(want).max()      # one unit of the 6th printed digit at most
