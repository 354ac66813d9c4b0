"""Host logic of the C++ drop-in layer (mm-admm_b200/host/dropin): the Eigen / nlohmann shims and the driver-layer mesh
generators, checked on CPU against the oracle's generators and a literal numpy restatement of the reference's
meshFromLevelSetFun (src/MeshUtils.h:404-538, including its quirks)."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "mm-admm_b200", "host", "dropin")


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    d = tmp_path_factory.mktemp("dropin")
    exe = str(d / "check")
    lib = os.path.join(ROOT, "mm-admm_b200", "libmmadmm_b200.so")
    if not os.path.exists(lib):
        pytest.skip("libmmadmm_b200.so not built")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-w", "-I" + DROPIN, "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "dropin_host_check.cpp"), "-o", exe,
                           "-L" + os.path.dirname(lib), "-lmmadmm_b200", "-Wl,-rpath," + os.path.dirname(lib)])
    spec = {"TestType": "SquareGrid", "nx": 40, "dt": 0.025, "CompMesh": False, "tau": 1e-1, "rho": 5, "Method": 0}
    json.dump(spec, open(d / "in.json", "w"), indent=4)
    out = subprocess.run([exe, str(d)], capture_output=True, text=True, check=True).stdout
    return d, out


def load(fn):
    with open(fn) as f:
        P, D, E, K, M = (int(v) for v in f.readline().split())
        V = np.array([[float(v) for v in f.readline().split()] for _ in range(P)])
        F = np.array([[int(v) for v in f.readline().split()] for _ in range(E)])
        mask = np.array([int(f.readline()) for _ in range(M)])
    return V, F, mask


def test_box_meshes_match_oracle_generator(checker):
    d, _ = checker
    V, F, mask = load(d / "box2.txt")
    Vo, Fo, mo = O.uniform_rect_mesh(2, 5, 5, 0, btype=0)
    assert np.array_equal(V, Vo) and np.array_equal(F, Fo) and np.array_equal(mask, mo)
    V, F, mask = load(d / "box3.txt")
    Vo, Fo, mo = O.uniform_rect_mesh(3, 3, 3, 3, btype=1)
    assert np.array_equal(V, Vo) and np.array_equal(F, Fo) and np.array_equal(mask, mo)


def reference_level_set_2d(nx, ny, btype):
    """src/MeshUtils.h:404-538 restated literally (quadratic loops and all); small sizes only."""
    EPS = 1e-12
    phi = lambda x, y: np.sqrt((x - 0.5) * (x - 0.5) + (y - 0.5) * (y - 0.5)) - 0.35
    Vp, F, _ = O.uniform_rect_mesh(2, nx, ny, 0, btype=btype)
    mask = np.full(len(Vp), 2)
    F = [list(r) for r in F]
    ids = [s for s, t in enumerate(F) if all(phi(*Vp[v]) > -EPS for v in t)]
    for s in sorted(ids, reverse=True):
        del F[s]
    used = sorted({v for t in F for v in t})
    for p in used:
        x = Vp[p].copy()
        ph = phi(*x)
        if abs(ph) < EPS or ph > 0:
            xv, yv = x[0] - 0.5, x[1] - 0.5
            g = np.array([xv / np.sqrt(xv * xv + yv * yv), yv / np.sqrt(xv * xv + yv * yv)])
            x = x - phi(*x) * g
            mask[p] = btype
        Vp[p] = x
    pmap = {off: i for i, off in enumerate(used)}
    Vn = np.array([Vp[off] for off in used])
    for pnt in used:                      # the reference's in-place relabelling sweep
        for t in F:
            for c in range(3):
                if t[c] == pnt:
                    t[c] = pmap[pnt]
    for p in range(len(Vn)):
        if abs(phi(*Vn[p])) < EPS:
            mask[p] = 1
    return Vn, np.array(F), mask


def test_level_set_mesh_matches_reference_algorithm(checker):
    d, _ = checker
    V, F, mask = load(d / "disc.txt")
    Vr, Fr, mr = reference_level_set_2d(12, 12, 0)
    assert np.array_equal(F, Fr)
    assert np.array_equal(V, Vr)
    assert np.array_equal(mask, mr)       # un-compacted mask, old-index marks included


def test_shims(checker):
    _, out = checker
    lines = dict((ln.split(" ", 1)[0], ln.split(" ", 1)[1]) for ln in out.strip().split("\n") if not ln.startswith(("mon", "huang", "meshio")))
    # MeshIO.h: the parallel writers give the bytes of the reference's per-line ostream loops (src/Mesh.cpp:1049-1095), the binary
    # twin round-trips exactly, readTriangles takes both formats
    assert "meshio text identical" in out and "meshio bin exact" in out and "meshio read ok" in out
    assert lines["removeRow"] == "ok"
    mons = [[float(v) for v in ln.split()[1:]] for ln in out.split("\n") if ln.startswith("mon ")]
    X = np.array([[0.1, 0.2], [0.5, 0.5], [0.9, 0.3]])
    want = O.monitor_eval(2, 1, X)        # the oracle's MEx1
    assert np.allclose(np.array(mons), want, rtol=1e-15, atol=0)
    a = [float(v) for v in lines["aniso"].replace("det", "").split()]
    assert np.allclose(a[:4], [1.75, 1.25, 1.25, 1.75], rtol=1e-15) and abs(a[4] - 1.5) < 1e-14
    assert lines["json"] == "40 0.025000000000000001 0 SquareGrid SquareGrid 0 1"   # cout precision 17
    # nlohmann's compact dump: keys sorted, integers stay integers
    assert lines["dump"] == '{"CompMesh":false,"Method":2,"TestType":"SquareGrid","dt":0.025,"nx":40,"rho":5,"tau":0.1}'


def test_huang_functional_virtual_interface(checker):
    """HuangFunctional<D>::G / dGdJ / dGddet / dGdM / dGdX (src/HuangFunctional.h:26-38, src/HuangFunctional.cpp:35-90) kept in the
    drop-in headers: values against a numpy restatement of the same formulas (p = 3/2, theta = 1/3 hard-coded; dGddet's
    theta = 1/2 quirk makes it 0), and dGdJ against a central difference of G at fixed determinant."""
    _, out = checker
    h = {}
    for ln in out.split("\n"):
        if ln.startswith("huang "):
            w = ln.split()
            h[w[1]] = w[2:]
    J = np.array([[1.3, 0.2], [-0.1, 0.9]]); Minv = np.array([[0.7, 0.1], [0.1, 1.4]])
    d, p, th = 2.0, 1.5, 1.0 / 3.0
    det = np.linalg.det(J)
    detM = np.sqrt(1.0 / np.linalg.det(Minv))
    tr = np.trace(J @ Minv @ J.T)
    G = th * detM * tr ** (d * p / 2) + (1 - 2 * th) * d ** (d * p / 2) * detM * (det / detM) ** p
    assert abs(float(h["G"][0]) - G) <= 1e-13 * abs(G)
    assert float(h["G"][2]) == 0.0 and [float(v) for v in h["G"][4:6]] == [0.0, 0.0]        # dGddet (quirk), dGdX
    dJ = d * p * th * detM * tr ** (d * p / 2 - 1) * (Minv @ J.T)
    assert np.allclose([float(v) for v in h["dGdJ"]], dJ.reshape(-1), rtol=1e-13)
    dM = (-0.5 * th * d * p * detM * tr ** (d * p / 2 - 1) * (Minv.T @ J.T @ J @ Minv) + 0.5 * th * detM * tr ** (d * p / 2) * Minv
          + ((0.5 - th) * (1 - p) * d ** (d * p / 2)) * detM ** (1 - p) * det ** p * Minv)
    assert np.allclose([float(v) for v in h["dGdM"]], dM.reshape(-1), rtol=1e-13)
    assert np.allclose([float(v) for v in h["fdJ"]], dJ.reshape(-1), rtol=1e-7)            # out(j, i) = dG/dJ(i, j)
