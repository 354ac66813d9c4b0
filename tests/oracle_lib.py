"""ctypes wrapper around oracle/liboracle.so -- TEST INFRASTRUCTURE (see oracle/mmadmm_oracle.cpp header).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_LIB = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


def _dp(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def build(force=False):
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "mmadmm_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_int, C.c_int, c_dp, c_dp, c_ip, c_ip, c_dp,
                                 C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_clone.restype = C.c_void_p
        L.orc_clone.argtypes = [C.c_void_p]
        L.orc_set_threads.argtypes = [C.c_void_p, C.c_int]
        L.orc_set_projection.argtypes = [C.c_void_p, C.c_int]
        L.orc_face_count.argtypes = [C.c_void_p]
        L.orc_update_monitor.argtypes = [C.c_void_p, c_dp]
        L.orc_project.argtypes = [C.c_void_p, C.c_int, c_dp]
        L.orc_step.restype = C.c_int
        L.orc_step.argtypes = [C.c_void_p, C.c_int, C.c_double, c_dp, c_ip, c_dp, c_dp]
        L.orc_euler_step.restype = C.c_int
        L.orc_euler_step.argtypes = [C.c_void_p, c_dp]
        L.orc_energy.restype = C.c_double
        L.orc_energy.argtypes = [C.c_void_p]
        for nm in ("orc_get_x", "orc_get_z", "orc_get_u", "orc_get_hess", "orc_get_grid"):
            getattr(L, nm).argtypes = [C.c_void_p, c_dp]
        L.orc_get_F.argtypes = [C.c_void_p, c_ip]
        L.orc_grid_dims.argtypes = [C.c_void_p, c_ip]
        L.orc_get_axes.argtypes = [C.c_void_p, c_dp, c_dp, c_dp]
        L.orc_get_stats.argtypes = [C.c_void_p, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        L.orc_w.restype = C.c_double
        L.orc_w.argtypes = [C.c_void_p]
        L.orc_eval_monitor.argtypes = [C.c_void_p, c_dp, C.c_int, c_dp]
        L.orc_block_grad.argtypes = [C.c_void_p, C.c_int, c_dp, c_dp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, c_dp]
        L.orc_monitor_eval.argtypes = [C.c_int, C.c_int, C.c_int, c_dp, c_dp]
        L.orc_lu_inverse.restype = C.c_int
        L.orc_lu_inverse.argtypes = [C.c_int, c_dp, c_dp]
        L.orc_euler_step_mod.restype = C.c_int
        L.orc_euler_step_mod.argtypes = [C.c_void_p, c_dp, c_dp, c_dp]
        L.orc_build_euler_jac.restype = C.c_int
        L.orc_build_euler_jac.argtypes = [C.c_void_p, c_ip, c_ip, c_dp]
        L.orc_set_x.argtypes = [C.c_void_p, c_dp]
        L.orc_uniform_rect_mesh.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_double] * 6 + [C.c_int, c_dp, c_ip, c_ip]
        _LIB = L
    return _LIB


def uniform_rect_mesh(D, nx, ny, nz=0, box=(0, 1, 0, 1, 0, 1), btype=1):
    """utils::generateUniformRectMesh<D> restated (src/MeshUtils.h:82-335)."""
    if D == 2:
        P, E = (nx + 1) * (ny + 1) + nx * ny, 4 * nx * ny
    else:
        P, E = (nx + 1) * (ny + 1) * (nz + 1) + nx * ny * nz, 12 * nx * ny * nz
    Vp = np.zeros((P, D)); F = np.zeros((E, D + 1), dtype=np.int32); mask = np.zeros(P, dtype=np.int32)
    lib().orc_uniform_rect_mesh(D, nx, ny, nz, *[float(b) for b in box], btype, _dp(Vp), _ip(F), _ip(mask))
    return Vp, F, mask


def monitor_eval(D, kind, X):
    X = np.ascontiguousarray(X, dtype=np.float64)
    out = np.zeros((X.shape[0], D * D))
    lib().orc_monitor_eval(D, kind, X.shape[0], _dp(X), _dp(out))
    return out


def lu_inverse(A):
    A = np.ascontiguousarray(A, dtype=np.float64)
    out = np.zeros_like(A)
    rc = lib().orc_lu_inverse(A.shape[0], _dp(A), _dp(out))
    assert rc == 0
    return out


class Oracle:
    """Mesh<D> + MeshIntegrator<D> of the reference, restated on the CPU."""

    def __init__(self, Vp, F, mask, mon_vertex, rho, tau, dt, Vc=None, grad_use=False, nthreads=1):
        self.L = lib()
        self.Vp = np.ascontiguousarray(Vp, dtype=np.float64)
        self.P, self.D = self.Vp.shape
        self.F = np.ascontiguousarray(F, dtype=np.int32)
        self.E = self.F.shape[0]
        self.n = self.D * (self.D + 1)
        self.mask = np.ascontiguousarray(mask, dtype=np.int32)[: self.P].copy()
        self.mon = np.ascontiguousarray(mon_vertex, dtype=np.float64)
        self.Vc = None if Vc is None else np.ascontiguousarray(Vc, dtype=np.float64)
        self.h = self.L.orc_create(self.D, self.P, self.E, _dp(self.Vp), _dp(self.Vc), _ip(self.F), _ip(self.mask),
                                   _dp(self.mon), rho, tau, dt, int(grad_use), nthreads)
        self._dt, self._tau = dt, tau
        self._m2_step_taken, self._m2_a = False, None

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def clone(self):
        """Deep copy of the whole oracle state (mesh, monitor grid, integrator vectors, inverse Hessians, counters)."""
        import copy
        other = copy.copy(self)
        other.h = self.L.orc_clone(self.h)
        return other

    def set_projection(self, mask):
        """Boundary sliding option (projectOntoBoundary, src/Mesh.cpp:118-241): bit 0 = after every element's BFGS solve in prox
        (:975-984), bit 1 = in updateAfterStep (:1020-1026).  0 (default) = what the reference executes: no projection."""
        self.L.orc_set_projection(self.h, int(mask))

    def project(self, node_id, pnt):
        p = np.ascontiguousarray(pnt, dtype=np.float64).copy()
        self.L.orc_project(self.h, int(node_id), _dp(p))
        return p

    def face_count(self):
        return self.L.orc_face_count(self.h)

    def update_monitor(self, mon_vertex):
        """Rebuild the monitor grid from the current node positions (the commented-out Mesh::setUp, src/Mesh.cpp:1006-1014)."""
        self.mon = np.ascontiguousarray(mon_vertex, dtype=np.float64)
        self.L.orc_update_monitor(self.h, _dp(self.mon))

    def set_threads(self, nthreads):
        self.L.orc_set_threads(self.h, int(nthreads))

    def step(self, n_iters, tol=1e-3):
        Ih = C.c_double(); it = C.c_int(); pr = C.c_double(); du = C.c_double()
        rc = self.L.orc_step(self.h, n_iters, tol, C.byref(Ih), C.byref(it), C.byref(pr), C.byref(du))
        return dict(Ih=Ih.value, iters=it.value, primal=pr.value, dual=du.value, status=rc)

    def euler_step(self):
        """Method 1 (MeshIntegrator::eulerStep): returns the energy at entry; Vp is not refreshed."""
        Ih = C.c_double()
        rc = self.L.orc_euler_step(self.h, C.byref(Ih))
        return dict(Ih=Ih.value, status=rc)

    def energy(self):
        return self.L.orc_energy(self.h)

    # ---- Method 2 (backward Euler): Mesh::backwardsEulerStep, src/Mesh.cpp:1263-1341 ----
    def euler_step_mod(self, x):
        """Mesh::eulerStepMod (src/Mesh.cpp:533-579): energy and INTERIOR-only assembled gradient at x."""
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1)
        g = np.zeros_like(x); Ih = C.c_double()
        rc = self.L.orc_euler_step_mod(self.h, _dp(x), _dp(g), C.byref(Ih))
        return Ih.value, g, rc

    def jacobian_entries(self):
        """The set_entry sequence of Mesh::buildMatrix (src/Mesh.cpp:305-337): per element, each rotation of its vertex
        list couples the D rows of the leading vertex with the D columns of every vertex of the element."""
        D = self.D
        Fo = self.F_oriented()
        rows, cols = [], []
        rr, cc = np.meshgrid(np.arange(D), np.arange(D), indexing="ij")
        for it in range(D + 1):
            lead = Fo[:, it]
            for nn in range(D + 1):
                other = Fo[:, (it + nn) % (D + 1)]
                rows.append((lead[:, None] * D + rr.reshape(1, -1)).reshape(self.E, -1))
                cols.append((other[:, None] * D + cc.reshape(1, -1)).reshape(self.E, -1))
        # element-major order like the reference's loop nest (the packed CSR does not depend on it)
        rows = np.concatenate(rows, axis=1).reshape(-1).astype(np.int32)
        cols = np.concatenate(cols, axis=1).reshape(-1).astype(np.int32)
        return rows, cols

    def build_euler_jac(self, ia, ja):
        """Mesh::buildEulerJac + FSubJac (src/Mesh.cpp:1111-1261) on the CSR pattern (ia, ja): I + (dt/tau) * FD Hessian
        taken at Vp."""
        a = np.zeros(int(ia[-1]))
        rc = self.L.orc_build_euler_jac(self.h, _ip(ia), _ip(ja), _dp(a))
        return a, rc

    def backwards_euler_step(self, solver, tol=1e-3, max_iters=1000):
        """One Mesh::backwardsEulerStep(dt, x, grad, tol) (src/Mesh.cpp:1263-1341) with `solver` doing what
        jac->sfac / jac->solve do there (tests/pcg_oracle.LASolverJac = the reference's own LASolver from oracle/_ref).
        Returns the energy of the last evaluated iterate, the Newton iteration count and the inner iteration counts."""
        dt_tau = self._dt / self._tau
        x = self.x().reshape(-1).copy()
        xn = x.copy()
        Ih, grad, rc = self.euler_step_mod(x)
        x -= dt_tau * grad
        if not self._m2_step_taken:
            self._m2_a, rc2 = self.build_euler_jac(solver.ia, solver.ja)
            rc |= rc2
            solver.sfac()
        n_iter, g1_prev, inner = 0, np.inf, []
        while True:
            Ih, grad, rc2 = self.euler_step_mod(x)
            rc |= rc2
            grad *= dt_tau
            grad += (x - xn)
            g1 = float(np.sum(np.abs(grad)))
            if g1 < 0.1 * tol:
                break
            if (not self._m2_step_taken) or abs(g1 - g1_prev) / g1 < 0.25:
                self._m2_a, rc2 = self.build_euler_jac(solver.ia, solver.ja)
                rc |= rc2
                self._m2_step_taken = True
            dx, nit = solver.solve(self._m2_a, -grad)
            assert nit > 0
            inner.append(nit)
            x += dx
            n_iter += 1
            g1_prev = g1
            if n_iter >= max_iters:
                break
        self.L.orc_set_x(self.h, _dp(x))
        return dict(Ih=Ih, newton_iters=n_iter, inner_iters=inner, res1=g1, status=rc)

    def _vec(self, fn, size):
        out = np.zeros(size)
        getattr(self.L, fn)(self.h, _dp(out))
        return out

    def x(self):
        return self._vec("orc_get_x", self.P * self.D).reshape(self.P, self.D)

    def z(self):
        return self._vec("orc_get_z", self.E * self.n).reshape(self.E, self.n)

    def u(self):
        return self._vec("orc_get_u", self.E * self.n).reshape(self.E, self.n)

    def hess(self):
        return self._vec("orc_get_hess", self.E * self.n * self.n).reshape(self.E, self.n, self.n)

    def F_oriented(self):
        out = np.zeros((self.E, self.D + 1), dtype=np.int32)
        self.L.orc_get_F(self.h, _ip(out))
        return out

    def grid_dims(self):
        d = np.zeros(3, dtype=np.int32)
        self.L.orc_grid_dims(self.h, _ip(d))
        return tuple(int(v) for v in d)

    def grid(self):
        nx, ny, nz = self.grid_dims()
        rows = (nx + 1) * (ny + 1) * (nz + 1)
        return self._vec("orc_get_grid", rows * self.D * self.D).reshape(rows, self.D * self.D)

    def axes(self):
        nx, ny, nz = self.grid_dims()
        ax = np.zeros(nx + 1); ay = np.zeros(ny + 1); az = np.zeros(nz + 1)
        self.L.orc_get_axes(self.h, _dp(ax), _dp(ay), _dp(az))
        return ax, ay, az

    def stats(self):
        hist = (C.c_longlong * 52)(); calls = C.c_longlong()
        self.L.orc_get_stats(self.h, hist, C.byref(calls))
        return np.array(list(hist), dtype=np.int64), calls.value

    def w(self):
        return self.L.orc_w(self.h)

    def eval_monitor(self, pts):
        pts = np.ascontiguousarray(pts, dtype=np.float64)
        out = np.zeros((pts.shape[0], self.D * self.D))
        self.L.orc_eval_monitor(self.h, _dp(pts), pts.shape[0], _dp(out))
        return out

    def block_grad(self, zs, dxpu=None, compute_grad=True, regularize=False, apply_mask=False):
        zs = np.ascontiguousarray(zs, dtype=np.float64)
        ne = zs.shape[0]
        dx = None if dxpu is None else np.ascontiguousarray(dxpu, dtype=np.float64)
        en = np.zeros(ne); g = np.zeros((ne, self.n)); igt = np.zeros(ne)
        self.L.orc_block_grad(self.h, ne, _dp(zs), _dp(dx), int(compute_grad), int(regularize), int(apply_mask),
                              _dp(en), _dp(g), _dp(igt))
        return en, g, igt


def read_triangles(D, tri_file, pnt_file, mask_file):
    """utils::readTriangles (src/MeshUtils.h:669-733)."""
    tri = []
    with open(tri_file) as f:
        for line in f:
            tri += [int(w) for w in line.strip().split(",") if w.strip()]
    pts = []
    with open(pnt_file) as f:
        for line in f:
            pts += [float(w) for w in line.strip().split(",") if w.strip()]
    with open(mask_file) as f:
        mask = [int(w) for w in f.read().split()]
    F = np.array(tri, dtype=np.int32).reshape(-1, D + 1)
    Vp = np.array(pts, dtype=np.float64).reshape(-1, D)
    mask = np.array(mask + [mask[-1]], dtype=np.int32)  # the reference's `while (fin)` loop pushes one extra entry
    return Vp, F, mask
