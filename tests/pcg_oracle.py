"""Oracle side of the general sparse x-update path (test infrastructure).

* `lasolver_solve`: the reference's own lib/LASolver, built from /root/reference into oracle/_ref/liblasolver_ref.so
  (ILU level 0 + CG accelerator, the configuration of src/Mesh.cpp:266-304 with iaccel = -1).
* `ilu0`, `pcg`: plain numpy restatements (ILU(0) in IKJ order on A's pattern = LASolver's level-0 factor,
  ILU_class.cpp:300-444; textbook PCG = accel_class.cpp:402-491) for factor-level and iteration-level comparisons.
* matrices: the ADMM system matrix t = M + dt^2 (WD)^T (WD) exactly as the reference assembles it (D x D blocks with
  explicit zeros, src/Mesh.cpp:676-753, src/MeshIntegrator.cpp:45-51) and a genuinely non-diagonal SPD one
  (tau I + dt^2 w^2 * vertex graph Laplacian blocks) on the same mesh."""
import ctypes as C
import os

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "liblasolver_ref.so")


def have_lasolver():
    return os.path.exists(REF_SO)


def lasolver_solve(A, b, tol, nitmax=10000, iaccel=-1, level=0):
    L = C.CDLL(REF_SO)
    A = sp.csr_matrix(A); A.sort_indices()
    ia = A.indptr.astype(np.int32); ja = A.indices.astype(np.int32); a = A.data.astype(np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros(A.shape[0]); it = C.c_int()
    ip, dp = C.POINTER(C.c_int), C.POINTER(C.c_double)
    L.lasolver_solve.argtypes = [C.c_int, ip, ip, dp, dp, C.c_int, C.c_int, C.c_double, C.c_int, dp, ip]
    rc = L.lasolver_solve(A.shape[0], ia.ctypes.data_as(ip), ja.ctypes.data_as(ip), a.ctypes.data_as(dp), b.ctypes.data_as(dp),
                          iaccel, level, float(tol), nitmax, x.ctypes.data_as(dp), C.byref(it))
    assert rc == 0
    return x, it.value


def ilu0(A):
    """ILU(0) on A's pattern, IKJ order; returns the factor values in A's (sorted) CSR layout."""
    A = sp.csr_matrix(A); A.sort_indices()
    ip, ix, lu = A.indptr, A.indices, A.data.astype(np.float64).copy()
    n = A.shape[0]
    diag = np.array([ip[i] + int(np.searchsorted(ix[ip[i]:ip[i + 1]], i)) for i in range(n)])
    for i in range(n):
        pos = {int(ix[k]): k for k in range(ip[i], ip[i + 1])}
        for k in range(ip[i], diag[i]):
            c = int(ix[k])
            lik = lu[k] / lu[diag[c]]
            lu[k] = lik
            for kk in range(diag[c] + 1, ip[c + 1]):
                pj = pos.get(int(ix[kk]))
                if pj is not None:
                    lu[pj] -= lik * lu[kk]
    return lu, diag


def ilu_apply(A, lu, diag, r):
    A = sp.csr_matrix(A); A.sort_indices()
    ip, ix = A.indptr, A.indices
    n = A.shape[0]
    y = np.zeros(n)
    for i in range(n):
        s = r[i]
        for k in range(ip[i], diag[i]):
            s -= lu[k] * y[ix[k]]
        y[i] = s
    x = np.zeros(n)
    for i in range(n - 1, -1, -1):
        s = y[i]
        for k in range(diag[i] + 1, ip[i + 1]):
            s -= lu[k] * x[ix[k]]
        x[i] = s / lu[diag[i]]
    return x


def pcg(A, b, tol, maxit, prec):
    x = np.zeros_like(b); r = b.copy(); z = prec(r); p = z.copy(); rz = r @ z
    bb = b @ b
    it = 0
    while r @ r > tol * tol * bb and it < maxit:
        q = A @ p
        alpha = rz / (p @ q)
        x += alpha * p; r -= alpha * q
        z = prec(r); rzn = r @ z
        p = z + (rzn / rz) * p; rz = rzn
        it += 1
    return x, it


def admm_matrix(F, P, D, tau, dt, rho):
    """t = M + dt^2 (WD)^T (WD) with the reference's structure: D selects z_e,v = x_{F[e,v]} with a D x D block per slot whose
    off-diagonal entries are explicit zeros (buildDMatrix, src/Mesh.cpp:703-753); W = w I, w = 0.5 sqrt(rho); M = tau I."""
    w = 0.5 * np.sqrt(rho)
    val = np.bincount(np.asarray(F).ravel(), minlength=P)
    d = tau + (dt * dt) * ((w * w) * val.astype(np.float64))
    # CSR with the D x D block of every vertex stored in full: (WD)^T(WD) keeps D's explicit zeros structurally
    indptr = np.arange(0, P * D * D + 1, D, dtype=np.int32)
    indices = (np.repeat(np.arange(P) * D, D * D).reshape(P, D, D) + np.arange(D)[None, None, :]).reshape(-1).astype(np.int32)
    data = np.zeros((P, D, D))
    for a in range(D):
        data[:, a, a] = d
    return sp.csr_matrix((data.reshape(-1), indices, indptr), shape=(P * D, P * D))


def laplacian_like_matrix(F, P, D, tau, dt, rho):
    """SPD and genuinely coupled: tau I + dt^2 w^2 sum over elements of the (nv x nv) element graph Laplacian, one copy per
    coordinate -- the sparsity pattern Method 2's Jacobian has (src/Mesh.cpp:306-338: all vertex pairs of every element)."""
    w = 0.5 * np.sqrt(rho)
    E, nv = F.shape
    rows, cols, vals = [], [], []
    for e in range(E):
        for i in range(nv):
            for j in range(nv):
                for a in range(D):
                    rows.append(F[e, i] * D + a); cols.append(F[e, j] * D + a); vals.append((nv - 1.0) if i == j else -1.0)
    L = sp.coo_matrix((vals, (rows, cols)), shape=(P * D, P * D)).tocsr()
    A = (tau * sp.identity(P * D, format="csr") + (dt * dt * w * w) * L).tocsr()
    A.sort_indices()
    return A


class LASolverJac:
    """The reference's LASolver matrix object driven as Method 2 drives it (oracle/lasolver_shim.cpp las_*): structure from
    the set_entry sequence of Mesh::buildMatrix, ParamIter of src/Mesh.cpp:266-304 (ILU(0), CGSTAB, resid_reduc 1e-6)."""

    def __init__(self, n, rows, cols):
        L = C.CDLL(REF_SO)
        L.las_open.restype = C.c_void_p
        L.las_open.argtypes = [C.c_int, C.c_longlong, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.las_nnz.argtypes = [C.c_void_p]
        L.las_pattern.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.las_sfac.argtypes = [C.c_void_p]
        L.las_solve.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                C.POINTER(C.c_int)]
        L.las_close.argtypes = [C.c_void_p]
        self.L, self.n = L, n
        rows = np.ascontiguousarray(rows, dtype=np.int32); cols = np.ascontiguousarray(cols, dtype=np.int32)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int))
        self.h = L.las_open(n, len(rows), ip(rows), ip(cols))
        assert self.h
        nnz = L.las_nnz(self.h)
        self.ia = np.zeros(n + 1, dtype=np.int32); self.ja = np.zeros(nnz, dtype=np.int32)
        L.las_pattern(self.h, ip(self.ia), ip(self.ja))

    def sfac(self):
        assert self.L.las_sfac(self.h) == 0

    def solve(self, a, b):
        dp = lambda v: v.ctypes.data_as(C.POINTER(C.c_double))
        a = np.ascontiguousarray(a, dtype=np.float64); b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(self.n); nit = C.c_int()
        assert self.L.las_solve(self.h, dp(a), dp(b), dp(x), C.byref(nit)) == 0
        return x, nit.value

    def __del__(self):
        if getattr(self, "h", None):
            self.L.las_close(self.h)
            self.h = None


def cgstab(A, b, ctol, nitmax, prec):
    """LASolver's scaler_cgstab restated (accel_class.cpp:229-273 init, :280-399 one iteration; driver loop
    MatrixIter.cpp:693-800 with a zero initial guess and toler = 0): right-preconditioned BiCGSTAB with rhat = r0,
    alpha = rholst = omega = 1 and p = A vbar = 0 at the start, `tiny` = 1e-300 guards in beta and omega, stop when
    ||r|| / ||r0|| < ctol (or when no component moved)."""
    tiny = 1.e-300
    n = len(b)
    x = np.zeros(n); res = b.astype(np.float64).copy(); res0 = res.copy()
    pvec = np.zeros(n); avbar = np.zeros(n)
    alpha = rholst = omega = 1.0
    rmsi = np.sqrt(res @ res)
    nitr = 0
    for _ in range(nitmax):
        nitr += 1
        rho = res0 @ res
        beta = rho / (rholst + tiny)
        beta *= alpha / (omega + tiny)
        rholst = rho
        pvec = res + beta * (pvec - omega * avbar)
        vbar = prec(pvec)
        avbar = A @ vbar
        alpha = rho / (res0 @ avbar)
        svec = res - alpha * avbar
        zvec = prec(svec)
        tvec = A @ zvec
        omega = (tvec @ svec) / (tvec @ tvec + tiny)
        step = alpha * vbar + omega * zvec
        x += step
        res = svec - omega * tvec
        rms = np.sqrt(res @ res)
        if not np.any(np.abs(step) > 0.0) or rms / rmsi < ctol:
            return x, nitr
    return x, -1


class RestatedJac:
    """Same interface as LASolverJac with nothing from the reference behind it: numpy ILU(0) (`ilu0`) applied through two
    scipy triangular solves, and `cgstab`.  Used where oracle/_ref is the thing being cross-checked, and as the Method 2
    linear-solve oracle that needs no reference build."""

    def __init__(self, n, rows, cols):
        pat = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(n, n))
        pat.sum_duplicates(); pat.sort_indices()
        self.n, self.ia, self.ja = n, pat.indptr.astype(np.int32), pat.indices.astype(np.int32)

    def sfac(self):
        pass  # level-0 fill: the factor lives on A's own pattern

    def solve(self, a, b, resid_reduc=1.e-6, nitmax=10000):
        from scipy.sparse.linalg import spsolve_triangular
        A = sp.csr_matrix((np.asarray(a, dtype=np.float64), self.ja, self.ia), shape=(self.n, self.n))
        lu, diag = ilu0(A)
        M = sp.csr_matrix((lu, self.ja, self.ia), shape=(self.n, self.n))
        Lm = (sp.tril(M, -1) + sp.identity(self.n)).tocsr()
        Um = sp.triu(M, 0).tocsr()
        prec = lambda r: spsolve_triangular(Um, spsolve_triangular(Lm, r, lower=True), lower=False)
        return cgstab(A, np.asarray(b, dtype=np.float64), resid_reduc, nitmax, prec)
