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
