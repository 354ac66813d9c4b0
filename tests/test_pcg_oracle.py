"""CPU checks of the oracle for the general sparse path: the reference's LASolver (oracle/_ref) against scipy and against
the numpy restatements the GPU tests compare with."""
import numpy as np
import pytest
import scipy.sparse.linalg as spla

import oracle_lib as O
import pcg_oracle as PO


def small_mesh(D):
    Vp, F, mask = O.uniform_rect_mesh(D, 6 if D == 2 else 3, 6 if D == 2 else 3, 3 if D == 3 else 0)
    return Vp, F.astype(np.int64), mask


@pytest.mark.parametrize("D", [2, 3])
def test_admm_matrix_is_diagonal_with_explicit_zero_blocks(D):
    """t = M + dt^2 (WD)^T(WD) = diag(tau + dt^2 w^2 valence) with P*D^2 stored entries."""
    Vp, F, _ = small_mesh(D)
    P = Vp.shape[0]
    t = PO.admm_matrix(F, P, D, 0.5, 0.025, 50.0)
    assert t.nnz == P * D * D
    val = np.bincount(F.ravel(), minlength=P)
    d = 0.5 + 0.025 ** 2 * (0.25 * 50.0) * np.repeat(val, D)
    assert np.allclose(t.diagonal(), d, rtol=1e-15)
    off = t - __import__("scipy.sparse", fromlist=["diags"]).diags(t.diagonal())
    assert abs(off).sum() == 0.0


@pytest.mark.skipif(not PO.have_lasolver(), reason="oracle/_ref/liblasolver_ref.so not built (needs /root/reference)")
@pytest.mark.parametrize("D", [2, 3])
def test_lasolver_matches_scipy_and_numpy_pcg(D):
    Vp, F, _ = small_mesh(D)
    P = Vp.shape[0]
    A = PO.laplacian_like_matrix(F, P, D, 0.5, 0.5, 50.0)
    rng = np.random.default_rng(1)
    b = rng.standard_normal(A.shape[0])
    x_ref = spla.spsolve(A.tocsc(), b)
    x_la, it_la = PO.lasolver_solve(A, b, 1e-12)
    assert np.abs(x_la - x_ref).max() <= 1e-10 * np.abs(x_ref).max()
    lu, diag = PO.ilu0(A)
    x_np, it_np = PO.pcg(A, b, 1e-12, 10000, lambda r: PO.ilu_apply(A, lu, diag, r))
    assert np.abs(x_np - x_ref).max() <= 1e-10 * np.abs(x_ref).max()
    assert abs(it_np - it_la) <= 1          # same algorithm, same stopping rule (||r||/||b|| < tol), rounding apart
    # on the ADMM matrix both need one iteration
    t = PO.admm_matrix(F, P, D, 0.5, 0.025, 50.0)
    x1, it1 = PO.lasolver_solve(t, b, 1e-14)
    assert it1 <= 2 and np.allclose(x1, b / t.diagonal(), rtol=1e-13)
