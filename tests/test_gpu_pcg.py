"""GPU parity of the general sparse x-update path (mm-admm_b200/csrc/pcg.cu, through the C ABI) against the reference's own
LASolver (oracle/_ref, when built), the numpy restatements in tests/pcg_oracle.py and scipy.
Tolerances: SpMV / preconditioner application 1e-13 relative (summation order differs), ILU(0) factor values 1e-14,
solutions 1e-9 relative (north star), iteration counts equal to the numpy PCG +-1."""
import importlib

import numpy as np
import pytest
import scipy.sparse.linalg as spla

import oracle_lib as O
import pcg_oracle as PO

pytestmark = pytest.mark.gpu
mm = importlib.import_module("mm-admm_b200")


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def mesh(D, nx):
    Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
    return Vp, F.astype(np.int64), mask


@pytest.mark.parametrize("D,nx", [(2, 12), (3, 4)])
def test_admm_matrix_jacobi_pcg_equals_the_diagonal_solve(D, nx):
    """The reference's x-update: Eigen CG + Jacobi on t (src/MeshIntegrator.cpp:54-55,138).  One iteration; the result is
    the diagonal solve the fused k_xupdate kernel performs (checked here against x = b / diag(t))."""
    Vp, F, _ = mesh(D, nx)
    P = Vp.shape[0]
    t = PO.admm_matrix(F, P, D, 0.5, 0.025, 50.0)
    assert t.nnz == P * D * D
    s = mm.PCG(t.indptr, t.indices, t.data, "jacobi")
    rng = np.random.default_rng(0)
    b = rng.standard_normal(P * D)
    x, it, rr = s.solve(b)           # Eigen defaults: tol = eps, maxit = 2n
    assert it <= 2 and rr <= 1e-15
    assert rel(x, b / t.diagonal()) < 1e-14
    assert rel(s.spmv(b), t @ b) < 1e-15
    s.close()


@pytest.mark.parametrize("D,nx,precond", [(2, 16, "ilu0"), (2, 16, "jacobi"), (3, 5, "ilu0")])
def test_pcg_on_coupled_spd_matrix(D, nx, precond):
    Vp, F, _ = mesh(D, nx)
    P = Vp.shape[0]
    A = PO.laplacian_like_matrix(F, P, D, 0.5, 0.5, 50.0)
    n = A.shape[0]
    rng = np.random.default_rng(2)
    b = rng.standard_normal(n)
    s = mm.PCG(A.indptr, A.indices, A.data, precond)
    info = s.info()
    assert info["lanes"] >= 2                       # sub-warp-per-row SpMV on this pattern (7..15 entries per row)
    # SpMV
    xs = rng.standard_normal(n)
    assert rel(s.spmv(xs), A @ xs) < 1e-13
    # preconditioner
    r = rng.standard_normal(n)
    if precond == "ilu0":
        lu, diag = PO.ilu0(A)
        assert rel(s.ilu(), lu) < 1e-14
        assert rel(s.precond(r), PO.ilu_apply(A, lu, diag, r)) < 1e-13
        assert info["lower_levels"] > 1 and info["upper_levels"] > 1
        prec = lambda v: PO.ilu_apply(A, lu, diag, v)
    else:
        assert rel(s.precond(r), r / A.diagonal()) < 1e-15
        prec = lambda v: v / A.diagonal()
    # solve
    tol = 1e-12
    x, it, rr = s.solve(b, tol=tol, maxit=10000)
    x_ref = spla.spsolve(A.tocsc(), b)
    assert rr <= tol and rel(x, x_ref) < 1e-9
    x_np, it_np = PO.pcg(A, b, tol, 10000, prec)
    assert abs(it - it_np) <= 1
    if precond == "ilu0" and PO.have_lasolver():    # the reference's own solver: ILU level 0 + CG accelerator
        x_la, it_la = PO.lasolver_solve(A, b, tol)
        assert rel(x, x_la) < 1e-9
        assert abs(it - it_la) <= 1
    s.close()


@pytest.mark.parametrize("D,nx", [(2, 16), (3, 5)])
def test_device_ilu0_factor_and_refactor_are_bit_identical_to_the_host_factor(D, nx):
    """The ILU(0) factor is computed on the device level by level (k_ilu_level) in LASolver's IKJ order with separately rounded
    multiply and subtract: the same bits as the numpy restatement (pcg_oracle.ilu0, itself within 1e-14 of the reference's
    factor).  mmadmm_pcg_refactor_device = LASolver's numeric `factor` on an unchanged pattern: new values from a device array,
    everything that holds values (Jacobi diagonal, factor, SELL copy) follows."""
    import torch
    Vp, F, _ = mesh(D, nx)
    A = PO.laplacian_like_matrix(F, Vp.shape[0], D, 0.5, 0.5, 50.0)
    A.sort_indices()
    n = A.shape[0]
    rng = np.random.default_rng(7)
    s = mm.PCG(A.indptr, A.indices, A.data, "ilu0")
    lu, diag = PO.ilu0(A)
    assert np.array_equal(s.ilu(), lu)
    A2 = A.copy()
    A2.data = A.data * (1.0 + 0.05 * rng.standard_normal(A.nnz))      # no longer symmetric: ILU(0) does not care
    A2.data[diag] = np.abs(A.data[diag]) * 1.2
    v = torch.from_numpy(A2.data.copy()).to("cuda:0")
    torch.cuda.synchronize()
    s.refactor_device(v.data_ptr())
    lu2, _ = PO.ilu0(A2)
    assert np.array_equal(s.ilu(), lu2)
    s.set_option("ilu_thread_kernel", 1)                                # the thread-per-row factor kernel (rows longer than 64 entries
    s.refactor_device(v.data_ptr())                                    # take it by themselves): same bits as the warp-per-row one
    assert np.array_equal(s.ilu(), lu2)
    s.set_option("ilu_thread_kernel", 0)
    fresh = mm.PCG(A2.indptr, A2.indices, A2.data, "ilu0")
    r = rng.standard_normal(n)
    assert np.array_equal(s.precond(r), fresh.precond(r))
    assert np.array_equal(s.spmv(r), fresh.spmv(r))
    xa, ita, _ = s.bicgstab(r, 1e-8, 2000); xb, itb, _ = fresh.bicgstab(r, 1e-8, 2000)
    assert ita == itb > 0 and np.array_equal(xa, xb)
    zf = fresh.precond(r)
    for bs in (2, 3, 4, 6):                                            # block rows keep serving the refreshed factor
        s.set_option("row_block", bs)
        assert np.array_equal(s.precond(-r), -zf)                      # (a different right-hand side first: the result buffer
        assert np.array_equal(s.precond(r), zf)                        # must be rewritten, not left over from the last call)
    s.set_option("tri_layout", 1)                                      # ... and so does the level-ELL copy of the factor
    for bs in (1, 2, 3):
        s.set_option("row_block", bs)
        assert np.array_equal(s.precond(-r), -zf)
        assert np.array_equal(s.precond(r), zf)
    s.set_option("tri_mode", 2)                                        # ... and the chunk-major copies of the dataflow solve
    for bs, blocks in ((1, 1), (2, 1), (D, 1), (D, 0), (4, 1)):        # (generic form and, where bs = D fits the pattern, block form)
        s.set_option("tri_blocks", blocks)
        s.set_option("row_block", bs)
        assert np.array_equal(s.precond(-r), -zf)
        assert np.array_equal(s.precond(r), zf)
    s.set_option("tri_blocks", 1)
    s.set_option("row_block", D)                                       # stays on for the re-factorisation below
    z0 = mm.PCG(A.indptr, A.indices, A.data, "ilu0")
    zA = z0.precond(r); z0.close()
    s.set_option("spmv_format", 1)                                     # the SELL copy is rebuilt from refreshed values
    assert rel(s.spmv(r), A2 @ r) < 1e-13
    v.copy_(torch.from_numpy(A.data))
    torch.cuda.synchronize()
    s.refactor_device(v.data_ptr())
    assert rel(s.spmv(r), A @ r) < 1e-13 and np.array_equal(s.ilu(), lu)
    assert np.array_equal(s.precond(r), zA)                            # level-ELL copy refreshed by the re-factorisation
    bad = A.data.copy(); bad[diag[0]] = 0.0
    v.copy_(torch.from_numpy(bad)); torch.cuda.synchronize()
    with pytest.raises(mm.MMADMMError) as ei:
        s.refactor_device(v.data_ptr())
    assert "ESINGULAR" in str(ei.value) or "zero pivot" in str(ei.value)
    s.close(); fresh.close()


def test_pcg_rejects_bad_input_and_reports_breakdown():
    indptr = np.array([0, 1, 2], dtype=np.int32); idx = np.array([1, 0], dtype=np.int32); val = np.array([1.0, 1.0])
    with pytest.raises(mm.MMADMMError):
        mm.PCG(indptr, idx, val, "jacobi")          # no stored diagonal


def test_spmv_bandwidth_on_c2_sized_admm_matrix():
    """SpMV on the C2-sized reference matrix (4.2 M rows, 8.4 M stored entries): algorithmic bytes 12 nnz + 20 rows
    (SURVEY 8d); prints the achieved GB/s (bench-style number for profiles/, no threshold beyond sanity)."""
    nx = 1024
    P = (nx + 1) ** 2 + nx * nx
    D = 2
    n = P * D
    indptr = np.arange(0, n * D + 1, D, dtype=np.int32)
    indices = (np.repeat(np.arange(P) * D, D * D).reshape(P, D, D) + np.arange(D)[None, None, :]).reshape(-1).astype(np.int32)
    data = np.zeros((P, D, D)); data[:, 0, 0] = 0.6; data[:, 1, 1] = 0.6
    s = mm.PCG(indptr, indices, data.reshape(-1), "jacobi")
    ms = s.time_spmv(50)
    gbs = (12 * n * D + 20 * n) / ms / 1e6
    print(f"\nSpMV+dot C2-sized ADMM matrix: {ms * 1e3:.1f} us, {gbs:.0f} GB/s algorithmic")
    assert gbs > 500
    s.close()


@pytest.mark.parametrize("D,nx", [(2, 16), (3, 5)])
def test_sell_spmv_and_solves_match_csr(D, nx):
    """SELL-32-4096 SpMV (option "spmv_format" = 1): same product as the CSR kernel and scipy to 1e-13, and the
    solvers run on it unchanged: PCG and LASolver-style BiCGSTAB give the CSR path's solution (1e-12) and iteration counts."""
    Vp, F, _ = mesh(D, nx)
    P = Vp.shape[0]
    A = PO.laplacian_like_matrix(F, P, D, 0.5, 0.5, 50.0)
    n = A.shape[0]
    rng = np.random.default_rng(4)
    xs, b = rng.standard_normal(n), rng.standard_normal(n)
    s = mm.PCG(A.indptr, A.indices, A.data, "ilu0")
    y_csr = s.spmv(xs)
    x_csr, it_csr, _ = s.solve(b, tol=1e-12, maxit=10000)
    xb_csr, itb_csr, _ = s.bicgstab(b, 1e-10, 10000)
    s.set_option("spmv_format", 1)
    pad = s.L.mmadmm_pcg_set_option(s.h, b"sell_padding_permille", 0)
    assert 0 <= pad < 400
    y_sell = s.spmv(xs)
    assert rel(y_sell, A @ xs) < 1e-13 and rel(y_sell, y_csr) < 1e-13
    x_sell, it_sell, rr = s.solve(b, tol=1e-12, maxit=10000)
    assert rr <= 1e-12 and abs(it_sell - it_csr) <= 1 and rel(x_sell, x_csr) < 1e-10
    xb_sell, itb_sell, _ = s.bicgstab(b, 1e-10, 10000)
    assert abs(itb_sell - itb_csr) <= 1 and rel(xb_sell, xb_csr) < 1e-8
    s.set_option("spmv_format", 0)
    assert np.array_equal(s.spmv(xs), y_csr)
    s.close()


def test_spmv_bandwidth_csr_vs_sell_on_c2_sized_method2_jacobian_pattern():
    """The matrix the general path really meets: the Method 2 Jacobian pattern of a SquareGrid (block rows, 10..18 entries per
    row) at 512^2 cells (2.1 M rows, 29 M entries).  Prints both formats' SpMV+dot times; algorithmic bytes 12 nnz + 20 rows."""
    nx = 512
    Vp, F, _ = mesh(2, nx)
    P = Vp.shape[0]
    A = PO.laplacian_like_matrix(F, P, 2, 0.5, 0.5, 50.0)
    n, nnz = A.shape[0], A.nnz
    s = mm.PCG(A.indptr, A.indices, A.data, "jacobi")
    t_csr = s.time_spmv(30)
    s.set_option("spmv_format", 1)
    pad = s.L.mmadmm_pcg_set_option(s.h, b"sell_padding_permille", 0)
    t_sell = s.time_spmv(30)
    by = 12 * nnz + 20 * n
    print(f"\nSpMV+dot {n} rows {nnz} nnz: CSR ({s.info()['lanes']} lanes/row) {t_csr * 1e3:.1f} us = {by / t_csr / 1e6:.0f} GB/s; "
          f"SELL-32-4096 {t_sell * 1e3:.1f} us = {by / t_sell / 1e6:.0f} GB/s algorithmic (padding {pad / 10:.1f} %)")
    assert by / t_sell / 1e6 > 500
    s.close()
