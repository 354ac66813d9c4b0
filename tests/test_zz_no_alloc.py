"""mmadmm_energy_noalloc (include/mmadmm.h): the allocation-free variant of mmadmm_energy, whose partial sums live in a
file-scope `static __device__` array.  No counterpart in the reference: the check is against the existing entry point
(mmadmm_energy) through the C ABI, bit for bit (same kernels, same fixed-order reduction; the tolerance of the unit tests,
1e-12, is the stated bound), on calls in a row that share the array.  Collected last on purpose (pytest -x)."""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
mm = importlib.import_module("mm-admm_b200")

TOL_UNIT = 1e-12


def solver_for(D, nx, kind, rho):
    Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
    mon = mm.TestMonitor(D, kind).evaluate_at_vertices(Vp)
    mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=rho, tau=0.5)
    return mesh, mm.MeshIntegrator(0.025, mesh)


def energy_pair(s, x_dev=None):
    """(mmadmm_energy, mmadmm_energy_noalloc) through the C ABI on the same state."""
    a, b = C.c_double(), C.c_double()
    assert s.L.mmadmm_energy(s.h, C.byref(a)) == 0, s.L.mmadmm_last_error(s.h)
    assert s.L.mmadmm_energy_noalloc(s.h, C.c_void_p(x_dev) if x_dev else None, C.byref(b)) == 0, s.L.mmadmm_last_error(s.h)
    return a.value, b.value


@pytest.mark.parametrize("D,nx,kind,rho", [(2, 48, 3, 100.0), (3, 8, 3, 50.0)])
def test_noalloc_energy_matches_energy_on_consecutive_calls(D, nx, kind, rho):
    mesh, s = solver_for(D, nx, kind, rho)
    seen = []
    for k in range(3):          # calls in a row share the static array: the state moves between them
        a, b = energy_pair(s)
        assert np.isfinite(a) and a > 0
        assert abs(a - b) <= TOL_UNIT * abs(a), (k, a, b)
        assert a == b, (k, a, b)                      # in fact bit-identical
        a2, b2 = energy_pair(s)                       # and again on the unchanged state
        assert (a2, b2) == (a, b)
        seen.append(a)
        s.step(10, 1e-3)
    assert len(set(seen)) == 3, seen                  # the compared values were three different energies
    s.close()


def test_noalloc_energy_with_caller_positions_and_interleaved_contexts():
    import torch
    mesh2, s2 = solver_for(2, 40, 4, 50.0)
    mesh3, s3 = solver_for(3, 6, 3, 50.0)
    s2.step(10, 1e-3)
    x = torch.empty((mesh2.P, 2), dtype=torch.float64, device="cuda:0")
    assert s2.L.mmadmm_get_x_device(s2.h, C.c_void_p(x.data_ptr())) == 0
    torch.cuda.synchronize()
    for _ in range(2):
        # a larger context and a smaller one take turns on the one array: stale partials beyond nblk are never read
        a2, b2 = energy_pair(s2, x.data_ptr())
        a3, b3 = energy_pair(s3)
        assert a2 == b2 and a3 == b3, (a2, b2, a3, b3)
    # positions other than the context's own: compare with a context that holds them
    s2.step(10, 1e-3)
    e_now = s2.getEnergy()
    e_old = s2.getEnergyNoAlloc(x.data_ptr())        # still the positions after the first step
    assert e_old == a2 and e_old != e_now
    s2.close(); s3.close()


def test_noalloc_energy_rejects_bad_arguments():
    mesh, s = solver_for(2, 8, 0, 50.0)
    assert s.L.mmadmm_energy_noalloc(None, None, None) == 1
    assert s.L.mmadmm_energy_noalloc(s.h, None, None) == 1
    s.close()


def test_noalloc_energy_rejects_meshes_beyond_the_static_scratch():
    """131072 partial sums x 128 elements per block = 16,777,216 elements; SquareGrid 2049^2 has 16,793,604."""
    nx = 2049
    Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
    assert F.shape[0] > 131072 * 128
    mon = mm.TestMonitor(2, 0).evaluate_at_vertices(Vp)
    mesh = mm.Mesh(Vp, F, mask, mon, rho=50.0, tau=0.5)
    s = mm.MeshIntegrator(0.025, mesh)
    v = C.c_double()
    assert s.L.mmadmm_energy(s.h, C.byref(v)) == 0
    assert s.L.mmadmm_energy_noalloc(s.h, None, C.byref(v)) == 1      # MMADMM_EINVAL
    msg = s.L.mmadmm_last_error(s.h).decode()
    assert "static scratch" in msg and "16777216" in msg, msg
    s.close()
