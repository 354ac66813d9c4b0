"""CPU-side checks of the drop-in boundary: the CUDA library loads without a GPU, exports every symbol that
include/mmadmm.h declares, and refuses to run without a device (there is no CPU fallback)."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
mm = importlib.import_module("mm-admm_b200")


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "mmadmm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mmadmm_[A-Za-z_0-9]+)\s*\(", src)))


def test_header_symbols_are_exported():
    if not os.path.exists(mm.LIB_PATH):
        pytest.skip("libmmadmm_b200.so not built (run __graft_entry__.build())")
    L = C.CDLL(mm.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 17
    for nm in names:
        assert hasattr(L, nm), f"{nm} declared in include/mmadmm.h but not exported"
    assert sorted(mm.ABI_SYMBOLS) == names, "python mirror and header disagree on the ABI surface"
    assert b"sm_100a" in C.cast(C.CFUNCTYPE(C.c_char_p)(("mmadmm_version", L))(), C.c_char_p).value


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    if not os.path.exists(mm.LIB_PATH):
        pytest.skip("library not built")
    Vp, F, mask = mm.generate_uniform_rect_mesh(2, 4, 4)
    mon = mm.TestMonitor(2, 0).evaluate_at_vertices(Vp)
    mesh = mm.Mesh(Vp, F, mask, mon, rho=50.0, tau=0.5)
    with pytest.raises(mm.MMADMMError) as ei:
        mm.MeshIntegrator(0.025, mesh)
    assert "ENODEVICE" in str(ei.value)


def test_bad_arguments_are_rejected_without_a_device():
    if not os.path.exists(mm.LIB_PATH):
        pytest.skip("library not built")
    L = mm.lib()
    h = C.c_void_p()
    assert L.mmadmm_create(4, 1, 1, None, None, None, None, None, 1.0, 1.0, 1.0, 0, 0, C.byref(h)) == 1  # EINVAL
    assert L.mmadmm_step(None, 1, 1e-3, None, None, None, None) == 1
    assert L.mmadmm_last_error(None) == b"null context"


def test_host_generator_sizes_and_mask():
    Vp, F, mask = mm.generate_uniform_rect_mesh(2, 5, 5)
    assert Vp.shape == (36 + 25, 2) and F.shape == (100, 3)
    assert (mask == mm.BOUNDARY_FIXED).sum() == 20 and (mask == mm.INTERIOR).sum() == 41
    Vp, F, mask = mm.generate_uniform_rect_mesh(3, 2, 2, 2)
    assert Vp.shape == (27 + 8, 3) and F.shape == (96, 4)
    assert F.max() == Vp.shape[0] - 1 and F.min() == 0
