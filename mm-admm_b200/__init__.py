"""mm-admm_b200: B200-native implicit MMPDE step (ADMM) of connortannahill/MM-ADMM.

This module is the Python-side mirror of the reference's host interface for the hot path
(`Mesh<D>` src/Mesh.h:16-127, `MeshIntegrator<D>` src/MeshIntegrator.h:12-50): same constructor
arguments, same method names, same return values.  It is a thin ctypes layer over the C ABI in
include/mmadmm.h (libmmadmm_b200.so, hand-written sm_100a CUDA).  There is NO CPU fallback:
if the CUDA library is missing or no GPU is present, constructing a MeshIntegrator raises.

The directory name contains a hyphen, so import it with
    import importlib; mm = importlib.import_module("mm-admm_b200")
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmmadmm_b200.so")
HOST_LIB_PATH = os.path.join(_HERE, "libmmadmm_host.so")

BOUNDARY_FREE, BOUNDARY_FIXED, INTERIOR = 0, 1, 2  # src/NodeType.h:4-8

_c_dp = C.POINTER(C.c_double)
_c_ip = C.POINTER(C.c_int)
_c_llp = C.POINTER(C.c_longlong)

ERRORS = {1: "EINVAL", 2: "ECUDA", 3: "EINVERTED", 4: "ESINGULAR", 5: "ENCCL", 6: "ENODEVICE", 7: "EPEER"}

# every symbol include/mmadmm.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "mmadmm_create", "mmadmm_set_dt", "mmadmm_step", "mmadmm_energy", "mmadmm_energy_noalloc", "mmadmm_get_x", "mmadmm_get_x_device", "mmadmm_get_x_colmajor", "mmadmm_host_register", "mmadmm_host_unregister", "mmadmm_get_F",
    "mmadmm_euler_step", "mmadmm_backward_euler_step", "mmadmm_backward_euler_jacobian", "mmadmm_grid_dims", "mmadmm_get_state", "mmadmm_eval_monitor", "mmadmm_block_grad", "mmadmm_block_grad_at", "mmadmm_get_stats",
    "mmadmm_get_timing", "mmadmm_set_option", "mmadmm_stream", "mmadmm_destroy", "mmadmm_last_error",
    "mmadmm_version", "mmadmm_plan_create", "mmadmm_plan_sizes", "mmadmm_plan_get", "mmadmm_plan_destroy",
    "mmadmm_comm_unique_id", "mmadmm_create_dist", "mmadmm_local_sizes", "mmadmm_get_x_local",
    "mmadmm_pcg_create", "mmadmm_pcg_solve", "mmadmm_pcg_solve_device", "mmadmm_pcg_bicgstab", "mmadmm_pcg_bicgstab_device", "mmadmm_pcg_spmv", "mmadmm_pcg_precond",
    "mmadmm_update_monitor", "mmadmm_set_writeback", "mmadmm_writeback_fence", "mmadmm_checkpoint", "mmadmm_restore",
    "mmadmm_pcg_get_ilu", "mmadmm_pcg_plan_levels", "mmadmm_pcg_refactor_device", "mmadmm_pcg_set_option", "mmadmm_pcg_info", "mmadmm_pcg_time_spmv", "mmadmm_pcg_time_precond", "mmadmm_pcg_last_error", "mmadmm_pcg_destroy", "mmadmm_measure_fp64_peak",
]


def pcg_plan_levels(rowptr, col, bs):
    """mmadmm_pcg_plan_levels (host only, no GPU needed): (levL, levU, info) of the group schedule that row_block = bs gives
    the ILU(0) triangular solves on this CSR pattern; info = dict(lower_levels, upper_levels, blockable)."""
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int32); col = np.ascontiguousarray(col, dtype=np.int32)
    n = len(rowptr) - 1
    ng = (n + bs - 1) // bs
    levL = np.zeros(ng, dtype=np.int32); levU = np.zeros(ng, dtype=np.int32); info = np.zeros(3, dtype=np.int32)
    rc = lib().mmadmm_pcg_plan_levels(n, _ip(rowptr), _ip(col), int(bs), _ip(levL), _ip(levU), _ip(info))
    if rc:
        raise MMADMMError(f"mmadmm_pcg_plan_levels error {ERRORS.get(rc, rc)}")
    return levL, levU, dict(lower_levels=int(info[0]), upper_levels=int(info[1]), blockable=bool(info[2]))


class MMADMMError(RuntimeError):
    pass


def _dp(a):
    return a.ctypes.data_as(_c_dp) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(_c_ip) if a is not None else None


_lib = None
_host = None


def lib():
    """The CUDA library.  Raises (loudly) when it has not been built: there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MMADMMError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(nvcc, sm_100a). mm-admm_b200 has no CPU fallback.")
        if "MMADMM_NCCL_LIB" not in os.environ:   # multi-GPU set-up: prefer the NCCL that ships with torch (one NCCL per process)
            try:
                import importlib.util
                spec = importlib.util.find_spec("nvidia.nccl")
                for d in (spec.submodule_search_locations if spec else []):
                    cand = os.path.join(d, "lib", "libnccl.so.2")
                    if os.path.exists(cand):
                        os.environ["MMADMM_NCCL_LIB"] = cand
                        break
            except Exception:
                pass
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.mmadmm_create.restype = C.c_int
        L.mmadmm_create.argtypes = [C.c_int, C.c_int, C.c_int, _c_dp, _c_dp, _c_ip, _c_ip, _c_dp, C.c_double, C.c_double,
                                    C.c_double, C.c_int, C.c_int, C.POINTER(vp)]
        L.mmadmm_step.argtypes = [vp, C.c_int, C.c_double, _c_dp, _c_ip, _c_dp, _c_dp]
        L.mmadmm_energy.argtypes = [vp, _c_dp]
        L.mmadmm_energy_noalloc.argtypes = [vp, vp, _c_dp]
        L.mmadmm_set_dt.argtypes = [vp, C.c_double]
        L.mmadmm_euler_step.argtypes = [vp, _c_dp]
        L.mmadmm_backward_euler_step.argtypes = [vp, C.c_double, C.c_double, _c_dp, _c_ip, _c_ip]
        L.mmadmm_backward_euler_jacobian.argtypes = [vp, _c_ip, _c_ip, _c_ip, _c_dp]
        L.mmadmm_pcg_set_option.argtypes = [vp, C.c_char_p, C.c_int]
        L.mmadmm_pcg_bicgstab.argtypes = [vp, _c_dp, _c_dp, C.c_double, C.c_int, _c_ip, _c_dp]
        L.mmadmm_get_x.argtypes = [vp, _c_dp]
        L.mmadmm_get_x_device.argtypes = [vp, vp]
        L.mmadmm_get_F.argtypes = [vp, _c_ip]
        L.mmadmm_get_x_colmajor.argtypes = [vp, _c_dp]
        L.mmadmm_update_monitor.argtypes = [vp, _c_dp]
        L.mmadmm_set_writeback.argtypes = [vp, _c_dp, C.c_int, C.c_int]
        L.mmadmm_writeback_fence.argtypes = [vp]
        L.mmadmm_checkpoint.argtypes = [vp]
        L.mmadmm_restore.argtypes = [vp]
        L.mmadmm_host_register.argtypes = [vp, C.c_ulong]
        L.mmadmm_host_unregister.argtypes = [vp]
        L.mmadmm_grid_dims.argtypes = [vp, _c_ip]
        L.mmadmm_get_state.argtypes = [vp, C.c_int, _c_dp]
        L.mmadmm_eval_monitor.argtypes = [vp, _c_dp, C.c_int, _c_dp]
        L.mmadmm_block_grad.argtypes = [vp, C.c_int, _c_dp, _c_dp, C.c_int, C.c_int, C.c_int, _c_dp, _c_dp, _c_dp]
        L.mmadmm_block_grad_at.argtypes = [vp, C.c_int, C.c_int, _c_dp, _c_dp, C.c_int, C.c_int, C.c_int, _c_dp, _c_dp, _c_dp]
        L.mmadmm_get_stats.argtypes = [vp, _c_llp, _c_llp, _c_llp, _c_llp]
        L.mmadmm_get_timing.argtypes = [vp, _c_dp, _c_llp, C.c_int]
        L.mmadmm_set_option.argtypes = [vp, C.c_char_p, C.c_int]
        L.mmadmm_stream.restype = vp
        L.mmadmm_stream.argtypes = [vp]
        L.mmadmm_destroy.restype = None
        L.mmadmm_destroy.argtypes = [vp]
        L.mmadmm_last_error.restype = C.c_char_p
        L.mmadmm_last_error.argtypes = [vp]
        L.mmadmm_version.restype = C.c_char_p
        L.mmadmm_plan_create.argtypes = [C.c_int, C.c_int, C.c_int, _c_ip, C.c_int, C.c_int, C.POINTER(vp)]
        L.mmadmm_plan_sizes.argtypes = [vp, _c_ip]
        L.mmadmm_plan_get.argtypes = [vp, C.c_int, _c_ip]
        L.mmadmm_plan_destroy.restype = None
        L.mmadmm_plan_destroy.argtypes = [vp]
        L.mmadmm_comm_unique_id.argtypes = [C.c_char_p]
        L.mmadmm_create_dist.argtypes = [C.c_int, C.c_int, C.c_int, _c_dp, _c_dp, _c_ip, _c_ip, _c_dp, C.c_double, C.c_double,
                                         C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(vp)]
        L.mmadmm_local_sizes.argtypes = [vp, _c_ip, _c_ip]
        L.mmadmm_get_x_local.argtypes = [vp, _c_dp]
        L.mmadmm_pcg_create.argtypes = [C.c_int, _c_ip, _c_ip, _c_dp, C.c_int, C.c_int, C.POINTER(vp)]
        L.mmadmm_pcg_solve.argtypes = [vp, _c_dp, _c_dp, C.c_double, C.c_int, _c_ip, _c_dp]
        L.mmadmm_pcg_solve_device.argtypes = [vp, vp, vp, C.c_double, C.c_int, _c_ip, _c_dp]
        L.mmadmm_pcg_spmv.argtypes = [vp, _c_dp, _c_dp]
        L.mmadmm_pcg_precond.argtypes = [vp, _c_dp, _c_dp]
        L.mmadmm_pcg_get_ilu.argtypes = [vp, _c_dp]
        L.mmadmm_pcg_plan_levels.argtypes = [C.c_int, _c_ip, _c_ip, C.c_int, _c_ip, _c_ip, _c_ip]
        L.mmadmm_pcg_refactor_device.argtypes = [vp, vp]
        L.mmadmm_pcg_info.argtypes = [vp, _c_ip]
        L.mmadmm_pcg_time_spmv.argtypes = [vp, C.c_int, _c_dp]
        L.mmadmm_pcg_time_precond.argtypes = [vp, C.c_int, _c_dp]
        L.mmadmm_pcg_last_error.restype = C.c_char_p
        L.mmadmm_pcg_last_error.argtypes = [vp]
        L.mmadmm_measure_fp64_peak.argtypes = [C.c_int, _c_dp]
        L.mmadmm_pcg_destroy.restype = None
        L.mmadmm_pcg_destroy.argtypes = [vp]
        _lib = L
    return _lib


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise MMADMMError(f"{HOST_LIB_PATH} is missing: run __graft_entry__.build()")
        H = C.CDLL(HOST_LIB_PATH)
        H.mmadmm_host_monitor_eval.argtypes = [C.c_int, C.c_int, C.c_int, _c_dp, _c_dp]
        H.mmadmm_host_uniform_rect_mesh.argtypes = [C.c_int] * 4 + [C.c_double] * 6 + [C.c_int, _c_dp, _c_ip, _c_ip]
        H.mmadmm_host_disc_mesh.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, _c_dp, _c_ip, _c_ip]
        _host = H
    return _host


# --------------------------------------------------------------------------------------------------
# host helpers (mesh generators / monitors of the reference driver; init-time only)
# --------------------------------------------------------------------------------------------------
def generate_uniform_rect_mesh(D, nx, ny, nz=0, box=(0, 1, 0, 1, 0, 1), boundary_type=BOUNDARY_FIXED):
    """utils::generateUniformRectMesh<D> (src/MeshUtils.h:82-335) -> (Vp[P,D], F[E,D+1] int32, mask[P] int32)."""
    if D == 2:
        P, E = (nx + 1) * (ny + 1) + nx * ny, 4 * nx * ny
    else:
        P, E = (nx + 1) * (ny + 1) * (nz + 1) + nx * ny * nz, 12 * nx * ny * nz
    Vp = np.zeros((P, D))
    F = np.zeros((E, D + 1), dtype=np.int32)
    mask = np.zeros(P, dtype=np.int32)
    rc = host_lib().mmadmm_host_uniform_rect_mesh(D, nx, ny, nz, *[float(b) for b in box], boundary_type,
                                                  _dp(Vp), _ip(F), _ip(mask))
    if rc:
        raise MMADMMError("bad mesh parameters")
    return Vp, F, mask


def generate_disc_mesh(N, R=1.0, center=(0.0, 0.0), boundary_type=BOUNDARY_FREE):
    """Disc of N hexagonal rings (6 N^2 triangles, 1 + 3N(N+1) points), outer ring marked `boundary_type`; the scalable
    stand-in for the reference's BaseCircle fixtures (BASELINE config 3).  -> (Vp, F, mask)."""
    P, E = 1 + 3 * N * (N + 1), 6 * N * N
    Vp = np.zeros((P, 2)); F = np.zeros((E, 3), dtype=np.int32); mask = np.zeros(P, dtype=np.int32)
    if host_lib().mmadmm_host_disc_mesh(N, float(R), float(center[0]), float(center[1]), boundary_type, _dp(Vp), _ip(F), _ip(mask)):
        raise MMADMMError("bad disc parameters")
    return Vp, F, mask


def partition_plan(D, P, E, F, nranks, rank):
    """Host-side element-block partition plan of `rank` (include/mmadmm.h, mmadmm_plan_*): no device needed."""
    L = lib()
    F = np.ascontiguousarray(F, dtype=np.int32)
    h = C.c_void_p()
    rc = L.mmadmm_plan_create(D, P, E, _ip(F), nranks, rank, C.byref(h))
    if rc:
        raise MMADMMError("bad partition arguments")
    try:
        sz = np.zeros(6, dtype=np.int32)
        L.mmadmm_plan_sizes(h, _ip(sz))
        e0, e1, npl, nsh, nnb, nitems = (int(v) for v in sz)
        lens = {0: npl, 1: (e1 - e0) * (D + 1), 2: npl, 3: nsh, 4: nnb, 5: nnb + 1, 6: nitems, 7: nsh + 1, 9: npl}
        names = {0: "verts", 1: "Floc", 2: "valence", 3: "shared", 4: "nbr_rank", 5: "nbr_ptr", 6: "nbr_list", 7: "contrib_ptr",
                 9: "owner"}
        out = dict(e0=e0, e1=e1)
        for k, n in lens.items():
            a = np.zeros(max(n, 1), dtype=np.int32)
            L.mmadmm_plan_get(h, k, _ip(a))
            out[names[k]] = a[:n]
        nc = int(out["contrib_ptr"][-1]) if nsh else 0
        a = np.zeros(max(nc, 1), dtype=np.int32)
        L.mmadmm_plan_get(h, 8, _ip(a))
        out["contrib"] = a[:nc]
        out["Floc"] = out["Floc"].reshape(-1, D + 1)
        return out
    finally:
        L.mmadmm_plan_destroy(h)


def comm_unique_id():
    """128-byte NCCL unique id (rank 0 creates it; broadcast it with torch.distributed or any other channel)."""
    buf = C.create_string_buffer(128)
    rc = lib().mmadmm_comm_unique_id(buf)
    if rc:
        raise MMADMMError("NCCL is not available")
    return buf.raw


def measure_fp64_peak(device=0):
    """DFMA microbenchmark on `device` -> TFLOP/s (2 flops per FMA)."""
    v = C.c_double()
    rc = lib().mmadmm_measure_fp64_peak(device, C.byref(v))
    if rc:
        raise MMADMMError(f"mmadmm_measure_fp64_peak: {ERRORS.get(rc, rc)}")
    return v.value


class PCG:
    """Preconditioned CG on a fixed CSR matrix (include/mmadmm.h, mmadmm_pcg_*): the general form of the ADMM x-update that
    the reference runs through Eigen::ConjugateGradient (Jacobi) / lib/LASolver (ILU(0) + CG).  precond: "jacobi" | "ilu0"."""

    def __init__(self, rowptr, col, val, precond="jacobi", device=0):
        self.L = lib()
        self.rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        self.col = np.ascontiguousarray(col, dtype=np.int32)
        self.val = np.ascontiguousarray(val, dtype=np.float64)
        self.n = self.rowptr.shape[0] - 1
        self.h = C.c_void_p()
        rc = self.L.mmadmm_pcg_create(self.n, _ip(self.rowptr), _ip(self.col), _dp(self.val), {"jacobi": 0, "ilu0": 1}[precond],
                                      device, C.byref(self.h))
        self._check(rc)

    def _check(self, rc):
        if rc:
            msg = self.L.mmadmm_pcg_last_error(self.h).decode() if self.h else ""
            raise MMADMMError(f"mmadmm_pcg error {ERRORS.get(rc, rc)}: {msg}")

    def solve(self, b, tol=2.220446049250313e-16, maxit=None):
        """-> (x, iterations, ||r||/||b||); defaults are Eigen's: tol = machine epsilon, maxit = 2n."""
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(self.n); it = C.c_int(); rr = C.c_double()
        self._check(self.L.mmadmm_pcg_solve(self.h, _dp(b), _dp(x), float(tol), int(maxit or 2 * self.n), C.byref(it), C.byref(rr)))
        return x, it.value, rr.value

    def bicgstab(self, b, resid_reduc=1e-6, nitmax=10000):
        """-> (x, iterations or -1, ||r||/||b||): LASolver's CGSTAB on the ILU(0) factors, as the reference's Method 2 solves
        (src/Mesh.cpp:266-304, 1311-1323)."""
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(self.n); it = C.c_int(); rr = C.c_double()
        self._check(self.L.mmadmm_pcg_bicgstab(self.h, _dp(b), _dp(x), float(resid_reduc), int(nitmax), C.byref(it), C.byref(rr)))
        return x, it.value, rr.value

    def set_option(self, name, value):
        self._check(self.L.mmadmm_pcg_set_option(self.h, name.encode(), int(value)))

    def spmv(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64); y = np.zeros(self.n)
        self._check(self.L.mmadmm_pcg_spmv(self.h, _dp(x), _dp(y)))
        return y

    def precond(self, r):
        r = np.ascontiguousarray(r, dtype=np.float64); z = np.zeros(self.n)
        self._check(self.L.mmadmm_pcg_precond(self.h, _dp(r), _dp(z)))
        return z

    def ilu(self):
        lu = np.zeros(self.val.shape[0])
        self._check(self.L.mmadmm_pcg_get_ilu(self.h, _dp(lu)))
        return lu

    def refactor_device(self, val_dev):
        """mmadmm_pcg_refactor_device: new values (device address, CSR order) on the same pattern; Jacobi diagonal and ILU(0)
        factor recomputed on the device (LASolver's numeric `factor`)."""
        self._check(self.L.mmadmm_pcg_refactor_device(self.h, C.c_void_p(int(val_dev))))

    def info(self):
        a = np.zeros(4, dtype=np.int32)
        self._check(self.L.mmadmm_pcg_info(self.h, _ip(a)))
        return dict(lanes=int(a[0]), lower_levels=int(a[1]), upper_levels=int(a[2]), launches=int(a[3]))

    def time_precond(self, reps=5):
        ms = C.c_double()
        self._check(self.L.mmadmm_pcg_time_precond(self.h, int(reps), C.byref(ms)))
        return ms.value

    def time_spmv(self, reps=20):
        ms = C.c_double()
        self._check(self.L.mmadmm_pcg_time_spmv(self.h, reps, C.byref(ms)))
        return ms.value

    def close(self):
        if getattr(self, "h", None):
            self.L.mmadmm_pcg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TestMonitor:
    """One of the reference's analytic monitors (Experiments/TestMonitors/MEx*.h) by its main.cpp slot number."""

    def __init__(self, D, kind):
        self.D, self.kind = D, kind

    def evaluate_at_vertices(self, X):
        X = np.ascontiguousarray(X, dtype=np.float64)
        out = np.zeros((X.shape[0], self.D * self.D))
        rc = host_lib().mmadmm_host_monitor_eval(self.D, self.kind, X.shape[0], _dp(X), _dp(out))
        if rc:
            raise MMADMMError(f"no monitor {self.kind} in {self.D}D")
        return out


# --------------------------------------------------------------------------------------------------
# Mesh / MeshIntegrator mirrors
# --------------------------------------------------------------------------------------------------
class Mesh:
    """Mirror of `Mesh<D>` (src/Mesh.h:22-25).  Holds caller-visible `Vp` (updated by the integrator like the
    reference's updateAfterStep does) and `F` (re-oriented in place at integrator construction, like
    reOrientElements).  `monitor` is an object with evaluate_at_vertices(X)->(P,D*D), or the table itself."""

    def __init__(self, Vp, F, boundary_mask, monitor, num_threads=1, rho=1.0, w=0.0, tau=1.0, integration_mode=0,
                 grad_use=False, Vc=None):
        self.Vp = np.ascontiguousarray(Vp, dtype=np.float64)
        self.P, self.D = self.Vp.shape
        if self.D not in (2, 3):
            raise ValueError("D must be 2 or 3")
        self.F = np.ascontiguousarray(F, dtype=np.int32)
        self.E = self.F.shape[0]
        self.boundary_mask = np.ascontiguousarray(boundary_mask, dtype=np.int32)[: self.P].copy()
        self.Vc = None if Vc is None else np.ascontiguousarray(Vc, dtype=np.float64)
        self.compMesh = Vc is not None
        self.rho, self.tau, self.gradUse = float(rho), float(tau), bool(grad_use)
        self.w_json = w  # ignored, as in the reference (src/Mesh.cpp:451 overrides it with 0.5*sqrt(rho))
        self.integrationMode = integration_mode
        if integration_mode not in (0, 1, 2):   # main.cpp:182-191: 0 ADMM step, 1 eulerStep, anything else backwardsEulerStep
            raise ValueError("integration_mode: 0 (ADMM), 1 (explicit Euler) or 2 (backward Euler)")
        mon = monitor.evaluate_at_vertices(self.Vp) if hasattr(monitor, "evaluate_at_vertices") else monitor
        self.mon_vertex = np.ascontiguousarray(mon, dtype=np.float64).reshape(self.P, self.D * self.D)

    def getNPnts(self):
        return self.P


class MeshIntegrator:
    """Mirror of `MeshIntegrator<D>` (src/MeshIntegrator.h:15-27): step / getEnergy / done."""

    def __init__(self, dt, mesh, device=0, sync_vp=True, comm=None):
        """comm = (rank, nranks, id128) makes this rank one member of an element-block-partitioned solve over nranks
        GPUs (one process per GPU; every rank passes the same global mesh).  step/getEnergy/done are then collective.

        sync_vp: what Mesh::updateAfterStep does after every step (src/Mesh.cpp:1030-1035: Vp <- x).  True / "strict": mesh.Vp
        is current when step() returns (the reference's semantics).  "deferred": the copy into mesh.Vp overlaps the next step
        (mmadmm_set_writeback mode 2); mesh.Vp may be read after done(), getEnergy() or fence().  False: no per-step
        write-back (done() still refreshes mesh.Vp)."""
        self.mesh, self.dt = mesh, float(dt)
        self._sync_vp = False
        self.L = lib()
        self.h = C.c_void_p()
        m = mesh
        self.dist = comm is not None and comm[1] > 1
        if self.dist:
            rank, nranks, id128 = comm
            rc = self.L.mmadmm_create_dist(m.D, m.P, m.E, _dp(m.Vp), _dp(m.Vc), _ip(m.F), _ip(m.boundary_mask),
                                           _dp(m.mon_vertex), m.rho, m.tau, self.dt, int(m.gradUse), device, rank, nranks,
                                           id128, C.byref(self.h))
            self._check(rc)
        else:
            rc = self.L.mmadmm_create(m.D, m.P, m.E, _dp(m.Vp), _dp(m.Vc), _ip(m.F), _ip(m.boundary_mask), _dp(m.mon_vertex),
                                      m.rho, m.tau, self.dt, int(m.gradUse), device, C.byref(self.h))
            self._check(rc)
            self._check(self.L.mmadmm_get_F(self.h, _ip(m.F)))  # the reference permutes the caller's F in place
        self.n = m.D * (m.D + 1)
        self.last = {}
        self.sync_vp = sync_vp

    @property
    def sync_vp(self):
        return self._sync_vp

    @sync_vp.setter
    def sync_vp(self, v):
        mode = {False: 0, None: 0, True: 1, "strict": 1, "deferred": 2}[v]
        self._check(self.L.mmadmm_set_writeback(self.h, _dp(self.mesh.Vp) if mode else None, 0, mode))
        self._sync_vp = v

    def update_monitor(self, monitor):
        """Rebuild the monitor grid from the current node positions (what the commented-out Mesh::setUp would do,
        src/Mesh.cpp:1006-1014).  `monitor`: object with evaluate_at_vertices(X), or the table (P, D*D) evaluated at x()."""
        mon = monitor.evaluate_at_vertices(self.x()) if hasattr(monitor, "evaluate_at_vertices") else monitor
        mon = np.ascontiguousarray(mon, dtype=np.float64).reshape(self.mesh.P, self.mesh.D ** 2)
        self._check(self.L.mmadmm_update_monitor(self.h, _dp(mon)))
        self.mesh.mon_vertex = mon

    def fence(self):
        """Wait for a deferred write-back of mesh.Vp to land."""
        self._check(self.L.mmadmm_writeback_fence(self.h))

    def checkpoint(self):
        self._check(self.L.mmadmm_checkpoint(self.h))

    def restore(self):
        self._check(self.L.mmadmm_restore(self.h))

    def _check(self, rc):
        if rc:
            msg = self.L.mmadmm_last_error(self.h).decode() if self.h else ""
            raise MMADMMError(f"mmadmm error {ERRORS.get(rc, rc)}: {msg}")

    def close(self):
        if getattr(self, "h", None):
            self.L.mmadmm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, nIters, tol=1e-3):
        Ih = C.c_double(); it = C.c_int(); pr = C.c_double(); du = C.c_double()
        self._check(self.L.mmadmm_step(self.h, int(nIters), float(tol), C.byref(Ih), C.byref(it), C.byref(pr), C.byref(du)))
        self.last = dict(Ih=Ih.value, iters=it.value, primal=pr.value, dual=du.value)
        # Mesh::updateAfterStep (Vp <- x) happened inside mmadmm_step when sync_vp is on (mmadmm_set_writeback); on a
        # distributed context this rank's vertices only -- done() assembles the global mesh
        return Ih.value

    def eulerStep(self, tol=1e-3):
        """Method 1, MeshIntegrator<D>::eulerStep(tol) (src/MeshIntegrator.cpp:87-95; `tol` is unused there too).  The caller's
        Vp is not refreshed on this path (no updateAfterStep in the reference): use x() or done()."""
        Ih = C.c_double()
        self._check(self.L.mmadmm_euler_step(self.h, C.byref(Ih)))
        return Ih.value

    def backwardsEulerStep(self, dt=None, tol=1e-3):
        """Method 2, MeshIntegrator<D>::backwardsEulerStep(dt, tol) (src/MeshIntegrator.cpp:68-76 -> Mesh::backwardsEulerStep,
        src/Mesh.cpp:1263-1341).  Returns the energy at the last Newton iterate; `last` holds the Newton and BiCGSTAB iteration
        counts.  As in the reference the caller's Vp is not refreshed on this path: use x() or done()."""
        Ih = C.c_double(); nit = C.c_int(); lit = C.c_int()
        self._check(self.L.mmadmm_backward_euler_step(self.h, float(dt if dt is not None else 0.0), float(tol), C.byref(Ih),
                                                      C.byref(nit), C.byref(lit)))
        self.last = dict(Ih=Ih.value, newton_iters=nit.value, lin_iters=lit.value)
        return Ih.value

    def backward_euler_jacobian(self):
        """-> (rowptr, col, val): the Method 2 Jacobian I + (dt/tau) * FD(grad) (Mesh::buildEulerJac, src/Mesh.cpp:1111-1261)."""
        sz = (C.c_int * 2)()
        self._check(self.L.mmadmm_backward_euler_jacobian(self.h, sz, None, None, None))
        rowptr = np.zeros(sz[0] + 1, dtype=np.int32); col = np.zeros(sz[1], dtype=np.int32); val = np.zeros(sz[1])
        self._check(self.L.mmadmm_backward_euler_jacobian(self.h, sz, _ip(rowptr), _ip(col), _dp(val)))
        return rowptr, col, val

    def local_sizes(self):
        p, e = C.c_int(), C.c_int()
        self._check(self.L.mmadmm_local_sizes(self.h, C.byref(p), C.byref(e)))
        return p.value, e.value

    def getEnergy(self):
        if self._sync_vp == "deferred":
            self.fence()
        v = C.c_double()
        self._check(self.L.mmadmm_energy(self.h, C.byref(v)))
        return v.value

    def getEnergyNoAlloc(self, x_dev=None):
        """mmadmm_energy_noalloc: the same energy through the static device scratch (no allocation in the call);
        x_dev = device address (int) of P x D row-major node positions, None = the context's current positions."""
        if self._sync_vp == "deferred":
            self.fence()
        v = C.c_double()
        self._check(self.L.mmadmm_energy_noalloc(self.h, C.c_void_p(x_dev) if x_dev else None, C.byref(v)))
        return v.value

    def done(self):
        self._check(self.L.mmadmm_get_x(self.h, _dp(self.mesh.Vp)))

    # ---- inspection (parity tests) ----
    def x(self):
        out = np.zeros((self.mesh.P, self.mesh.D))
        self._check(self.L.mmadmm_get_x(self.h, _dp(out)))
        return out

    def state(self, which):
        m = self.mesh
        sizes = {0: (m.E, self.n), 1: (m.E, self.n), 2: (m.E, self.n, self.n), 3: (m.P, m.D), 4: (m.P, m.D)}
        if which == 5:
            nx, ny, nz = self.grid_dims()
            shape = ((nx + 1) * (ny + 1) * ((nz + 1) if m.D == 3 else 1), m.D * m.D)
        else:
            shape = sizes[which]
        out = np.zeros(shape)
        self._check(self.L.mmadmm_get_state(self.h, which, _dp(out)))
        return out

    def grid_dims(self):
        d = np.zeros(3, dtype=np.int32)
        self._check(self.L.mmadmm_grid_dims(self.h, _ip(d)))
        return tuple(int(v) for v in d)

    def eval_monitor(self, pts):
        pts = np.ascontiguousarray(pts, dtype=np.float64)
        out = np.zeros((pts.shape[0], self.mesh.D ** 2))
        self._check(self.L.mmadmm_eval_monitor(self.h, _dp(pts), pts.shape[0], _dp(out)))
        return out

    def block_grad(self, zs, dxpu=None, compute_grad=True, regularize=False, apply_mask=False):
        zs = np.ascontiguousarray(zs, dtype=np.float64)
        ne = zs.shape[0]
        dx = None if dxpu is None else np.ascontiguousarray(dxpu, dtype=np.float64)
        en = np.zeros(ne); g = np.zeros((ne, self.n)); igt = np.zeros(ne)
        self._check(self.L.mmadmm_block_grad(self.h, ne, _dp(zs), _dp(dx), int(compute_grad), int(regularize),
                                             int(apply_mask), _dp(en), _dp(g), _dp(igt)))
        return en, g, igt

    def stats(self):
        hist = (C.c_longlong * 52)(); pc = C.c_longlong(); ai = C.c_longlong(); kl = C.c_longlong()
        self._check(self.L.mmadmm_get_stats(self.h, hist, C.byref(pc), C.byref(ai), C.byref(kl)))
        return dict(bfgs_hist=np.array(list(hist), dtype=np.int64), prox_calls=pc.value, admm_iters=ai.value,
                    kernel_launches=kl.value)

    def set_option(self, name, value):
        self._check(self.L.mmadmm_set_option(self.h, name.encode(), int(value)))

    def set_dt(self, dt):
        """Change the step size (the reference's driver assigns MeshIntegrator::dt between steps)."""
        self._check(self.L.mmadmm_set_dt(self.h, float(dt)))
        self.dt = float(dt)

    def p2p_active(self):
        """True when a distributed context exchanges halo sums / scalars by NVLink peer stores (CUDA IPC) instead of NCCL."""
        return self.L.mmadmm_set_option(self.h, b"p2p_active", 0) == 0

    def iso_grid_active(self):
        """True when the monitor grid is isotropic (m*I everywhere) and lookups use the scalar grid."""
        return self.L.mmadmm_set_option(self.h, b"iso_grid_active", 0) == 0

    def timing(self, reset=False):
        ms = (C.c_double * 5)(); cnt = (C.c_longlong * 5)()
        self._check(self.L.mmadmm_get_timing(self.h, ms, cnt, int(reset)))
        names = ["predict", "prox", "xupdate", "residual", "hess_init"]
        return {k: (ms[i], cnt[i]) for i, k in enumerate(names)}

    def stream(self):
        return self.L.mmadmm_stream(self.h)
