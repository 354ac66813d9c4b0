"""Per-chunk time stamps of the dataflow triangular solve (profiling build, -DMMADMM_TRI_PROF): where a level's time goes.
Stamps per chunk: 0 claimed, 1 factor planes + early operands requested, 2 proxy operand seen, 3 all operands of all lanes seen,
4 results stored.  usage: [MMADMM_SF_PROF_UPPER=1] python scripts/m2_sf_prof.py [nx] [row_block]"""
import ctypes as C, importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
mm.LIB_PATH = os.path.join(ROOT, "build", "prof", "libmmadmm_b200_prof.so")
nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 2
upper = 1 if os.environ.get("MMADMM_SF_PROF_UPPER") else 0
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
pcg.set_option("tri_mode", 2); pcg.set_option("row_block", bs)
r = np.random.default_rng(1).standard_normal(len(rowptr) - 1)
pcg.precond(r); pcg.precond(r)
L = mm.lib()
lv = np.zeros(1 << 16, dtype=np.int32)
L.mmadmm_pcg_debug_levels.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
nlv = L.mmadmm_pcg_debug_levels(pcg.h, upper, lv.ctypes.data_as(C.c_void_p), len(lv))
off = lv[:nlv].astype(np.int64)
widths = np.diff(off)
starts = np.concatenate([[0], np.cumsum((widths + 31) // 32)])        # first chunk of every level
buf = np.zeros(5 * 65536, dtype=np.uint64)
assert L.mmadmm_pcg_debug_sf_prof(buf.ctypes.data_as(C.c_void_p)) == 0
t = buf.reshape(65536, 5).astype(np.int64)
nl = int(np.searchsorted(starts, 65536, side="right")) - 1               # levels fully inside the stamped window
done = np.array([t[starts[l]:starts[l + 1], 4].max() for l in range(nl)])
first = np.array([t[starts[l]:starts[l + 1], 4].min() for l in range(nl)])
print(json.dumps({"solve": "upper" if upper else "lower", "levels": int(nlv - 1), "levels_stamped": nl, "chunks": int(starts[-1]),
                  "width_mean": float(widths[:nl].mean()), "width_max": int(widths[:nl].max())}))
lo, hi = min(100, nl // 4), nl - 2
hop = np.diff(done[lo:hi])
print("level l done - level l-1 done (us): mean %.3f  median %.3f  p10 %.3f  p90 %.3f" % (hop.mean() / 1e3, np.median(hop) / 1e3,
      np.percentile(hop, 10) / 1e3, np.percentile(hop, 90) / 1e3))
spread = (done - first)[lo:hi]
print("spread of a level's chunk completions (us): mean %.3f  p90 %.3f" % (spread.mean() / 1e3, np.percentile(spread, 90) / 1e3))
w = t[starts[lo]:starts[hi]]
print("per chunk, mean us: claim->requested %.2f  requested->proxy %.1f  proxy->all operands %.2f  operands->stored %.2f" % tuple(
    float((w[:, k + 1] - w[:, k]).mean()) / 1e3 for k in range(4)))
# when a chunk of level l has its operands relative to the completion of level l-1 (negative: before the previous level was complete)
rel3 = np.concatenate([t[starts[l]:starts[l + 1], 3] - done[l - 1] for l in range(lo, hi)])
rel2 = np.concatenate([t[starts[l]:starts[l + 1], 2] - done[l - 1] for l in range(lo, hi)])
print("all operands seen - previous level done (us): mean %.3f  p10 %.3f  p90 %.3f" % (rel3.mean() / 1e3, np.percentile(rel3, 10) / 1e3, np.percentile(rel3, 90) / 1e3))
print("proxy seen - previous level done (us): mean %.3f  p10 %.3f  p90 %.3f" % (rel2.mean() / 1e3, np.percentile(rel2, 10) / 1e3, np.percentile(rel2, 90) / 1e3))
sp = np.zeros(65536, dtype=np.int32)
if L.mmadmm_pcg_debug_sf_spins(sp.ctypes.data_as(C.c_void_p)) == 0:
    ws = sp[starts[lo]:starts[hi]]
    dur = (w[:, 3] - w[:, 2]) / 1e3
    ok = ws > 0
    print("per-lane wait of lane 0: polls per chunk mean %.1f  median %.0f;  us per poll %.3f  (chunks with polls: %d of %d)" % (
        ws[ok].mean(), np.median(ws[ok]), float(dur[ok].sum() / ws[ok].sum()), int(ok.sum()), len(ws)))
for l in range(lo + 200, lo + 204):
    seg = t[starts[l]:starts[l + 1]]
    print("level", l, "width", int(widths[l]), "chunk stored times rel. to previous level done (us):", np.round((seg[:, 4] - done[l - 1]) / 1e3, 2).tolist())
