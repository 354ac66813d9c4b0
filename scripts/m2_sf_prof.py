"""Per-chunk time stamps of the dataflow triangular solve (profiling build, -DMMADMM_TRI_PROF): where a level's time goes.
usage: python scripts/m2_sf_prof.py [nx] [row_block]"""
import ctypes as C, importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
mm = importlib.import_module("mm-admm_b200")
mm.LIB_PATH = os.path.join(ROOT, "build", "prof", "libmmadmm_b200_prof.so")
nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 2
Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
solver = mm.MeshIntegrator(0.025, mesh)
rowptr, col, val = solver.backward_euler_jacobian()
pcg = mm.PCG(rowptr, col, val, precond="ilu0")
pcg.set_option("tri_mode", 2); pcg.set_option("row_block", bs)
r = np.random.default_rng(1).standard_normal(len(rowptr) - 1)
pcg.precond(r); pcg.precond(r)
buf = np.zeros(5 * 65536, dtype=np.uint64)
L = mm.lib()
assert L.mmadmm_pcg_debug_sf_prof(buf.ctypes.data_as(C.c_void_p)) == 0
t = buf.reshape(65536, 5).astype(np.int64)
# L solve on the (nx+1)^2 lattice, vertices numbered row by row: level l holds the groups with i + j = l (bs = 2: one vertex per group)
per = 2 // bs if bs <= 2 else 1
widths = [min(l // per + 1, nx + 1, 2 * nx + 1 - l // per) for l in range((2 * nx + 1) * per)]
starts = np.concatenate([[0], np.cumsum([(w + 31) // 32 for w in widths])])
t0 = t[:, 0].min()
rows = []
for l in range(600 * per, 606 * per):
    a, b = starts[l], starts[l + 1]
    if b > 65536: break
    seg = t[a:b]
    prev = t[starts[l - 1]:starts[l], 4].max()
    rows.append({"level": l, "chunks": int(b - a), "claim_to_loaded_us": round(float((seg[:, 1] - seg[:, 0]).mean()) / 1e3, 2),
                 "loaded_to_proxy_us": round(float((seg[:, 2] - seg[:, 1]).mean()) / 1e3, 2),
                 "proxy_to_all_us": round(float((seg[:, 3] - seg[:, 2]).mean()) / 1e3, 2),
                 "all_to_stored_us": round(float((seg[:, 4] - seg[:, 3]).mean()) / 1e3, 2),
                 "prev_level_last_store_to_proxy_seen_us": round(float((seg[:, 2] - prev).mean()) / 1e3, 2),
                 "level_done_after_prev_us": round(float(seg[:, 4].max() - prev) / 1e3, 2),
                 "claimed_before_prev_done_us": round(float((prev - seg[:, 0]).mean()) / 1e3, 2)})
for rw in rows: print(json.dumps(rw))
w = t[starts[300 * per]:starts[900 * per]]
print("chunks in levels 300..900: mean us  claim->loaded %.2f  loaded->proxy %.1f  proxy->all operands %.2f  operands->stored %.2f" % tuple(
    float((w[:, k + 1] - w[:, k]).mean()) / 1e3 for k in range(4)))
lv = [t[starts[l]:starts[l + 1], 4].max() for l in range(200 * per, 900 * per) if starts[l + 1] <= 65536]
print("mean us per level over levels 200..900:", round(float(np.diff(lv).mean()) / 1e3, 3))
