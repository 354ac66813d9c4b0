"""CPU baseline of Method 2 on the same workload as scripts/m2_probe.py: the oracle's backward-Euler step (serial eulerStepMod /
FSubJac restatement, like the reference's serial loops) with the linear solves done by the REFERENCE'S OWN LASolver (oracle/_ref,
ILU(0) + CGSTAB).  Test infrastructure timed as a baseline, like bench.py's cpu_baseline leg.
usage: python scripts/m2_cpu_baseline.py [nx] [steps]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O      # noqa: E402
import pcg_oracle as PO     # noqa: E402

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
Vp, F, mask = O.uniform_rect_mesh(2, nx, nx)
mon = O.monitor_eval(2, 4, Vp)
orc = O.Oracle(Vp, F, mask, mon, 50.0, 0.5, 0.025)
t = time.perf_counter()
rows, cols = orc.jacobian_entries()
S = PO.LASolverJac(orc.P * orc.D, rows, cols) if PO.have_lasolver() else PO.RestatedJac(orc.P * orc.D, rows, cols)
t_struct = time.perf_counter() - t
out = {"workload": f"SquareGrid 2D nx={nx}", "elements": int(F.shape[0]), "solver": type(S).__name__, "cores": 1,
       "structure_s": round(t_struct, 2), "steps": []}
for s in range(steps):
    t = time.perf_counter()
    r = orc.backwards_euler_step(S, 1e-3)
    out["steps"].append({"ms": round((time.perf_counter() - t) * 1e3, 1), "Ih": r["Ih"], "newton_iters": r["newton_iters"],
                         "lin_iters": sum(r["inner_iters"])})
print(json.dumps(out))
