#!/usr/bin/env python
"""bench.py -- ADMM steps/sec & element-prox updates/sec of the implicit MMPDE step on B200.

A "step" is one MeshIntegrator::step (reference src/MeshIntegrator.cpp:101-191) on the workload named in
`config.workload`.  At N=1 that is BASELINE.json configs[1]: the 2D unit square, SquareGrid nx=ny=1024
(4,194,304 triangles -- the generator emits 4 triangles per cell), shock-layer monitor MEx4, rho=50, tau=0.5,
dt=0.025, AdmmIter=10, tol=1e-3.  Warm-up steps 0..W-1 contain the reference's start-up work (finite-difference
Hessians, gradient predictor); the timed steps are steady state (extrapolation predictor).

  value  : element-prox updates/s (= E x ADMM iterations executed / time), state resident in HBM
  e2e    : same metric through the reference-facing host API (MeshIntegrator.step with the per-step write-back of the caller's
           host Vp that Mesh::updateAfterStep does: D2H of P*D doubles into pinned memory, overlapped with the next step and
           fenced INSIDE the timed region; `e2e.strict` is the same with the copy fenced after every step)
  roofline: k_prox (the dominant kernel) against the measured HBM copy bandwidth
  workloads: the same record for BASELINE configs[2] (c3: disc with BOUNDARY_FREE rim) and configs[3] (c4: 3DMonitor280)
  parity : (N > 1) a mid-size 2D and 3D case solved on all N ranks and by the CPU oracle on rank 0 before anything is timed;
           the run fails above 1e-9
  cpu_baseline / --impl reference: the CPU oracle port (oracle/, OpenMP over elements like the reference's -D THREADS build)
           on THE SAME mesh and step window as the GPU arm, on this box's host cores.
"""
import argparse
import importlib
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "ADMM steps/sec & element-prox updates/sec"
UNIT = "element-prox updates/s"
PARAMS = dict(admm_iter=10, tol=1e-3)
WEAK_NX_2D = {1: 1024, 2: 1448, 4: 2048, 8: 2896}   # SURVEY 8(d) C5 weak-scaling ladder
WEAK_N_DISC = {1: 816, 2: 1154, 4: 1632, 8: 2308}   # C3: 6 N^2 triangles = 4.0 / 8.0 / 16.0 / 32.0 M
WEAK_NX_3D = {1: 80, 2: 101, 4: 127, 8: 160}        # C4: 12 nx^3 tets = 6.1 / 12.4 / 24.6 / 49.2 M
PARITY_TOL = 1e-9


def workload(n_gpus, name, nx=0):
    """BASELINE.json configs: c2 = configs[1] (the N=1 bench workload; N>1: the C5 weak-scaling ladder), c3 = configs[2]
    (disc with BOUNDARY_FREE rim), c4 = configs[3] (3DMonitor280)."""
    if name == "c4":
        nx = nx or WEAK_NX_3D.get(n_gpus, 80)
        return dict(name=name, D=3, nx=nx, kind="box", mon=3, btype=1, rho=50.0, tau=0.5, dt=0.025, safe=10**9,
                    label=f"SquareGrid 3D nx=ny=nz={nx} ({12 * nx ** 3:,} tets), MEx33D, rho=50 tau=0.5 dt=0.025 AdmmIter=10")
    if name == "c3":
        N = nx or WEAK_N_DISC.get(n_gpus, 816)
        return dict(name=name, D=2, nx=N, kind="disc", mon=5, btype=0, rho=5.0, tau=0.1, dt=0.05, safe=10**9,
                    label=f"unit disc of {N} hexagonal rings ({6 * N * N:,} triangles), rim nodes BOUNDARY_FREE (reference semantics: "
                          f"gradient live, no projection), MEx5 spiral monitor, rho=5 tau=0.1 dt=0.05 AdmmIter=10 (Monitor310's parameters)")
    nx = nx or WEAK_NX_2D.get(n_gpus, 1024)
    # The reference algorithm breaks down on the shock-layer workload after a dozen outer steps (0/0 in the BFGS update once
    # an element's step underflows, src/Mesh.cpp:846-848; reproduced identically by the oracle and by this library:
    # tests/test_gpu_parity_large.py::test_reference_breakdown_is_reproduced).  A timed region therefore never runs past
    # outer step 8: the K timed steps are taken in epochs of <= 9 - W steps, every epoch restarting from the same
    # post-warm-up checkpoint (mmadmm_checkpoint / mmadmm_restore; the oracle arm: orc_clone).
    return dict(name=name, D=2, nx=nx, kind="box", mon=4, btype=1, rho=50.0, tau=0.5, dt=0.025, safe=9,
                label=f"SquareGrid 2D nx=ny={nx} ({4 * nx * nx:,} triangles), MEx4 shock-layer monitor, "
                      f"rho=50 tau=0.5 dt=0.025 AdmmIter=10 tol=1e-3")


def make_mesh(wl, nx=None):
    """(Vp, F, mask, monitor table) of a workload, from the package's host-side generators."""
    mm = importlib.import_module("mm-admm_b200")
    D, nx = wl["D"], nx or wl["nx"]
    if wl["kind"] == "disc":
        Vp, F, mask = mm.generate_disc_mesh(nx, 1.0, (0.0, 0.0), wl["btype"])
    else:
        Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0, boundary_type=wl["btype"])
    return Vp, F, mask, mm.TestMonitor(D, wl["mon"]).evaluate_at_vertices(Vp)


def algorithmic_bytes_prox(D, P, E, g):
    """B_prox(D) = 8(2n^2 + 3n) + 4(D+1) per element (z in/out, DXpU in, Bkinv in/out, F)
    plus the compulsory monitor-grid read 8 D^2 (g+1)^D / E."""
    n = D * (D + 1)
    return 8 * (2 * n * n + 3 * n) + 4 * (D + 1) + 8 * D * D * (g + 1) ** D / E


def ncu_traffic(workload_name):
    """dram__bytes_read.sum + dram__bytes_write.sum of the prox kernel, per launch, from the committed `ncu --set full`
    capture of this workload (profiles/*_prof_<workload>_raw.json, newest round), or None."""
    import glob
    units = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", f"*_prof_{workload_name}_raw.json")), reverse=True):
        try:
            for m in json.load(open(f)):
                if "k_prox" in m.get("kernel", ""):
                    rd, wr = m["dram__bytes_read.sum"], m["dram__bytes_write.sum"]
                    return float(rd[0]) * units[rd[1]] + float(wr[0]) * units[wr[1]], os.path.relpath(f, ROOT)
        except Exception:
            pass
    return None, None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons sampled through NVML every ~2 ms DURING the timed region (the region is a few
    milliseconds long, far below nvidia-smi's polling period), in a background thread."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index=0, period_s=0.002):
        import threading
        self.period_s = period_s
        self.sm, self.mask, self.max_mhz, self.stop_flag = [], 0, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
            return
        self._sample()
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _sample(self):
        try:
            self.sm.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            self.mask |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        except Exception:
            pass

    def _run(self):
        while not self.stop_flag:
            self._sample()
            time.sleep(self.period_s)

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=self.max_mhz, reasons=[], samples=0)
        if self.nv is None:
            return out
        self.stop_flag = True
        self.t.join(timeout=2)
        self._sample()
        if self.sm:
            out["sm_mhz"] = statistics.median(self.sm)
        out["reasons"] = sorted(v for k, v in self.REASONS.items() if self.mask & k)
        out["samples"] = len(self.sm)
        return out


# ------------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the same mesh, the same warm-up and the same epochs as the GPU arm
# ------------------------------------------------------------------------------------------------------------------------
def run_oracle(wl, steps, warmup, threads, one_thread_steps=0):
    """-> dict(value, ms_per_step, sample, elements, iters, kbar, one_thread).  Every epoch of <= wl['safe'] - warmup steady-state
    steps restarts from a deep copy of the post-warm-up state (orc_clone), as the GPU arm restores its checkpoint."""
    import numpy as np
    import oracle_lib as O
    Vp, F, mask, mon = make_mesh(wl)
    t0 = time.perf_counter()
    base = O.Oracle(Vp, F, mask, mon, wl["rho"], wl["tau"], wl["dt"], nthreads=threads)
    init_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in range(warmup):
        r = base.step(PARAMS["admm_iter"], PARAMS["tol"])
        if r["status"]:
            raise SystemExit("oracle reported an inverted element during the warm-up")
    warm_s = time.perf_counter() - t0
    safe = max(1, wl["safe"] - warmup)
    h0, c0 = base.stats()
    iters, dt, done, epochs, hist, calls = 0, 0.0, 0, 0, 0, 0
    one = None
    while done < steps:
        kk = min(safe, steps - done)
        orc = base.clone() if (done + kk < steps or one_thread_steps) else base
        t0 = time.perf_counter()
        for _ in range(kk):
            r = orc.step(PARAMS["admm_iter"], PARAMS["tol"])
            if r["status"]:
                raise SystemExit("oracle reported an inverted element inside the timed region")
            iters += r["iters"]
        dt += time.perf_counter() - t0
        h1, c1 = orc.stats()
        hist, calls = hist + (h1 - h0), calls + (c1 - c0)
        done += kk
        epochs += 1
        del orc
    E = F.shape[0]
    if one_thread_steps:   # the same loop on ONE host thread (a bounded sample: it runs at ~1 M element-prox/s)
        orc = base.clone()
        orc.set_threads(1)
        it1 = 0
        t0 = time.perf_counter()
        for _ in range(min(one_thread_steps, safe)):
            it1 += orc.step(PARAMS["admm_iter"], PARAMS["tol"])["iters"]
        d1 = time.perf_counter() - t0
        one = dict(value=E * it1 / d1, unit=UNIT, cores=1, ms_per_step=d1 / min(one_thread_steps, safe) * 1e3,
                   sample=f"{min(one_thread_steps, safe)} steady-state outer steps from the same post-warm-up state")
        del orc
    shape = f"disc of {wl['nx']} rings" if wl["kind"] == "disc" else f"SquareGrid {wl['D']}D nx={wl['nx']}"
    kbar = float((hist * np.arange(52)).sum() / max(hist.sum(), 1))
    desc = (f"oracle port (C++/OpenMP prox loop, {threads} threads) on {shape} ({E:,} elements) = the GPU arm's mesh, monitor and "
            f"parameters; {steps} steady-state outer steps {warmup}..{warmup + min(safe, steps) - 1} in {epochs} epoch(s) from one post-warm-up "
            f"state ({iters} ADMM iterations, {kbar:.3f} BFGS iterations per element-prox); init {init_s:.1f} s + {warmup} warm-up steps {warm_s:.1f} s untimed")
    return dict(value=E * iters / dt, ms_per_step=dt / steps * 1e3, sample=desc, elements=E, points=Vp.shape[0], iters=iters,
                kbar=kbar, one_thread=one)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wl = workload(args.gpus, args.workload, args.nx)
    threads = os.cpu_count() or 1
    W = min(max(args.warmup, 3), 6)
    r = run_oracle(wl, args.steps, W, threads, one_thread_steps=2)
    v, ms = r["value"], r["ms_per_step"]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl["label"], "elements": r["elements"], "points": r["points"]},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": r["sample"],
                             "mean_bfgs_iters_per_element_prox": r["kbar"], "admm_iters": r["iters"], "one_thread": r["one_thread"]},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "steps_per_sec": 1e3 / ms, "admm_iters": r["iters"], "mean_bfgs_iters_per_element_prox": r["kbar"], "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------------------
class Env:
    """torch / torch.distributed plumbing: device, barrier, max-over-ranks, NCCL id broadcast.  Nothing here computes."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: mm-admm_b200 has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:   # one process per GPU (torchrun)
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.mm = importlib.import_module("mm-admm_b200")

    def new_comm(self):
        if self.world == 1:
            return None
        torch = self.torch
        t = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if self.rank == 0:
            t.copy_(torch.frombuffer(bytearray(self.mm.comm_unique_id()), dtype=torch.uint8))
        self.dist.broadcast(t, 0)
        return (self.rank, self.world, bytes(t.cpu().numpy().tobytes()))

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if not self.dist:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, v):
        if not self.dist:
            return v
        t = self.torch.tensor([v], dtype=self.torch.int64, device="cuda")
        self.dist.all_reduce(t)
        return int(t.item())

    def bcast_flag(self, ok):
        if not self.dist:
            return ok
        t = self.torch.tensor([1 if ok else 0], dtype=self.torch.int64, device="cuda")
        self.dist.broadcast(t, 0)
        return bool(t.item())


def multi_gpu_parity(env):
    """N > 1: a 2D and a 3D mid-size case on all N ranks against the CPU oracle (rank 0), 6 outer steps each (finite-difference
    Hessians, gradient predictor, several ADMM iterations, then steady state).  -> record; raises SystemExit above 1e-9."""
    import numpy as np
    mm = env.mm
    rec = {"tolerance": PARITY_TOL, "cases": []}
    worst_dx = worst_ih = 0.0
    iters_equal = True
    for D, nx, kind in ((2, 192, 3), (3, 16, 3)):
        Vp, F, mask = mm.generate_uniform_rect_mesh(D, nx, nx, nx if D == 3 else 0)
        mon = mm.TestMonitor(D, kind).evaluate_at_vertices(Vp)
        mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5)
        s = mm.MeshIntegrator(0.025, mesh, device=env.local, comm=env.new_comm())
        ihs, its = [], []
        for _ in range(6):
            ihs.append(s.step(10, 1e-3))
            its.append(s.last["iters"])
        e1 = s.getEnergy()
        s.done()
        p2p = s.p2p_active()
        s.close()
        if env.rank == 0:
            import oracle_lib as O
            orc = O.Oracle(Vp, F, mask, mon, 50.0, 0.5, 0.025, nthreads=os.cpu_count() or 1)
            ro = [orc.step(10, 1e-3) for _ in range(6)]
            xo = orc.x()
            dx = float(np.abs(mesh.Vp - xo).max() / np.abs(xo).max())
            dih = max(abs(a - b["Ih"]) / abs(b["Ih"]) for a, b in zip(ihs, ro))
            dih = max(dih, abs(e1 - orc.energy()) / abs(orc.energy()))
            eq = its == [b["iters"] for b in ro] and all(b["status"] == 0 for b in ro)
            worst_dx, worst_ih, iters_equal = max(worst_dx, dx), max(worst_ih, dih), iters_equal and eq
            rec["cases"].append(dict(case=f"SquareGrid {D}D nx={nx} MEx{kind} ({F.shape[0]:,} elements), 6 outer steps, {env.world} ranks vs CPU oracle",
                                     max_rel_dx=dx, max_rel_Ih=dih, admm_iters=its, iters_equal=eq, p2p=p2p))
    rec.update(max_rel_dx=worst_dx, max_rel_Ih=worst_ih, iters_equal=iters_equal)
    ok = (worst_dx <= PARITY_TOL and worst_ih <= PARITY_TOL and iters_equal) if env.rank == 0 else True
    rec["ok"] = ok
    if not env.bcast_flag(ok):
        if env.rank == 0:
            print(json.dumps({"error": "multi-GPU parity against the CPU oracle failed", "parity": rec}))
        raise SystemExit(3)
    return rec


def run_workload(env, args, wl, K, W, cpu_baseline):
    """One workload on env.world GPUs -> the bench record (rank 0 fills the reporting fields)."""
    import numpy as np
    torch, mm, dist = env.torch, env.mm, env.dist
    D = wl["D"]
    Vp, F, mask, mon = make_mesh(wl)
    P, E = Vp.shape[0], F.shape[0]
    # caller-owned Vp in pinned host memory (what the reference's Mesh writes back every step)
    vp_pinned = torch.empty((P, D), dtype=torch.float64).pin_memory()
    vp_pinned.numpy()[:] = Vp
    mesh = mm.Mesh(vp_pinned.numpy(), F.copy(), mask, mon, rho=wl["rho"], tau=wl["tau"])
    mesh.Vp = vp_pinned.numpy()
    t0 = time.perf_counter()
    solver = mm.MeshIntegrator(wl["dt"], mesh, device=env.local, sync_vp=False, comm=env.new_comm())
    init_s = time.perf_counter() - t0
    if args.prox_variant >= 0:
        solver.set_option("prox_variant", args.prox_variant)
    if not args.iso_grid:
        solver.set_option("iso_grid", 0)
    for kv in args.opt:
        solver.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    e0 = solver.getEnergy()
    warm = []
    for _ in range(W):   # untimed warm-up: outer steps 0..W-1
        t1 = time.perf_counter()
        solver.step(PARAMS["admm_iter"], PARAMS["tol"])
        warm.append(dict(ms=(time.perf_counter() - t1) * 1e3, **solver.last))
    solver.checkpoint()
    SAFE = max(1, wl["safe"] - W)
    stream = torch.cuda.ExternalStream(solver.stream())
    nit, tol = PARAMS["admm_iter"], PARAMS["tol"]

    def timed(k, sync_vp):
        tot = dict(ms=0.0, iters=0, ih=[], launches=0, timing={}, hist=0, clocks=[], epochs=0, step_ms=[])
        done = 0
        while done < k:
            solver.sync_vp = False
            solver.restore()
            kk = min(SAFE, k - done)
            solver.sync_vp = sync_vp
            # per-phase CUDA events only in the device-timed run (and on request in the e2e runs: they cost ~0.1 ms per step there)
            solver.set_option("timing", 1 if (not sync_vp or args.e2e_phases) else 0)
            solver.timing(reset=True)
            st0 = solver.stats()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            # NVML set-up (nvmlInit takes milliseconds) happens BEFORE the barrier: anything between the barrier and ev0 on
            # one rank shows up as waiting time inside the first collective of every other rank.
            sampler = ClockSampler(env.local, args.clock_period_ms * 1e-3) if (env.rank == 0 and args.clock_period_ms > 0) else None   # NVML polling from 8 processes contends on a driver lock
            env.barrier()
            marks = [torch.cuda.Event(enable_timing=True) for _ in range(kk)] if args.per_step else []
            ev0.record(stream)
            for i in range(kk):
                tot["ih"].append(solver.step(nit, tol))
                tot["iters"] += solver.last["iters"]
                if marks:
                    marks[i].record(stream)
            if sync_vp:
                solver.fence()      # the last step's write-back has landed in the caller's Vp: inside the timed region
            ev1.record(stream)
            env.barrier()
            tot["clocks"].append(sampler.stop() if sampler else {})
            ms = ev0.elapsed_time(ev1)
            if marks:
                tot["step_ms"] += [round(a.elapsed_time(b), 4) for a, b in zip([ev0] + marks[:-1], marks)]   # this rank, per step
            tot["ms"] += env.max_over_ranks(ms)   # device time, max over ranks
            st1 = solver.stats()
            tot["launches"] += st1["kernel_launches"] - st0["kernel_launches"]
            tot["hist"] = tot["hist"] + (st1["bfgs_hist"] - st0["bfgs_hist"])
            for kname, (pms, cnt) in solver.timing(reset=True).items():
                a, b = tot["timing"].get(kname, (0.0, 0))
                tot["timing"][kname] = (a + pms, b + cnt)
            solver.set_option("timing", 0)
            done += kk
            tot["epochs"] += 1
        cl = [c for c in tot["clocks"] if c.get("sm_mhz")]
        tot["clocks"] = dict(sm_mhz=statistics.median([c["sm_mhz"] for c in cl]) if cl else None,
                             sm_max_mhz=cl[0]["sm_max_mhz"] if cl else None,
                             reasons=sorted(set(sum([c.get("reasons", []) for c in tot["clocks"]], []))),
                             samples=sum(c.get("samples", 0) for c in tot["clocks"]))
        return tot

    r = timed(K, False)
    r2 = timed(K, "deferred")
    r3 = timed(K, "strict")
    # the write-back really wrote the caller's array: it equals the device state at the end of the last e2e epoch
    wb_ok = None
    if env.world == 1:
        wb_ok = bool(np.array_equal(mesh.Vp, solver.x()))
    g = solver.grid_dims()[0]
    world = env.world
    Ploc, Eloc = solver.local_sizes()
    launches = env.sum_over_ranks(r["launches"])
    value = E * r["iters"] / (r["ms"] / 1e3)
    e2e = E * r2["iters"] / (r2["ms"] / 1e3)
    e2e_strict = E * r3["iters"] / (r3["ms"] / 1e3)
    prox_ms, prox_cnt = r["timing"]["prox"]
    peak, peak_src = peaks()
    bytes_per_launch = algorithmic_bytes_prox(D, P, E, g) * Eloc   # this rank's launch (rank 0 reports)
    achieved = bytes_per_launch / (prox_ms / max(prox_cnt, 1) / 1e3) / 1e9
    traffic, traffic_src = ncu_traffic(wl["name"]) if world == 1 and not args.nx else (None, None)
    hist = r["hist"]
    kbar = float((hist * np.arange(52)).sum() / max(hist.sum(), 1))
    n = D * (D + 1)
    rec = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": K, "warmup": W,
        "ms_per_step": r["ms"] / K, "higher_is_better": True, "scaling": "strong" if (args.nx and world > 1) else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["label"], "elements": E, "points": P, "monitor_grid": g,
                   "monitor_grid_isotropic_fast_path": bool(args.iso_grid) and solver.iso_grid_active(),
                   "parallelism": (f"element-block partition over {world} GPUs, one process per GPU; halo sums of shared-vertex partials + "
                                   f"5-double all-gather per ADMM iteration by "
                                   + ("NVLink peer stores into the neighbours' mailboxes (CUDA IPC) with epoch flags"
                                      if solver.p2p_active() else "NCCL send/recv + ncclAllGather")
                                   + f"; rank 0 holds {Eloc:,} elements")
                   if world > 1 else "single GPU",
                   "l2": f"inputs_exceed_L2 (element state {8 * E * n * (n + 2) / 1e9:.2f} GB >> 126 MB L2)",
                   "timed_steps": (f"{K} steady-state outer steps in {r['epochs']} epoch(s) of <= {SAFE} (outer steps {W}..{W + min(SAFE, K) - 1}), every epoch "
                                   f"restored from one post-warm-up checkpoint"
                                   + ("; the reference algorithm itself breaks down at outer step ~12 on this workload, see DESIGN.md" if wl["safe"] < 10**9 else ""))},
        "steps_per_sec": K / (r["ms"] / 1e3), "admm_iters_per_sec": r["iters"] / (r["ms"] / 1e3),
        "admm_iters": r["iters"], "mean_bfgs_iters_per_element_prox": kbar,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": Ploc * D * 8 + 32,
                "ms_per_step": r2["ms"] / K, "phase_ms_per_step": {k: v[0] / K for k, v in r2["timing"].items()} if args.e2e_phases else None,
                "strict": {"value": e2e_strict, "ms_per_step": r3["ms"] / K,
                           "note": "write-back fenced after every step (the reference's updateAfterStep semantics)"},
                "writeback_verified": wb_ok, "step_ms_rank0": r2["step_ms"],
                "note": "MeshIntegrator.step incl. the per-step write-back of Vp into the caller's pinned host array: device-side snapshot + "
                        "copy on a copy stream overlapped with the next step (mmadmm_set_writeback mode 2), last copy fenced inside the timed "
                        "region.  h2d is 0 by design: like the reference's step(), the call takes no arrays after construction (state is "
                        "device-resident), scalars only"},
        "gpu_launches": launches,
        "roofline": {"kernel": "k_prox", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_per_launch,
                     "avg_launch_ms": prox_ms / max(prox_cnt, 1), "launches_timed": prox_cnt},
        "phase_ms_per_step": {k: v[0] / K for k, v in r["timing"].items()},
        "step_ms_rank0": r["step_ms"],
        "clocks": r["clocks"],
        "init_seconds": init_s, "energy_initial": e0, "energy_last": r2["ih"][-1],
        "warmup_steps": warm,
    }
    solver.sync_vp = False
    solver.close()
    del solver, mesh, vp_pinned
    if cpu_baseline and env.rank == 0:
        threads = os.cpu_count() or 1
        c = run_oracle(wl, min(K, 6), W, threads, one_thread_steps=1)
        rec["cpu_baseline"] = {"value": c["value"], "unit": UNIT, "cores": threads, "kind": "port", "sample": c["sample"],
                               "ms_per_step": c["ms_per_step"], "mean_bfgs_iters_per_element_prox": c["kbar"], "admm_iters": c["iters"],
                               "one_thread": c["one_thread"]}
    return rec


def method2_line(args):
    """`--method 2`: MeshIntegrator::backwardsEulerStep (Method 2, src/Mesh.cpp:1263-1341: Newton on the backward-Euler residual,
    finite-difference Jacobian, ILU(0) + BiCGSTAB as LASolver runs them) on SquareGrid 2D nx (default C2's 1024), MEx4.  One GPU.
    The first step builds the Jacobian and the ILU(0) factor and is the warm-up.  The Newton loop is host-driven, so the timed
    region is wall clock around K steps bracketed by device synchronises (CUDA events on the library's stream agree with it)."""
    env = Env()
    if env.world > 1:
        raise SystemExit("--method 2 runs on one GPU (mmadmm_backward_euler_step is single-GPU)")
    import numpy as np
    torch, mm = env.torch, env.mm
    nx = args.nx or 1024
    Vp, F, mask = mm.generate_uniform_rect_mesh(2, nx, nx, 0)
    mon = mm.TestMonitor(2, 4).evaluate_at_vertices(Vp)
    mesh = mm.Mesh(Vp.copy(), F.copy(), mask, mon, rho=50.0, tau=0.5, integration_mode=2)
    solver = mm.MeshIntegrator(0.025, mesh, sync_vp=False)
    for kv in args.opt:                                  # e.g. m2_row_block=1 (row per thread), 2 (default: one vertex), 4
        solver.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    warm = []
    for _ in range(max(1, min(args.warmup, 2))):
        t0 = time.perf_counter()
        solver.backwardsEulerStep(0.025, 1e-3)
        warm.append(dict(ms=(time.perf_counter() - t0) * 1e3, **solver.last))
    stream = torch.cuda.ExternalStream(solver.stream())
    K = args.steps
    st0 = solver.stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(env.local, 0.002)
    torch.cuda.synchronize()
    newton = lin = 0
    t0 = time.perf_counter()
    ev0.record(stream)
    for _ in range(K):
        ih = solver.backwardsEulerStep(0.025, 1e-3)
        newton += solver.last["newton_iters"]; lin += solver.last["lin_iters"]
    ev1.record(stream)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    rowptr, col, val = solver.backward_euler_jacobian()
    n, nnz = len(rowptr) - 1, len(val)
    launches = solver.stats()["kernel_launches"] - st0["kernel_launches"]
    # algorithmic bytes of one BiCGSTAB iteration: 2 ILU(0) applications (factor values + column indices once each, lower and
    # upper half together = nnz entries) + 2 SpMV (nnz entries) + ~14 vector passes
    bytes_iter = 2 * nnz * 12 + 2 * nnz * 12 + 14 * n * 8
    peak, peak_src = peaks()
    achieved = bytes_iter * lin / (ms * 1e-3) / 1e9 if lin else 0.0
    line = {"metric": "backward-Euler steps/sec (Method 2)", "value": K / (ms * 1e-3), "unit": "outer steps/s", "n_gpus": 1, "steps": K,
            "warmup": len(warm), "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"Method 2 (backward Euler) on SquareGrid 2D nx=ny={nx} ({F.shape[0]:,} triangles), MEx4, dt=0.025 tau=0.5, tol 1e-3",
                       "unknowns": n, "jacobian_nnz": nnz, "newton_iters": newton, "bicgstab_iters": lin, "options": args.opt},
            "wall_ms_per_step": wall / K * 1e3, "gpu_launches": launches,
            "e2e": {"value": K / wall, "unit": "outer steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 64 * max(lin, 1) // K,
                    "note": "wall clock through MeshIntegrator.backwardsEulerStep; as in the reference this path does not write Vp back per step"},
            "roofline": {"kernel": "ILU(0)+BiCGSTAB iteration (4 triangular solves -- k_tri_sf_block, the dataflow solve, by default --, k_spmv_dot x2, vector kernels)", "bound": "hbm", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": bytes_iter,
                         "note": "a dependency chain, not a stream: natural-order ILU(0) on a structured mesh has thousands of levels of a few "
                                 "hundred rows and what bounds a solve is the latency of one hop of the wavefront (DESIGN.md 3.5); the fraction "
                                 "is reported, not claimed"},
            "clocks": clocks, "warmup_steps": warm, "energy_last": ih}
    solver.close()
    if not args.no_cpu_baseline:
        import oracle_lib as O
        import pcg_oracle as PO
        cnx = 256
        Vc, Fc, mc = O.uniform_rect_mesh(2, cnx, cnx)
        orc = O.Oracle(Vc, Fc, mc, O.monitor_eval(2, 4, Vc), 50.0, 0.5, 0.025)
        rows, cols = orc.jacobian_entries()
        S = PO.LASolverJac(orc.P * orc.D, rows, cols) if PO.have_lasolver() else PO.RestatedJac(orc.P * orc.D, rows, cols)
        orc.backwards_euler_step(S, 1e-3)
        t0 = time.perf_counter()
        for _ in range(2):
            orc.backwards_euler_step(S, 1e-3)
        cms = (time.perf_counter() - t0) / 2 * 1e3
        line["cpu_baseline"] = {"value": 1e3 / cms, "unit": "outer steps/s", "cores": 1, "kind": "reference" if PO.have_lasolver() else "port",
                                "ms_per_step": cms,
                                "sample": f"oracle backward-Euler step (serial, like the reference's loops) with the linear solves by "
                                          f"{type(S).__name__} on SquareGrid 2D nx={cnx} ({Fc.shape[0]:,} triangles = 1/{(nx * nx) // (cnx * cnx)} of the GPU "
                                          f"arm's mesh: a bounded sample), 2 steps after 1 warm-up"}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4"])
    ap.add_argument("--method", type=int, default=0, choices=[0, 2], help="0: ADMM step (the benchmark); 2: backward Euler (Method 2) line")
    ap.add_argument("--nx", type=int, default=0, help="override the mesh resolution (with --gpus N: strong scaling at that size, e.g. SURVEY C5 = --nx 2048)")
    ap.add_argument("--per-step", action="store_true", help="also record one event per step (step_ms_rank0); adds host work between steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-workloads", action="store_true", help="skip the c3 / c4 sub-records of the default (c2) line")
    ap.add_argument("--clock-period-ms", type=float, default=2.0, help="NVML clock sampling period inside the timed region (0: off)")
    ap.add_argument("--e2e-phases", action="store_true", help="per-phase CUDA events inside the e2e runs too (diagnostics)")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the multi-GPU parity check against the oracle")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value (A/B runs), e.g. usum_period=0")
    ap.add_argument("--iso-grid", type=int, default=1, help="0: disable the scalar monitor-grid fast path (A/B runs)")
    ap.add_argument("--prox-variant", type=int, default=-1, help="prox kernel tile/occupancy variant (debugging; -1 = library default)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.method == 2:
        return method2_line(args)

    env = Env()
    W = min(max(args.warmup, 3), 6)
    K = args.steps
    parity = None
    if env.world > 1 and not args.no_parity:
        parity = multi_gpu_parity(env)
    wl = workload(args.gpus, args.workload, args.nx)
    line = run_workload(env, args, wl, K, W, cpu_baseline=(env.world == 1 and not args.no_cpu_baseline))
    line["fp64"] = {"peak_tflops_measured": env.mm.measure_fp64_peak(env.local),
                    "how": "DFMA microbenchmark, 8 independent chains per thread (mmadmm_measure_fp64_peak); the prox kernel's FP64 pipe "
                           "utilisation is the ncu figure sm__inst_executed_pipe_fp64 in profiles/"}
    if parity is not None:
        line["parity"] = parity
    if args.workload == "c2" and not args.nx and not args.no_extra_workloads:
        # BASELINE configs[2] and configs[3] at the same N, same timing rules (driver-run numbers for the FREE-rim disc and 3D)
        line["workloads"] = {}
        for name in ("c3", "c4"):
            sub = run_workload(env, args, workload(args.gpus, name), K, W, cpu_baseline=False)
            line["workloads"][name] = {k: sub[k] for k in ("value", "unit", "ms_per_step", "steps_per_sec", "admm_iters", "mean_bfgs_iters_per_element_prox",
                                                           "roofline", "e2e", "config", "phase_ms_per_step", "gpu_launches", "clocks", "init_seconds")}
        line["gpu_launches_all_workloads"] = line["gpu_launches"] + sum(w["gpu_launches"] for w in line["workloads"].values())
    if env.rank == 0:
        print(json.dumps(line))
    if env.dist:
        env.dist.barrier()
        env.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
