#!/usr/bin/env python
"""bench.py -- ADMM steps/sec & element-prox updates/sec of the implicit MMPDE step on B200.

A "step" is one MeshIntegrator::step (reference src/MeshIntegrator.cpp:101-191) on the workload named in
`config.workload`.  At N=1 that is BASELINE.json configs[1]: the 2D unit square, SquareGrid nx=ny=1024
(4,194,304 triangles -- the generator emits 4 triangles per cell), shock-layer monitor MEx4, rho=50, tau=0.5,
dt=0.025, AdmmIter=10, tol=1e-3.  Warm-up steps 0..W-1 contain the reference's start-up work (finite-difference
Hessians, gradient predictor); the timed steps are steady state (extrapolation predictor).

  value  : element-prox updates/s (= E x ADMM iterations executed / time), state resident in HBM
  e2e    : same metric through the reference-facing host API (MeshIntegrator.step), which like the reference's
           Mesh::updateAfterStep writes the caller's host Vp every step (D2H of P*D doubles into pinned memory)
  roofline: k_prox (the dominant kernel) against the measured HBM copy bandwidth
  cpu_baseline / --impl reference: the CPU oracle port (oracle/, OpenMP over elements like the reference's
           -D THREADS build) on a bounded sample of the same workload, on this box's host cores.
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "ADMM steps/sec & element-prox updates/sec"
UNIT = "element-prox updates/s"
PARAMS = dict(rho=50.0, tau=0.5, dt=0.025, admm_iter=10, tol=1e-3, mon2d=4, mon3d=3)
WEAK_NX_2D = {1: 1024, 2: 1448, 4: 2048, 8: 2896}   # SURVEY 8(d) C5 weak-scaling ladder
WEAK_N_DISC = {1: 816, 2: 1154, 4: 1632, 8: 2308}   # C3: 6 N^2 triangles = 4.0 / 8.0 / 16.0 / 32.0 M
WEAK_NX_3D = {1: 80, 2: 101, 4: 127, 8: 160}        # C4: 12 nx^3 tets = 6.1 / 12.4 / 24.6 / 49.2 M


def workload(n_gpus, name):
    """BASELINE.json configs: c2 = configs[1] (the N=1 bench workload; N>1: the C5 weak-scaling ladder), c3 = configs[2]
    (disc with BOUNDARY_FREE rim), c4 = configs[3] (3DMonitor280)."""
    if name == "c4":
        nx = WEAK_NX_3D.get(n_gpus, 80)
        return dict(D=3, nx=nx, kind="box", mon=3, btype=1, rho=50.0, tau=0.5, dt=0.025, cpu_nx=20,
                    label=f"SquareGrid 3D nx=ny=nz={nx} ({12 * nx ** 3:,} tets), MEx33D, rho=50 tau=0.5 dt=0.025 AdmmIter=10")
    if name == "c3":
        N = WEAK_N_DISC.get(n_gpus, 816)
        return dict(D=2, nx=N, kind="disc", mon=5, btype=0, rho=5.0, tau=0.1, dt=0.05, cpu_nx=200,
                    label=f"unit disc of {N} hexagonal rings ({6 * N * N:,} triangles), rim nodes BOUNDARY_FREE (reference semantics: "
                          f"gradient live, no projection), MEx5 spiral monitor, rho=5 tau=0.1 dt=0.05 AdmmIter=10 (Monitor310's parameters)")
    nx = WEAK_NX_2D.get(n_gpus, 1024)
    return dict(D=2, nx=nx, kind="box", mon=4, btype=1, rho=50.0, tau=0.5, dt=0.025, cpu_nx=256,
                label=f"SquareGrid 2D nx=ny={nx} ({4 * nx * nx:,} triangles), MEx4 shock-layer monitor, "
                      f"rho=50 tau=0.5 dt=0.025 AdmmIter=10 tol=1e-3")


def make_mesh(wl, nx):
    """(Vp, F, mask, monitor table) of a workload at resolution nx, from the package's host-side generators."""
    mm = importlib.import_module("mm-admm_b200")
    D = wl["D"]
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

    def __init__(self, index=0):
        import threading
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
            time.sleep(0.002)

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


def run_oracle_sample(wl, nx, steps, warmup, threads, safe):
    """CPU oracle port on a bounded sample, in the same epochs of <= `safe` steady-state steps as the GPU arm;
    returns (element-prox updates/s over the timed steps, ms/step, description, E, ADMM iterations)."""
    import oracle_lib as O
    D = wl["D"]
    Vp, F, mask, mon = make_mesh(wl, nx)
    iters, dt, done, epochs = 0, 0.0, 0, 0
    while done < steps:
        orc = O.Oracle(Vp, F, mask, mon, wl["rho"], wl["tau"], wl["dt"], nthreads=threads)
        for _ in range(warmup):
            orc.step(PARAMS["admm_iter"], PARAMS["tol"])
        kk = min(safe, steps - done)
        t0 = time.perf_counter()
        for _ in range(kk):
            r = orc.step(PARAMS["admm_iter"], PARAMS["tol"])
            if r["status"]:
                raise SystemExit("oracle reported an inverted element inside the timed region")
            iters += r["iters"]
        dt += time.perf_counter() - t0
        done += kk
        epochs += 1
    E = F.shape[0]
    shape = f"disc of {nx} rings" if wl["kind"] == "disc" else f"SquareGrid {D}D nx={nx}"
    desc = (f"oracle port (C++/OpenMP, {threads} threads) on {shape} ({E:,} elements), same monitor/"
            f"parameters, {steps} steady-state outer steps in {epochs} epoch(s) after {warmup} warm-up steps ({iters} ADMM iterations)")
    return E * iters / dt, dt / steps * 1e3, desc, E, iters


def safe_steps(workload_name, W):
    # see main(): the reference algorithm breaks down at outer step ~12 on the shock-layer workload
    return max(1, 9 - W) if workload_name == "c2" else 10**9


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wl = workload(args.gpus, args.workload)
    D = wl["D"]
    threads = os.cpu_count() or 1
    nx = args.cpu_nx or wl["cpu_nx"]
    W = min(max(args.warmup, 3), 6)
    v, ms, desc, E, iters = run_oracle_sample(wl, nx, args.steps, W, threads, safe_steps(args.workload, W))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl["label"], "sample": desc},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "steps_per_sec": args.steps / (ms * args.steps / 1e3), "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c4"])
    ap.add_argument("--nx", type=int, default=0, help="override the mesh resolution (with --gpus N: strong scaling at that size, e.g. SURVEY C5 = --nx 2048)")
    ap.add_argument("--per-step", action="store_true", help="also record one event per step (step_ms_rank0); adds host work between steps")
    ap.add_argument("--cpu-nx", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value (A/B runs), e.g. b_evict_first=1")
    ap.add_argument("--iso-grid", type=int, default=1, help="0: disable the scalar monitor-grid fast path (A/B runs)")
    ap.add_argument("--prox-variant", type=int, default=-1, help="prox kernel tile/occupancy variant (debugging; -1 = library default)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import numpy as np
    import torch
    mm = importlib.import_module("mm-admm_b200")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: mm-admm_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:   # one process per GPU (torchrun); torch.distributed is plumbing only: id broadcast, barrier, max-over-ranks
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def new_comm():
        if world == 1:
            return None
        t = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(mm.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        return (rank, world, bytes(t.cpu().numpy().tobytes()))
    W = min(max(args.warmup, 3), 6)
    K = args.steps
    wl = workload(args.gpus, args.workload)
    D = wl["D"]
    nx = args.nx or wl["nx"]
    Vp, F, mask, mon = make_mesh(wl, nx)
    P, E = Vp.shape[0], F.shape[0]
    # caller-owned Vp in pinned host memory (what the reference's Mesh writes back every step)
    vp_pinned = torch.empty((P, D), dtype=torch.float64).pin_memory()
    state = {}

    def fresh_solver():
        """New Mesh + MeshIntegrator from the initial mesh, then the W untimed warm-up steps (steps 0..W-1)."""
        if state.get("solver") is not None:
            state["solver"].close()
        vp_pinned.numpy()[:] = Vp
        mesh = mm.Mesh(vp_pinned.numpy(), F.copy(), mask, mon, rho=wl["rho"], tau=wl["tau"])
        mesh.Vp = vp_pinned.numpy()
        t0 = time.perf_counter()
        solver = mm.MeshIntegrator(wl["dt"], mesh, device=local, sync_vp=False, comm=new_comm())
        state["init_s"] = time.perf_counter() - t0
        if args.prox_variant >= 0:
            solver.set_option("prox_variant", args.prox_variant)
        if not args.iso_grid:
            solver.set_option("iso_grid", 0)
        for kv in args.opt:
            solver.set_option(kv.split("=")[0], int(kv.split("=")[1]))
        state["e0"] = solver.getEnergy()
        warm = []
        for _ in range(W):
            t1 = time.perf_counter()
            solver.step(PARAMS["admm_iter"], PARAMS["tol"])
            warm.append(dict(ms=(time.perf_counter() - t1) * 1e3, **solver.last))
        state["warm"] = warm
        state["solver"] = solver
        return solver

    # The reference algorithm breaks down on this workload at outer step ~12 (0/0 in the BFGS update once an
    # element's step underflows, src/Mesh.cpp:846-848; reproduced identically by the oracle and by this library --
    # tests/test_gpu_parity.py::test_reference_breakdown_is_reproduced).  So a timed region never runs past outer
    # step W+SAFE-1: K timed steps are taken in epochs of <= SAFE steady-state steps, each epoch restarting from the
    # initial mesh with its own untimed warm-up; every epoch is bracketed by a synchronize on both sides.
    SAFE = safe_steps(args.workload, W)

    def timed(k, sync_vp):
        tot = dict(ms=0.0, iters=0, ih=[], launches=0, timing={}, hist=0, clocks=[], epochs=0, step_ms=[])
        done = 0
        while done < k:
            solver = fresh_solver()
            stream = torch.cuda.ExternalStream(solver.stream())
            kk = min(SAFE, k - done)
            solver.sync_vp = sync_vp
            solver.set_option("timing", 1)
            solver.timing(reset=True)
            st0 = solver.stats()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            # NVML set-up (nvmlInit takes milliseconds) happens BEFORE the barrier: anything between the barrier and ev0 on
            # one rank shows up as waiting time inside the first collective of every other rank.
            sampler = ClockSampler(local) if rank == 0 else None   # rank 0 reports; NVML polling from 8 processes contends on a driver lock
            torch.cuda.synchronize()
            if dist:
                dist.barrier()
                torch.cuda.synchronize()
            marks = [torch.cuda.Event(enable_timing=True) for _ in range(kk)] if args.per_step else []
            nit, tol = PARAMS["admm_iter"], PARAMS["tol"]
            ev0.record(stream)
            for i in range(kk):
                tot["ih"].append(solver.step(nit, tol))
                tot["iters"] += solver.last["iters"]
                if marks:
                    marks[i].record(stream)
            ev1.record(stream)
            torch.cuda.synchronize()
            if dist:
                dist.barrier()
                torch.cuda.synchronize()
            tot["clocks"].append(sampler.stop() if sampler else {})
            ms = ev0.elapsed_time(ev1)
            if marks:
                tot["step_ms"] += [round(a.elapsed_time(b), 4) for a, b in zip([ev0] + marks[:-1], marks)]   # this rank, per step
            if dist:   # device time, max over ranks
                tms = torch.tensor([ms], dtype=torch.float64, device="cuda")
                dist.all_reduce(tms, op=dist.ReduceOp.MAX)
                ms = float(tms.item())
            tot["ms"] += ms
            st1 = solver.stats()
            tot["launches"] += st1["kernel_launches"] - st0["kernel_launches"]
            tot["hist"] = tot["hist"] + (st1["bfgs_hist"] - st0["bfgs_hist"])
            for kname, (ms, cnt) in solver.timing(reset=True).items():
                a, b = tot["timing"].get(kname, (0.0, 0))
                tot["timing"][kname] = (a + ms, b + cnt)
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
    r2 = timed(K, True)
    solver = state["solver"]
    g = solver.grid_dims()[0]
    init_s, e0, warm = state["init_s"], state["e0"], state["warm"]
    launches = r["launches"]
    Ploc, Eloc = solver.local_sizes()
    if dist:
        tl = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(tl)
        launches = int(tl.item())
    value = E * r["iters"] / (r["ms"] / 1e3)
    e2e = E * r2["iters"] / (r2["ms"] / 1e3)
    prox_ms, prox_cnt = r["timing"]["prox"]
    peak, peak_src = peaks()
    fp64_peak = mm.measure_fp64_peak(local)
    bytes_per_launch = algorithmic_bytes_prox(D, P, E, g) * Eloc   # this rank's launch (rank 0 reports)
    achieved = bytes_per_launch / (prox_ms / max(prox_cnt, 1) / 1e3) / 1e9
    traffic, traffic_src = ncu_traffic(args.workload) if world == 1 and not args.nx else (None, None)
    hist = r["hist"]
    kbar = float((hist * np.arange(52)).sum() / max(hist.sum(), 1))
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": K, "warmup": W,
        "ms_per_step": r["ms"] / K, "higher_is_better": True, "scaling": "strong" if (args.nx and world > 1) else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": wl["label"], "elements": E, "points": P, "monitor_grid": g, "monitor_grid_isotropic_fast_path": bool(args.iso_grid) and solver.iso_grid_active(),
                   "parallelism": (f"element-block partition over {world} GPUs, one process per GPU; halo sums of shared-vertex partials + "
                                   f"5-double all-gather per ADMM iteration by "
                                   + ("NVLink peer stores into the neighbours' mailboxes (CUDA IPC) with epoch flags"
                                      if solver.p2p_active() else "NCCL send/recv + ncclAllGather")
                                   + f"; rank 0 holds {Eloc:,} elements")
                   if world > 1 else "single GPU", "l2": "inputs_exceed_L2 "
                   f"(element state {8 * E * (D * (D + 1)) * (D * (D + 1) + 2) / 1e9:.2f} GB >> 126 MB L2)",
                   "timed_steps": (f"{K} steady-state outer steps in {r['epochs']} epoch(s) of <= {SAFE} (outer steps {W}..{W + min(SAFE, K) - 1} after a "
                                   f"fresh start; the reference algorithm itself breaks down at outer step ~12 on this workload, see DESIGN.md)")
                   if args.workload == "c2" else f"{K} steady-state outer steps {W}..{W + K - 1} after {W} warm-up steps"},
        "steps_per_sec": K / (r["ms"] / 1e3), "admm_iters_per_sec": r["iters"] / (r["ms"] / 1e3),
        "admm_iters": r["iters"], "mean_bfgs_iters_per_element_prox": kbar,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": Ploc * D * 8 + 32,
                "ms_per_step": r2["ms"] / K, "note": "MeshIntegrator.step incl. per-step write-back of Vp to pinned host memory; "
                "the step takes no array inputs after construction (state is device-resident), scalars only"},
        "gpu_launches": launches,
        "roofline": {"kernel": "k_prox", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_per_launch,
                     "avg_launch_ms": prox_ms / max(prox_cnt, 1), "launches_timed": prox_cnt},
        "fp64": {"peak_tflops_measured": fp64_peak, "how": "DFMA microbenchmark, 8 independent chains per thread (mmadmm_measure_fp64_peak); "
                 "the prox kernel's FP64 pipe utilisation is the ncu figure sm__inst_executed_pipe_fp64 in profiles/"},
        "phase_ms_per_step": {k: v[0] / K for k, v in r["timing"].items()},
        "step_ms_rank0": r["step_ms"],
        "clocks": r["clocks"],
        "init_seconds": init_s, "energy_initial": e0, "energy_last": r2["ih"][-1],
        "warmup_steps": warm,
    }
    if rank == 0 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        cnx = args.cpu_nx or wl["cpu_nx"]
        v, ms, desc, _, _ = run_oracle_sample(wl, cnx, 6, 3, threads, safe_steps(args.workload, 3))
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc, "ms_per_step": ms}
    if rank == 0:
        print(json.dumps(line))
    solver.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
