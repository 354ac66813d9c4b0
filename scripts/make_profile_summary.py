#!/usr/bin/env python
"""Turn one scripts/gpu_profile.sh output directory (gpurun_out/<tag>) into the tracked evidence under profiles/:
bench lines, ncu launch lists (csv), raw-metric excerpts of the full captures, and a markdown summary.
usage: make_profile_summary.py <gpurun_out/tag> <profiles prefix, e.g. r01_d>"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

src, pre = sys.argv[1], sys.argv[2]
out = "profiles"
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum"]


def raw(rep):
    p = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
    rows = list(csv.reader(p.stdout.splitlines()))
    if len(rows) < 3:
        return []
    h, u = rows[0], rows[1]
    res = []
    for d in rows[2:]:
        m = {"kernel": d[h.index("Kernel Name")]}
        for i, n in enumerate(h):
            short = n.split(".", 2)[-1] if n.startswith(("FBSP.", "breakdown:")) else n   # section-prefixed names of the raw page
            for cand in (n, short):
                if cand in KEYS or ("issue_stalled" in cand and "per_issue_active" in cand):
                    m[cand] = (d[i], u[i])
                    break
        res.append(m)
    return res


def launches(path):
    rows = list(csv.DictReader(l for l in open(path) if not l.startswith("==")))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        k = r["Kernel Name"]
        agg[k][0] += 1
        agg[k][1] += float(r["Metric Value"].replace(",", ""))
    tot = sum(v[1] for v in agg.values())
    return sorted(((k, n, t / n / 1e3, 100 * t / tot) for k, (n, t) in agg.items()), key=lambda x: -x[3]), tot


def load_line(f):
    return json.loads([ln for ln in open(f).read().split("\n") if ln.startswith("{")][-1])


md = [f"# Round {int(pre[1:3])}, capture {pre[-1].upper()} (`profiles/{pre}_*`)\n",
      "Produced by `scripts/gpu_profile.sh` on one B200 (bench lines first, each ncu pass only after its command had exited 0 "
      "without ncu) and `scripts/make_profile_summary.py`.\n"]
for wl in ("c2", "c3", "c4", "ref"):
    f = os.path.join(src, f"bench_{wl}.json")
    if not os.path.exists(f) or not open(f).read().strip():
        continue
    shutil.copy(f, os.path.join(out, f"{pre}_bench_{wl}.json"))
    d = load_line(f)
    if wl == "ref":
        md.append(f"* reference arm (`--impl reference`): {d['value'] / 1e6:.2f} M element-prox/s, {d['cpu_baseline']['sample']}\n")
        continue
    r = d["roofline"]
    md.append(f"## {wl}: {d['config']['workload']}\n")
    md.append(f"| quantity | value |\n|---|---|\n| value (state resident) | {d['value'] / 1e9:.3f} G element-prox/s, {d['ms_per_step']:.3f} ms per outer step "
              f"({d['admm_iters']} ADMM iterations in {d['steps']} steps, k̄ = {d['mean_bfgs_iters_per_element_prox']:.2f}) |\n"
              f"| e2e (Vp written back to pinned host memory every step, {d['e2e']['d2h_bytes_per_step'] / 1e6:.1f} MB D2H) | {d['e2e']['value'] / 1e9:.3f} G/s, {d['e2e']['ms_per_step']:.3f} ms per step |\n"
              f"| prox kernel, CUDA events in the timed region | {r['avg_launch_ms']:.4f} ms per launch, {r['achieved']:.0f} GB/s algorithmic = {r['frac']:.3f} of {r['peak']:.0f} GB/s ({r['peak_source']}) |\n"
              f"| phases (ms per step) | {', '.join(f'{k} {v:.4f}' for k, v in d['phase_ms_per_step'].items())} |\n"
              f"| clocks during the timed region | {d['clocks']} |\n"
              f"| FP64 FMA peak measured | {d.get('fp64', {}).get('peak_tflops_measured', float('nan')):.1f} TFLOP/s |\n"
              + (f"| CPU oracle port | {d['cpu_baseline']['value'] / 1e6:.2f} M element-prox/s on {d['cpu_baseline']['cores']} threads ({d['cpu_baseline']['sample']}) |\n" if "cpu_baseline" in d else ""))
for wl in ("c2", "c4"):
    f = os.path.join(src, f"launches_{wl}.csv")
    if not os.path.exists(f):
        continue
    shutil.copy(f, os.path.join(out, f"{pre}_launches_{wl}.csv"))
    rows, tot = launches(f)
    md.append(f"### ncu launch list, {wl} (`{pre}_launches_{wl}.csv`; cold-cache, serialised launches: shares, not absolutes)\n")
    md.append("| kernel | launches | avg µs | share |\n|---|---|---|---|\n" +
              "".join(f"| `{k[:90]}` | {n} | {a:.1f} | {s:.1f} % |\n" for k, n, a, s in rows[:12]))
    # share of the prox kernel within one steady-state outer step (prox + the two x-updates + predictor + finalize),
    # to be compared with the phase timers of the bench line (k_dfma_peak is the FP64 microbenchmark, outside any step)
    avg = {k: a for k, n, a, s in rows}
    pick = lambda pat: next((a for k, a in avg.items() if pat in k), 0.0)
    D = "3" if wl == "c4" else "2"
    prox = (pick("k_prox_q") or pick("k_prox_vw")) if wl == "c4" else pick("k_prox_p")
    # (k_xfirst<D,1> = predictor + first x-solve, k_xupdate<D,0,0,1,1> = the x-update after the sweep on the dual sums; the
    # pre-"usum" names are kept so that older captures still summarise)
    step = (prox + pick(f"k_xupdate<{D}, 0, 0, 1, 1>") + pick(f"k_xfirst<{D}, 1>") + pick(f"k_xupdate<{D}, 0, 0, 1>(") +
            pick(f"k_xupdate<{D}, 1, 0, 0>(") + pick("k_predict_extrap") + pick("k_finalize"))
    if step > 0:
        bj = load_line(os.path.join(src, f"bench_{wl}.json"))
        ph = bj["phase_ms_per_step"]
        md.append(f"Steady-state step from this list: prox {prox:.0f} of {step:.0f} µs = {100 * prox / step:.1f} %; phase timers of the bench line: "
                  f"{100 * ph['prox'] / (ph['prox'] + ph['xupdate'] + ph['predict'] + ph['residual']):.1f} %.\n")
for name in ("prof_c2", "prof_c4", "prof_xup", "prof_tri"):
    rep = os.path.join(src, name + ".ncu-rep")
    if not os.path.exists(rep):
        continue
    res = raw(rep)
    with open(os.path.join(out, f"{pre}_{name}_raw.json"), "w") as fo:
        json.dump(res, fo, indent=1)
    for m in res:
        md.append(f"### `ncu --set full`: `{m['kernel'][:100]}` (`{pre}_{name}_raw.json`)\n")
        g = lambda k: m.get(k, ("?", ""))[0]
        stalls = sorted(((float(v[0]), k.split("issue_stalled_")[1].split("_per_issue")[0]) for k, v in m.items()
                         if "issue_stalled" in k and v[0] not in ("", "n/a")), reverse=True)[:6]
        md.append(f"* duration {g('gpu__time_duration.sum')} {m.get('gpu__time_duration.sum', ('', ''))[1]}; DRAM read {g('dram__bytes_read.sum')} + write {g('dram__bytes_write.sum')} "
                  f"{m.get('dram__bytes_read.sum', ('', ''))[1]} (= `roofline.traffic`); DRAM throughput {g('dram__throughput.avg.pct_of_peak_sustained_elapsed')} % of peak\n"
                  f"* {g('launch__registers_per_thread')} registers/thread, grid {g('launch__grid_size')} x {g('launch__block_size')}, "
                  f"warps active {g('sm__warps_active.avg.pct_of_peak_sustained_active')} %, local loads/stores {g('sass__inst_executed_local_loads')}/{g('sass__inst_executed_local_stores')}\n"
                  f"* FP64 pipe {g('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active')} %, issue slots {g('smsp__issue_active.avg.pct_of_peak_sustained_active')} %, "
                  f"L1 data pipe {g('l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed')} %, L2 hit {g('lts__t_sector_hit_rate.pct')} %\n"
                  f"* stalls per issue: {', '.join(f'{n} {v:.2f}' for v, n in stalls)}\n")
# SASS evidence that the library is Blackwell-native code: bulk-copy (TMA) engine, mbarriers, cp.async, cluster barriers, FP64 FMAs;
# and what must NOT be there (tensor-core MMAs: nothing here is a contraction)
so = os.path.join("mm-admm_b200", "libmmadmm_b200.so")
if os.path.exists(so):
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    import re
    pats = ["UBLKCP.S.G", "UBLKCP.G.S", "UTMALDG", "SYNCS", "LDGSTS", "UCGABAR_ARV", "UCGABAR_WAIT", "DFMA", "DMUL", "DADD", "MUFU.RSQ64H", "SHFL",
            "UTCHMMA", "UTCQMMA", "HMMA", "IMMA", "DMMA", "MEMBAR.SC.SYS", "ST.E.STRONG.SYS", "LD.E.STRONG.SYS"]
    counts = {k: len(re.findall(r"\b" + re.escape(k), sass)) for k in pats}
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    json.dump({"library": so, "arch": arch, "instruction_counts": counts}, open(os.path.join(out, f"{pre}_sass_counts.json"), "w"), indent=1)
    md.append(f"### SASS of `{so}` (`cuobjdump -sass`, `{pre}_sass_counts.json`): arch {arch}\n")
    md.append("| mnemonic | count | what it is |\n|---|---|---|\n" + "".join(
        f"| `{k}` | {counts[k]} | {w} |\n" for k, w in [
            ("UBLKCP.S.G", "cp.async.bulk global -> shared (TMA bulk copy engine): inverse-Hessian tiles in"),
            ("UBLKCP.G.S", "cp.async.bulk shared -> global: inverse-Hessian tiles out"),
            ("UTMALDG", "tensor-map TMA loads (none: one contiguous bulk copy per tile needs no tensor map)"),
            ("SYNCS", "mbarrier operations (transaction barriers of the bulk copies)"),
            ("LDGSTS", "cp.async per-thread gathers into shared memory (next tile's vertices, connectivity)"),
            ("UCGABAR_ARV", "thread-block-cluster barrier arrive (k_tri_cluster)"), ("UCGABAR_WAIT", "cluster barrier wait"),
            ("DFMA", "FP64 fused multiply-add"), ("SHFL", "warp shuffles (3D prox exchange, reductions)"),
            ("UTCHMMA", "tcgen05 MMA (none: no contraction on this path)"), ("HMMA", "mma.sync (none)"), ("DMMA", "FP64 mma (none)")]))
open(os.path.join(out, f"{pre}_summary.md"), "w").write("\n".join(md))
print("wrote", os.path.join(out, f"{pre}_summary.md"))
