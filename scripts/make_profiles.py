#!/usr/bin/env python
"""Turn the scratch ncu reports (gpurun_out/prof_*.ncu-rep) and the built library into the tracked evidence under
profiles/: one Markdown summary per kernel (key ncu metrics, hottest SASS instructions with their stall reason) and
one SASS listing per hot kernel with the mnemonic counts that prove tcgen05 / TMA / peer-memory use."""
import csv
import io
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_algorithm_pyspark_b200", "_lib", "libdapb200.so")
OUT = os.path.join(ROOT, "profiles")
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__bytes_read.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum"]
MNEMONICS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UBLKCP", "LDTM", "STTM", "SYNCS", "RED", "ATOM", "HMMA", "LDGSTS", "FMNMX3", "LDG", "STG",
             "LDGMC", "ACQBULK", "PREEXIT", "ST.E", "LD.E"]


def ncu(args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True).stdout


def summarize_report(name, rep):
    raw = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    if len(raw) < 3:
        return None
    hdr, units, vals = raw[0], raw[1], raw[2]
    kern = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else name
    lines = [f"# ncu capture: {name}", "", f"Kernel: `{kern}`  ", f"Source report: `gpurun_out/{os.path.basename(rep)}` "
             "(`ncu --set full --clock-control none --import-source on`, one launch, on a B200 via gpurun).", "",
             "| metric | value | unit |", "|---|---|---|"]
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            lines.append(f"| {k} | {vals[i]} | {units[i]} |")
    src = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv"]))))
    if len(src) > 3:
        h = src[1]
        ix = {c: i for i, c in enumerate(h)}
        stalls = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
        items, tot = [], 0.0
        for r in src[2:]:
            try:
                s = float(r[ix["Warp Stall Sampling (All Samples)"]])
            except Exception:
                continue
            tot += s
            items.append((s, r))
        agg = Counter()
        for s, r in items:
            for c in stalls:
                try:
                    agg[c] += float(r[ix[c]])
                except Exception:
                    pass
        lines += ["", f"Warp-stall samples: {int(tot)}.  Mix: " +
                  ", ".join(f"{k} {100 * v / max(tot, 1):.1f}%" for k, v in agg.most_common(6)), "",
                  "Hottest SASS instructions (samples, share, executions, dominant stall):", "", "```"]
        items.sort(key=lambda x: -x[0])
        for s, r in items[:25]:
            dom = max(stalls, key=lambda c: float(r[ix[c]] or 0))
            lines.append(f"{int(s):7d} {100 * s / max(tot, 1):5.1f}%  exec={r[ix['Instructions Executed']]:>11}  {dom:22s} {r[ix['Source']].strip()[:90]}")
        lines.append("```")
    open(os.path.join(OUT, f"ncu_{name}.md"), "w").write("\n".join(lines) + "\n")
    return kern


def sass_listings():
    os.makedirs(os.path.join(OUT, "sass"), exist_ok=True)
    names = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs = re.findall(r"Function : (\S+)", names)
    wanted = {"kmeans_tc_kernelILi128": "kmeans_tc_kernel_d128", "gram_tc_kernel": "gram_tc_kernel",
              "lr_step_kernelI13__nv_bfloat16Li32": "lr_step_kernel_bf16_d1024", "lr_parity_kernel": "lr_parity_kernel",
              "pagerank_step_kernel": "pagerank_step_kernel", "mc_pi_kernel": "mc_pi_kernel", "tc_expand_kernel": "tc_expand_kernel",
              "kmeans_finalize_kernelIfE": "kmeans_finalize_kernel_f32", "lr_tc_kernel": "lr_tc_kernel",
              "pagerank_run_kernelIfE": "pagerank_run_kernel_f32", "pagerank_run_kernelIdE": "pagerank_run_kernel_f64",
              "tc_closure_kernel": "tc_closure_kernel", "chol_coop_kernelIfE": "chol_coop_kernel_f32",
              "eigh_jacobi_kernelIdE": "eigh_jacobi_kernel_f64", "twoshot_kernel": "twoshot_allreduce_kernel",
              "allgather_kernelIfE": "allgather_kernel_f32"}
    summary = ["# SASS evidence per hot kernel", "",
               "`cuobjdump -sass` of `distributed_algorithm_pyspark_b200/_lib/libdapb200.so` (sm_100a).  `UTCHMMA` = tcgen05.mma, "
               "`LDTM` = tcgen05.ld, `UTMALDG` = TMA tensor load, `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier ops, "
               "`RED` = no-return atomics (L2 / peer).  Peer-memory pushes are ordinary `ST.E`/`RED` on mapped peer pointers "
               "followed by `fence.sys` (`MEMBAR.SC.SYS` / `ERRBAR`) and a release store of the epoch flag.  NVLS: "
               "`multimem.ld_reduce` = `LDGMC.E.ADD...`, `multimem.st` = `STG.E...STRONG.SYS` on the multicast mapping "
               "(column `STG.SYS` counts the `STG` with `.STRONG.SYS`: multimem stores and release flag stores).  "
               "Programmatic dependent launch: `griddepcontrol.wait` = `ACQBULK`, `griddepcontrol.launch_dependents` = `PREEXIT`.", "",
               "| kernel | instructions | " + " | ".join(MNEMONICS) + " | STG.SYS |", "|---|---|" + "---|" * (len(MNEMONICS) + 1)]
    for f in funcs:
        for key, short in wanted.items():
            if key in f:
                txt = subprocess.run(["cuobjdump", "-sass", "-fun", f, LIB], capture_output=True, text=True).stdout
                body = [l for l in txt.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
                ops = Counter()
                for l in body:
                    t = re.sub(r"/\*.*?\*/", "", l).strip()
                    t = re.sub(r"^@!?U?P\d\s+", "", t)
                    if t:
                        ops[t.split()[0]] += 1
                counts = [sum(v for k, v in ops.items() if k.startswith(m)) for m in MNEMONICS]
                stg_sys = sum(v for k, v in ops.items() if k.startswith("STG") and "STRONG.SYS" in k)
                summary.append(f"| {short} | {len(body)} | " + " | ".join(str(c) for c in counts) + f" | {stg_sys} |")
                clean = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in txt.splitlines() if not re.match(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", l)]
                open(os.path.join(OUT, "sass", short + ".sass"), "w").write("\n".join(clean) + "\n")
    open(os.path.join(OUT, "SASS_SUMMARY.md"), "w").write("\n".join(summary) + "\n")


def main():
    os.makedirs(OUT, exist_ok=True)
    for rep in sorted(os.listdir(os.path.join(ROOT, "gpurun_out"))):
        if rep.startswith("prof_") and rep.endswith(".ncu-rep"):
            name = rep[len("prof_"):-len(".ncu-rep")]
            print(name, summarize_report(name, os.path.join(ROOT, "gpurun_out", rep)))
    sass_listings()


if __name__ == "__main__":
    main()
