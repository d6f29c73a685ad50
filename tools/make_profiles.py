"""Turn gpurun_out/ evidence into the tracked summaries under profiles/."""
import csv
import glob
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GO = os.path.join(ROOT, "gpurun_out")
os.makedirs(os.path.join(OUT, "sass"), exist_ok=True)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__cluster_size", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "sm__cycles_elapsed.max",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum"]


_units = {}


def ncu_raw(rep):
    """`rep` is a .ncu-rep (converted here) or the raw-page CSV already extracted on the GPU box (<name>_raw.csv)."""
    if rep.endswith(".csv"):
        text = open(rep).read()
    else:
        text = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(text.splitlines()))
    if len(rows) < 3:
        return [], []
    global _units
    _units = dict(zip(rows[0], rows[1]))
    return rows[0], rows[2:]


def summarise(rep, name):
    hdr, data = ncu_raw(rep)
    if not hdr:
        return
    cols = [k for k in KEYS if k in hdr]
    lines = [f"# ncu --set full --clock-control none: {os.path.basename(rep)}", "",
             f"Roofline denominators (MEASURED_PEAKS.json): HBM {peaks.get('hbm_gbs')} GB/s, bf16 {peaks.get('bf16_tflops')} TFLOP/s (of measured).", "",
             "| kernel | grid | " + " | ".join(c.replace('smsp__average_warps_issue_stalled_', 'stall_').replace('_per_issue_active.ratio', '')
                                               + (f" [{_units.get(c)}]" if _units.get(c) else "") for c in cols) + " |",
             "|---|---|" + "---|" * len(cols)]
    ki, gi = hdr.index("Kernel Name"), hdr.index("Grid Size")
    for d in data:
        nm = re.sub(r"\(.*", "", d[ki]).replace("void <unnamed>::", "").replace("<unnamed>::", "")
        if "at::" in nm or nm.startswith("elementwise_kernel") or "cub::" in nm:
            continue                      # torch helper launches of the driver script, not this library's kernels
        lines.append(f"| {nm} | {d[gi]} | " + " | ".join(d[hdr.index(c)][:12] for c in cols) + " |")
    open(os.path.join(OUT, name), "w").write("\n".join(lines) + "\n")


for rep, name in (("prof_gemm_tc.ncu-rep", "ncu_gemm_tc_summary.md"), ("prof_small.ncu-rep", "ncu_small_kernels_summary.md"),
                  ("prof_fedavg.ncu-rep", "ncu_fedavg_summary.md"), ("prof_misc.ncu-rep", "ncu_misc_kernels_summary.md"),
                  ("prof_gemm_x3.ncu-rep", "ncu_gemm_tc_fp32x3_summary.md")):
    p = os.path.join(GO, rep)
    pc = os.path.join(GO, rep.replace(".ncu-rep", "_raw.csv"))
    if os.path.exists(pc) and os.path.getsize(pc) > 100:
        summarise(pc, name)
    elif os.path.exists(p):
        summarise(p, name)

# per-launch durations of one iteration
for prec_tag in ("tf32", "fp32x3"):
  lp = os.path.join(GO, f"launches_{prec_tag}.csv")
  if os.path.exists(lp):
      lines = [l for l in open(lp) if not l.startswith("==")]
      rows = list(csv.DictReader(lines))
      with open(os.path.join(OUT, f"launches_iteration_{prec_tag}.md"), "w") as fh:
          fh.write("# Every launch of 2 training iterations + one 8192-row generation chunk (ncu gpu__time_duration, cold cache, serialised)\n\n| # | kernel | grid | block | us |\n|---|---|---|---|---|\n")
          tot = 0.0
          for i, r in enumerate(rows):
              us = float(r["Metric Value"].replace(",", "")) / 1000.0
              tot += us
              nm = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void <unnamed>::", "")
              fh.write(f"| {i} | {nm} | {r['Grid Size']} | {r['Block Size']} | {us:.1f} |\n")
          fh.write(f"\nsum = {tot:.1f} us over {len(rows)} launches\n")

def compact_sass(text):
    """cuobjdump listing without the hex encodings (instruction address + mnemonic + operands only): ~6x smaller."""
    out = []
    for line in text.splitlines():
        if re.match(r"^\s*/\* 0x[0-9a-f]+ \*/\s*$", line):
            continue
        line = re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line)
        m = re.match(r"^\s+(/\*[0-9a-f]+\*/)\s+(.*)$", line)
        out.append(f"{m.group(1)} {m.group(2).rstrip()}" if m else line.rstrip())
    return "\n".join(out) + "\n"


# SASS evidence
objs = sorted(glob.glob(os.path.join(ROOT, "fed_tgan_b200", "_native", "obj", "*.cu.o")))
hist_lines = ["# SASS mnemonic evidence per kernel (cuobjdump -sass, sm_100a)", "",
              "| object | kernel | instrs | UTCHMMA | UTMALDG | LDTM | UTCBAR | SYNCS | UCGABAR/CGA | LDG/STG | RED/ATOM | LDGMC (multimem) | STAS (st.async) |", "|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
for o in objs:
    sass = subprocess.run(["cuobjdump", "-sass", o], stdout=subprocess.PIPE, text=True).stdout
    cur, stats = None, {}
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
            cur = re.sub(r"^void ", "", cur.replace("(anonymous namespace)::", "")); cur = re.sub(r"\(.*", "", cur)
            stats[cur] = dict(n=0, mma=0, tma=0, ldtm=0, utcbar=0, syncs=0, cga=0, ldst=0, atom=0, mc=0, stas=0)
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            s = stats[cur]
            s["n"] += 1
            s["mma"] += op.startswith("UTCHMMA") or op.startswith("UTCQMMA")
            s["tma"] += op.startswith("UTMALDG") or op.startswith("UBLKCP")
            s["ldtm"] += op.startswith("LDTM")
            s["utcbar"] += op.startswith("UTCBAR")
            s["syncs"] += op.startswith("SYNCS")
            s["cga"] += op.startswith("UCGABAR") or "CGA" in op
            s["ldst"] += op.startswith("LDG") or op.startswith("STG")
            s["atom"] += op.startswith("RED") or op.startswith("ATOM")
            s["mc"] += op.startswith("LDGMC")
            s["stas"] += op.startswith("STAS")
    for k, s in stats.items():
        hist_lines.append(f"| {os.path.basename(o)} | {k[:70]} | {s['n']} | {s['mma']} | {s['tma']} | {s['ldtm']} | {s['utcbar']} | {s['syncs']} | {s['cga']} | {s['ldst']} | {s['atom']} | {s['mc']} | {s['stas']} |")
    base_o = os.path.basename(o).replace(".cu.o", "")
    if "gemm_tc" not in o:
        # every kernel of this translation unit, in full
        open(os.path.join(OUT, "sass", f"{base_o}_all_kernels.sass"), "w").write(compact_sass(sass))
    else:
        # the tcgen05 GEMM has ~100 template instantiations: keep the representative ones in full --
        # fp32x3 + fused BatchNorm epilogue (K-major), and the persistent double-buffered-TMEM kernel (bias epilogue)
        for tag, pat in (("k_gemm_tc_kmajor_bn_fp32x3", "k_gemm_tcILb0ELb0ELi6ELb1E"), ("k_gemm_tc_persist_kmajor_bias", "k_gemm_tc_persistILb0ELb0ELi2E"),
                         ("k_gemm_tc_wgrad_mnmajor_plain_fp32x3", "k_gemm_tcILb1ELb1ELi0ELb1E")):
            keep2, on2 = [], False
            for line in sass.splitlines():
                if "Function :" in line:
                    on2 = pat in line
                if on2:
                    keep2.append(line)
            if keep2:
                open(os.path.join(OUT, "sass", tag + ".sass"), "w").write(compact_sass("\n".join(keep2)))
    if "gemm_tc" in o:
        # full SASS of the forward (K-major, bias epilogue) variant
        keep, on = [], False
        for line in sass.splitlines():
            if "Function :" in line:
                on = "k_gemm_tcILb0ELb0ELi2ELb0E" in line
            if on:
                keep.append(line)
        open(os.path.join(OUT, "sass", "k_gemm_tc_kmajor_bias.sass"), "w").write(compact_sass("\n".join(keep)))
    if "fedavg" in o:
        keep, on = [], False
        for line in sass.splitlines():
            if "Function :" in line:
                on = "twoshot" in line
            if on:
                keep.append(line)
        pass   # (all fedavg kernels are in fedavg_all_kernels.sass)
open(os.path.join(OUT, "sass", "mnemonic_histogram.md"), "w").write("\n".join(hist_lines) + "\n")
for f in ("profile_step_tf32_adult.json", "profile_step_fp32x3_adult.json", "profile_step_fp32x3_wide.json", "profile_step_tf32_wide.json",
          "gemm_timeline.json", "fedavg_mgpu_2.json", "fedavg_mgpu_8.json", "fedavg_mgpu_4.json", "fedavg_mgpu_2_nvls.json",
          "fedavg_mgpu_8_nvls.json", "round_timeline_1.json", "round_timeline_2.json", "round_timeline_8.json", "spmd_quality_2.json",
          "parity_adult_native_fp32x3.json", "parity_adult_native_tf32.json", "parity_adult_ref_cpu20.json"):
    p = os.path.join(GO, f)
    if os.path.exists(p):
        open(os.path.join(OUT, f), "w").write(open(p).read())
print(sorted(os.listdir(OUT)), sorted(os.listdir(os.path.join(OUT, "sass"))))
