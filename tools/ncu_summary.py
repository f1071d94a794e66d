#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into small text files for profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv          > profiles/<name>.txt
  python tools/ncu_summary.py full gpurun_out/prof.ncu-rep [metric ...] > profiles/<name>.txt
  python tools/ncu_summary.py source gpurun_out/prof.ncu-rep [regex]     > profiles/<name>.txt
      (warp-state samples of each captured kernel by source line and by SASS opcode; needs -lineinfo and --import-source on)
"""
import collections
import csv
import re
import subprocess
import sys

FULL_METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "dram__bytes_read.sum",
    "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "sm__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sectors_srcunit_tex_op_write.sum", "smsp__inst_executed_op_ldgsts.sum",
]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    unit = ""
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("fb::<unnamed>::", "").replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        if unit == "us":
            v *= 1e3
        elif unit == "ms":
            v *= 1e6
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(a[1] for a in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)")
    print("# %d launches, total %.3f ms" % (n, tot / 1e6))
    print("%-44s %6s %12s %8s %12s" % ("kernel", "count", "total_ms", "share%", "avg_us"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-44s %6d %12.3f %8.2f %12.1f" % (k[:44], a[0], a[1] / 1e6, 100 * a[1] / tot, a[1] / a[0] / 1e3))


def source(path, pattern):
    """Per kernel in the report: where the sampled warps were (source line, stall reason, SASS opcode)."""
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True,
                         text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    kernels = []  # (name, header, body rows)
    i = 0
    cur_file = None
    while i < len(rows):
        r = rows[i]
        if len(r) == 2 and r[0] == "File Path":
            cur_file = r[1]
        elif len(r) == 2 and r[0] == "Function Name":
            kernels.append([r[1], cur_file, None, []])
        elif r and r[0] == "Line No" and kernels:
            kernels[-1][2] = r
        elif kernels and kernels[-1][2] is not None and len(r) == len(kernels[-1][2]):
            kernels[-1][3].append(r)
        i += 1
    # the page has one section per (kernel instance, source file): merge the sections of one instance
    merged = []
    for name, fpath, hdr, body in kernels:
        if merged and merged[-1][0] == name and fpath not in merged[-1][1]:
            merged[-1][1].append(fpath)
            merged[-1][2].append((fpath, hdr, body))
        else:
            merged.append([name, [fpath], [(fpath, hdr, body)]])
    for name, files, sections in merged:
        if pattern and not re.search(pattern, name):
            continue
        by_line = collections.Counter()
        by_line_stall = collections.defaultdict(collections.Counter)
        by_op = collections.Counter()
        by_op_stall = collections.defaultdict(collections.Counter)
        by_stall = collections.Counter()
        by_file = collections.Counter()
        line_src = {}
        total = 0
        for fpath, hdr, body in sections:
            short = fpath.rsplit("/", 1)[-1]
            col = {h: k for k, h in enumerate(hdr)}
            ci_samp = col.get("# Samples")
            stall_cols = [(h, k) for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
            cur_line = None
            for r in body:
                if r[0] != "":  # a source line; the SASS rows that follow carry its samples (counted once, from SASS)
                    cur_line = "%s:%s" % (short, r[0])
                    line_src[cur_line] = r[1].strip()[:80]
                    continue
                try:
                    n = int(r[ci_samp])
                except (ValueError, TypeError):
                    continue
                if not n:
                    continue
                total += n
                op = r[3].strip().split()
                op = (op[1] if op and op[0].startswith("@") and len(op) > 1 else (op[0] if op else "?"))
                by_line[cur_line] += n
                by_op[op] += n
                by_file[short] += n
                for h, k in stall_cols:
                    try:
                        v = int(r[k])
                    except ValueError:
                        v = 0
                    if v:
                        by_stall[h] += v
                        by_line_stall[cur_line][h] += v
                        by_op_stall[op][h] += v
        pct = lambda v: 100.0 * v / max(total, 1)
        print("--- %s" % name[:120])
        print("  warp-state samples %d;  by file: %s" % (total, ", ".join("%s %.1f%%" % (f, pct(v)) for f, v in by_file.most_common(6))))
        print("  by stall reason: " + ", ".join("%s %.1f%%" % (h[6:], pct(v)) for h, v in by_stall.most_common(10)))
        print("  by SASS opcode (top 16):")
        for op, n in by_op.most_common(16):
            top = ", ".join("%s %.1f" % (h[6:], pct(v)) for h, v in by_op_stall[op].most_common(4))
            print("    %-22s %5.1f%%   %s" % (op, pct(n), top))
        print("  by source line (top 24):")
        for ln, n in by_line.most_common(24):
            top = ", ".join("%s %.1f" % (h[6:], pct(v)) for h, v in by_line_stall[ln].most_common(3))
            print("    %-20s %5.1f%%   %-80s %s" % (ln, pct(n), line_src.get(ln, ""), top))


def full(path, extra):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print("# ncu --set full --clock-control none, raw page of %s" % path)
    for d in data:
        print("--- %s" % d[hdr.index("Kernel Name")][:100])
        for w in FULL_METRICS + extra:
            if w in hdr:
                print("  %-82s %20s %s" % (w, d[hdr.index(w)], units[hdr.index(w)]))


if __name__ == "__main__":
    if sys.argv[1] == "source":
        source(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
    elif sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], sys.argv[3:])
