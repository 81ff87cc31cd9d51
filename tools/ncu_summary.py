#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`): headline counters, stall ratios and the
executed-opcode histogram of every kernel in the report.
usage: tools/ncu_summary.py file.ncu-rep [--samples N]   (N = samples processed: adds a per-sample column)"""
import collections
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]


def page(rep, name, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    samples = int(sys.argv[sys.argv.index("--samples") + 1]) if "--samples" in sys.argv else 0
    raw = page(rep, "raw")
    h, units = raw[0], raw[1]
    ik = h.index("Kernel Name")
    for v in raw[2:]:
        print("kernel:", v[ik])
        for i, n in enumerate(h):
            if n in WANT:
                print(f"  {n:70s} {v[i]:>18s} {units[i]}")
        pre, suf = "smsp__average_warps_issue_stalled_", "_per_issue_active.ratio"
        st = [(float(v[i].replace(",", "")), n[len(pre):-len(suf)]) for i, n in enumerate(h)
              if n.startswith(pre) and n.endswith(suf) and "not_issued" not in n]
        print("  stalled warps per issue:", ", ".join(f"{k} {x:.2f}" for x, k in sorted(st, reverse=True)[:7]))
    src = page(rep, "source")
    ops, tot, name = collections.Counter(), 0, None

    def flush():
        if name is None:
            return
        print(f"opcodes of {name[:70]}: {tot} warp instructions" + (f", {tot * 32 / samples:.1f} thread-instr/sample" if samples else ""))
        for k, n in ops.most_common(24):
            print(f"    {k:18s} {n:14d} {100.0 * n / max(tot, 1):5.1f}%" + (f"  {n * 32 / samples:6.2f}/sample" if samples else ""))

    hh = None
    for r in src:
        if r and r[0] == "Kernel Name":
            flush()
            ops, tot, name, hh = collections.Counter(), 0, r[1], None
        elif r and r[0] == "Address":
            hh = r
            iS, iI = hh.index("Source"), hh.index("Instructions Executed")
        elif hh and len(r) >= len(hh):
            n = int(r[iI])
            toks = [t for t in r[iS].split() if not t.startswith("@")]
            ops[toks[0].rstrip(";") if toks else "?"] += n
            tot += n
    flush()


if __name__ == "__main__":
    main()
