#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`): headline counters + opcode / stall histograms.
usage: tools/ncu_summary.py file.ncu-rep [--lines N]"""
import collections, csv, io, subprocess, sys

def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))

def main():
    rep = sys.argv[1]
    raw = page(rep, "raw")
    h, units, v = raw[0], raw[1], raw[2]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum", "sm__cycles_elapsed.max",
            "lts__t_sector_hit_rate.pct", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]
    print("kernel:", v[h.index("Kernel Name")] if "Kernel Name" in h else "?")
    for i, n in enumerate(h):
        if n in want:
            print(f"  {n:70s} {v[i]:>16s} {units[i]}")
    src = page(rep, "source")
    hh = src[1]
    iS, iI, iSm = hh.index("Source"), hh.index("Instructions Executed"), hh.index("# Samples")
    stall_cols = [(k, i) for i, k in enumerate(hh) if k.startswith("stall_") and "Not Issued" not in k]
    ops, samp, stalls, tot = collections.Counter(), collections.Counter(), collections.Counter(), 0
    for r in src[2:]:
        if len(r) < len(hh):
            continue
        n = int(r[iI]); tot += n
        toks = [t for t in r[iS].split() if not t.startswith("@")]
        op = toks[0].split(".")[0].rstrip(";") if toks else "?"
        ops[op] += n; samp[op] += int(r[iSm])
        for k, i in stall_cols:
            stalls[k] += int(r[i])
    print(f"  warp instructions executed: {tot}  (static {len(src) - 2})")
    for k, n in ops.most_common(28):
        print(f"    {k:10s} {n:14d} {100.0 * n / max(tot, 1):5.1f}%  samples {samp[k]}")
    ts = sum(stalls.values())
    print("  stall samples:", ", ".join(f"{k[6:]} {100.0 * n / max(ts, 1):.1f}%" for k, n in stalls.most_common(9)))

if __name__ == "__main__":
    main()
