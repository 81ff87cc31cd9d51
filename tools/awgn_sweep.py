#!/usr/bin/env python
"""AWGN Monte-Carlo PER/BER sweep (BASELINE.json configs[4]): SNR grid x SF7..12, every cell's packets sharded over
the ranks by contiguous packet range (shard.packet_range) -- no collective on the data path, one 64-byte counter
all_reduce per cell at the end.  Single GPU: `python tools/awgn_sweep.py`; N GPUs:
`python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/awgn_sweep.py`.
Writes a CSV (one row per cell) and prints one JSON line with the packet throughput (CUDA events, max over ranks).
The reference's TEST_PLAN.md names a tests/awgn_sweep.py that is not shipped; this is the device-side equivalent."""
import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

COLS = ("packets", "packet_errors", "dropped", "bit_errors", "bits", "symbol_errors", "symbols", "sync_errors")


def cells(args):
    sfs = [int(s) for s in args.sf.split(",")]
    n = int(round((args.snr_stop - args.snr_start) / args.snr_step)) + 1
    return [(sf, i, args.snr_start + i * args.snr_step) for sf in sfs for i in range(n)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--packets", type=int, default=20000, help="packets per (SF, SNR) cell, whole job")
    ap.add_argument("--payload", type=int, default=16)
    ap.add_argument("--cr", type=int, default=4, help="redundancy bits 0..4 (coding rate 4/4..4/8)")
    ap.add_argument("--sf", default="7,8,9,10,11,12")
    ap.add_argument("--snr-start", type=float, default=-20.0)
    ap.add_argument("--snr-stop", type=float, default=0.0)
    ap.add_argument("--snr-step", type=float, default=2.0)
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0xA76E)
    ap.add_argument("--out", default="gpurun_out/awgn_sweep.csv")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    import lora_phy_b200 as lp

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    phy = lp.LoraPhyCuda(local)
    phy.set_stream(torch.cuda.current_stream().cuda_stream)      # the events below must see the driver's launches
    lo, hi = lp.packet_range(args.packets, rank, world)
    grid = cells(args)
    totals = torch.zeros((len(grid), 8), dtype=torch.int64, device=phy.device)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # warm-up: builds tables / scratch outside the timed region
    phy.awgn_roundtrip(lp.make_chain_cfg(grid[0][0], cr=args.cr), 0.0, min(64, max(hi - lo, 1)), args.payload)
    phy.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = phy.launch_count()
    start.record(torch.cuda.current_stream())
    for k, (sf, idx, snr) in enumerate(grid):
        if hi > lo:
            phy.awgn_roundtrip(lp.make_chain_cfg(sf, cr=args.cr), snr, hi - lo, args.payload, seed=args.seed,
                               cell_id=sf << 8 | idx, first_packet=lo, counts=totals[k])
    stop.record(torch.cuda.current_stream())
    torch.cuda.synchronize()
    ms = torch.tensor([start.elapsed_time(stop)], device=phy.device)
    launches = phy.launch_count() - l0
    if world > 1:
        dist.all_reduce(totals)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        t = totals.cpu().numpy()
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            f.write("sf,snr_db,cr,payload_bytes," + ",".join(COLS) + ",per,ber,ser\n")
            for (sf, idx, snr), c in zip(grid, t):
                per = c[1] / max(c[0], 1)
                ber = c[3] / max(c[4], 1)
                ser = c[5] / max(c[6], 1)
                f.write(f"{sf},{snr:g},{args.cr},{args.payload}," + ",".join(str(int(v)) for v in c) +
                        f",{per:.6g},{ber:.6g},{ser:.6g}\n")
        total_packets = int(t[:, 0].sum())
        print(json.dumps({"metric": "awgn_roundtrip_packets_per_s", "value": total_packets / (ms.item() * 1e-3),
                          "unit": "packets/s", "n_gpus": world, "cells": len(grid), "packets_per_cell": args.packets,
                          "total_packets": total_packets, "ms": ms.item(), "gpu_launches": int(launches),
                          "config": {"workload": "cfg5 awgn sweep", "sf": args.sf, "cr": args.cr,
                                     "payload_bytes": args.payload, "snr_db": [args.snr_start, args.snr_stop, args.snr_step]},
                          "csv": args.out}))
    phy.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
