#!/usr/bin/env python
"""The reference's performance log from the GPU path (SURVEY.md section 5 "tracing"; VERDICT r1 missing #8).

tests/performance_test.cpp:59-160 times, per profile of tests/profiles.yaml, 1000 packets x 32 bytes through
lora_modulate -> caller-side dechirp -> lora_demodulate and writes logs/performance_<RUN_ID>.csv with the columns
run_id,profile,sf,N,pps,cycles_per_symbol, which scripts/compare_perf.py diffs (regression = pps down or cycles up).
This tool writes the SAME file for the batched B200 path: the same three stages per profile on a batch of packets,
timed with CUDA events; cycles_per_symbol = elapsed SM clocks (at the clock nvidia-smi reports under load) per
demodulated symbol; extra columns (symbols_per_s, msamples_per_s, demod_hbm_frac) are ignored by compare_perf.py.
    RUN_ID=b200 python tools/perf_csv.py [--packets N]    ->  logs/performance_b200.csv
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def profiles(path):
    out, cur = [], None
    for line in open(path):
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        if line.startswith("-"):
            cur = {}
            out.append(cur)
            continue
        k, v = [x.strip() for x in line.split(":", 1)]
        cur[k] = v
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--packets", type=int, default=0, help="packets per profile (default: ~4 GB of IQ)")
    ap.add_argument("--out", default="logs")
    args = ap.parse_args()
    import torch
    import lora_phy_b200 as lp
    phy = lp.LoraPhyCuda(0)
    run_id = os.environ.get("RUN_ID", "run")
    os.makedirs(args.out, exist_ok=True)
    path = os.path.join(args.out, "performance_%s.csv" % run_id)
    try:
        mhz = float(subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=clocks.max.sm", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip())
    except Exception:
        mhz = 1965.0
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    rows = ["run_id,profile,sf,N,pps,cycles_per_symbol,symbols_per_s,msamples_per_s,demod_hbm_frac"]
    for p in profiles(os.path.join(ROOT, "tests", "golden", "profiles.yaml")):
        sf, bw = int(p["sf"]), int(p["bw"])
        N = 1 << sf
        nsym = 66
        P = args.packets or max(256, int(4e9 // (nsym * N * 8)))
        cfg = lp.make_cfg(sf, bw_hz=bw)
        payload = (torch.arange(32, dtype=torch.int32) & 0xff).to(torch.uint8).repeat(P, 1).to(phy.device)   # performance_test.cpp:80
        sym = phy.encode_batch(payload)
        iq = torch.empty((P, nsym * N), dtype=torch.complex64, device=phy.device)
        out = phy.alloc_demod_outputs(P, nsym - 2)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        for it in range(3):
            ev[0].record()
            phy.modulate_batch(cfg, sym, iq)
            phy.dechirp_batch(cfg, iq)
            ev[1].record()
            phy.demodulate_batch(cfg, iq, out=out)
            ev[2].record()
            torch.cuda.synchronize()
        s_all, s_dem = ev[0].elapsed_time(ev[2]) * 1e-3, ev[1].elapsed_time(ev[2]) * 1e-3
        pps = P / s_all
        cyc = s_all * mhz * 1e6 / (P * nsym)
        frac = P * nsym * (8 * N + 2) / s_dem / 1e9 / peak
        rows.append("%s,%s,%d,%d,%.1f,%.4f,%.1f,%.1f,%.4f" % (run_id, p["name"], sf, N, pps, cyc, P * nsym / s_all, P * nsym * N / s_all / 1e6, frac))
        del iq, out, sym, payload
        torch.cuda.empty_cache()
    open(path, "w").write("\n".join(rows) + "\n")
    print("\n".join(rows))
    phy.close()


if __name__ == "__main__":
    main()
