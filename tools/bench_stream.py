#!/usr/bin/env python
"""Throughput of the streaming preamble / sync-word detector (BASELINE.json configs[3]; K2, csrc/stream_sync.cu).
Streams are built on the device: frames in the upstream format (modulate_frames_batch) laid back to back -- the
padding symbols of each frame are the gap to the next -- plus complex Gaussian noise.  Timed with CUDA events on
the launching stream; every frame must be found (n_frames == frames per stream) or the number is not reported."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sf", default="7,8,9,10,11,12")
    ap.add_argument("--samples-per-stream", type=int, default=1 << 21)
    ap.add_argument("--bytes", type=float, default=8e9, help="IQ bytes per spreading factor")
    ap.add_argument("--symbols", type=int, default=32, help="data symbols per frame")
    ap.add_argument("--noise", type=float, default=0.05)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import torch
    import lora_phy_b200 as lp
    phy = lp.LoraPhyCuda(0)
    phy.set_stream(torch.cuda.current_stream().cuda_stream)
    for sf in [int(s) for s in args.sf.split(",")]:
        N = 1 << sf
        cfg = lp.make_cfg(sf, amplitude=1.0)
        per_frame = (14 + args.symbols + 2) * N + N // 4
        F = max(1, args.samples_per_stream // per_frame)
        S = max(1, int(args.bytes / 8 / (F * per_frame)))
        g = torch.Generator(device=phy.device).manual_seed(sf)
        sym = torch.randint(0, N, (S * F, args.symbols), device=phy.device, generator=g).to(torch.uint16)
        iq = phy.modulate_frames_batch(cfg, sym, padding=2).reshape(S, F * per_frame)
        iq = torch.cat([iq, torch.zeros((S, 3 * N), dtype=iq.dtype, device=iq.device)], dim=1).contiguous()
        noise = torch.randn((S, iq.shape[1], 2), device=phy.device, generator=g) * args.noise
        iq = (iq + torch.view_as_complex(noise)).contiguous()
        del noise
        out = None
        times = []
        for it in range(args.steps + 1):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = phy.detect_stream(sf, iq, thresh_db=-8.0, mtu=args.symbols, max_frames=F + 1)
            b.record()
            torch.cuda.synchronize()
            if it:
                times.append(a.elapsed_time(b))
        ok = bool((out["n_frames"] == F).all())
        got = out["symbols"][:, :F, :].reshape(S * F, args.symbols).to(torch.int32)
        sym_ok = float((got == sym.to(torch.int32)).float().mean())
        ms = sum(times) / len(times)
        n_samples = S * iq.shape[1]
        print(json.dumps({"metric": "stream_detect_msamples_per_s", "sf": sf, "value": n_samples / ms / 1e3, "unit": "Msamples/s",
                          "ms": ms, "streams": S, "frames_per_stream": F, "samples_per_stream": iq.shape[1],
                          "all_frames_found": ok, "symbol_match": sym_ok, "iq_gbytes": n_samples * 8 / 1e9,
                          "hbm_gbps": n_samples * 8 / ms / 1e6}), flush=True)
        del iq, out, sym, got
        torch.cuda.empty_cache()
    phy.close()


if __name__ == "__main__":
    main()
