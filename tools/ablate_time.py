#!/usr/bin/env python
"""Times lora_phy_cuda_demodulate_batch on a device-resident batch WITHOUT checking the symbols: the
measurement-only variants of tools/ablate.sh (LORA_PHY_LIB=...) produce wrong symbols by construction.
Same batch construction as bench.py (our own encoder / modulator / dechirp), CUDA events, one line out."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import lora_phy_b200 as lp

ap = argparse.ArgumentParser()
ap.add_argument("--sf", type=int, default=7)
ap.add_argument("--payload", type=int, default=32)
ap.add_argument("--gb", type=float, default=35.0, help="IQ resident per step")
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
a = ap.parse_args()
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
phy = lp.LoraPhyCuda(0)
N, nsym = 1 << a.sf, 2 * a.payload + 2
spp = nsym * N
P = int(a.gb * 1e9 // (spp * 8))
cfg = lp.make_cfg(a.sf)
iq = torch.empty((P, spp), dtype=torch.complex64, device=dev)
g = torch.Generator(device="cpu").manual_seed(7)
chunk = max(1, min(P, (1 << 28) // spp))
for lo in range(0, P, chunk):
    hi = min(P, lo + chunk)
    sym = phy.encode_batch(torch.randint(0, 256, (hi - lo, a.payload), dtype=torch.uint8, generator=g).to(dev))
    phy.modulate_batch(cfg, sym, iq[lo:hi])
    phy.dechirp_batch(cfg, iq[lo:hi])
out = phy.alloc_demod_outputs(P, nsym - 2)
for _ in range(a.warmup):
    phy.demodulate_batch(cfg, iq, out=out)
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(a.steps):
    phy.demodulate_batch(cfg, iq, out=out)
t1.record()
torch.cuda.synchronize()
ms = t0.elapsed_time(t1) / a.steps
gbs = P * nsym * (8 * N + 2) / ms / 1e6
print(f"{os.path.basename(lp.LIB_PATH)} SF{a.sf} P={P} {ms:.3f} ms/step {gbs:.0f} GB/s algorithmic")
