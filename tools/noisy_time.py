#!/usr/bin/env python
"""Demodulator throughput on NOISY pre-dechirped batches (8 GB resident): python tools/noisy_time.py SF SNR_DB
(LORA_PHY_NO_LEAN=1 in the environment = the certified kernel alone).  Prints ms per launch, GB/s, the share of windows
settled by the pruned transform, the share of packets the streaming pass handed over, the share replayed exactly."""
import os, sys, torch
sys.path.insert(0, "/root/repo")
import lora_phy_b200 as lp
sf = int(sys.argv[1]); snr = float(sys.argv[2])
phy = lp.LoraPhyCuda(0); dev = phy.device
N = 1 << sf; payload = 32; nsym = 2 * payload + 2; spp = nsym * N
P = int(8e9 // (spp * 8))
cfg = lp.make_cfg(sf)
g = torch.Generator(device="cpu").manual_seed(3)
iq = torch.empty((P, spp), dtype=torch.complex64, device=dev)
chunk = max(1, min(P, (1 << 27) // spp))
for lo in range(0, P, chunk):
    hi = min(P, lo + chunk)
    sym = phy.encode_batch(torch.randint(0, 256, (hi - lo, payload), dtype=torch.uint8, generator=g).to(dev))
    phy.modulate_batch(cfg, sym, iq[lo:hi]); phy.dechirp_batch(cfg, iq[lo:hi])
    sg = 10 ** (-snr / 20) / 2 ** 0.5
    iq[lo:hi] += torch.view_as_complex(torch.randn((hi - lo, spp, 2), device=dev) * sg)
out = phy.alloc_demod_outputs(P, nsym - 2)
for _ in range(2): phy.demodulate_batch(cfg, iq, out=out)
torch.cuda.synchronize()
c0 = phy.cert_stats()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3): phy.demodulate_batch(cfg, iq, out=out)
b.record(); torch.cuda.synchronize()
c1 = phy.cert_stats()
ms = a.elapsed_time(b) / 3
print("SF%d snr %+.0f dB lean=%s  %.3f ms  %.0f GB/s  pruned %.2f  flagged packets %.2f  replayed %.4f" % (sf, snr, "off" if os.environ.get("LORA_PHY_NO_LEAN") else "on", ms, P * spp * 8 / ms / 1e6,
      (c1[4] - c0[4]) / max(1, c1[0] - c0[0]), (c1[5] - c0[5]) / 3 / P, (c1[1] - c0[1]) / max(1, c1[0] - c0[0])))
