#!/usr/bin/env python
"""One AWGN Monte-Carlo cell (SF, SNR, packets) -- the command the ncu launch list of the Monte-Carlo driver is taken on."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lora_phy_b200 as lp
sf = int(sys.argv[1]) if len(sys.argv) > 1 else 7
n = int(sys.argv[2]) if len(sys.argv) > 2 else 200000
phy = lp.LoraPhyCuda(0)
cc = lp.make_chain_cfg(sf, cr=4)
phy.awgn_roundtrip(cc, -6.0, 2048, 16, cell_id=1)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
c, _ = phy.awgn_roundtrip(cc, -6.0, n, 16, cell_id=2)
b.record()
torch.cuda.synchronize()
print("sf%d %d packets %.2f ms %.0f packets/s PER %.4f" % (sf, n, a.elapsed_time(b), n / a.elapsed_time(b) * 1e3, c[1].item() / max(c[0].item(), 1)))
