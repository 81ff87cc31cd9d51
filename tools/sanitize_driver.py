#!/usr/bin/env python
"""Small invocation of EVERY kernel family for tools/sanitize.sh (compute-sanitizer memcheck / racecheck / synccheck).
Sizes are tiny on purpose (the tools slow kernels down 10-100x); parity is the tests' job, this only has to reach
every kernel variant: demodulator SF7-12 x {default certified path, F_EXACT, F_TWO_KERNEL}, F_DECHIRP, Hann,
ragged / oversized counts, 8-byte-offset batches, osr 2, SF2-6; detector; modulator (shipped + frame format);
dechirp; shipped coder; coding chain; stream detector (proof and serial paths); offsets; Monte-Carlo driver."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import lora_phy_b200 as lp


def main():
    phy = lp.LoraPhyCuda(0)
    dev = phy.device
    g = torch.Generator(device="cpu").manual_seed(1)
    n_kernels = 0
    for sf in range(7, 13):
        N = 1 << sf
        P, S = (5, 5) if sf <= 9 else (2, 3)
        cfg = lp.make_cfg(sf)
        sym = torch.randint(0, N, (P, S), dtype=torch.int32, generator=g).to(torch.uint16).to(dev)
        iq_raw = phy.modulate_batch(cfg, sym)
        iq = iq_raw.clone()
        phy.dechirp_batch(cfg, iq)
        noisy = iq + 0.5 * torch.randn(iq.shape, dtype=torch.complex64, device=dev)
        loud = iq * 2.5
        noise = 0.3 * torch.randn(iq.shape, dtype=torch.complex64, device=dev)
        spp = iq.shape[1]
        counts = torch.tensor([spp, spp + 7, N + 3, 0, 2 * N][:P] + [spp] * max(0, P - 5), dtype=torch.int64).to(torch.uint32).to(dev)
        flat = torch.zeros(P * spp + 1, dtype=torch.complex64, device=dev)
        odd = flat[1:].view(P, spp)
        odd.copy_(noisy)
        for flags in (0, lp.F_EXACT, lp.F_TWO_KERNEL):
            for x in (iq, noisy, loud, noise, odd):
                out = phy.demodulate_batch(lp.make_cfg(sf, flags=flags), x)
            phy.demodulate_batch(lp.make_cfg(sf, flags=flags | lp.F_NO_SCRATCH), loud)
            phy.demodulate_batch(lp.make_cfg(sf, flags=flags, window=1), noisy)
            phy.demodulate_batch(lp.make_cfg(sf, flags=flags), noisy, sample_counts=counts)
            phy.demodulate_batch(lp.make_cfg(sf, flags=flags | lp.F_DECHIRP), iq_raw)
            phy.demodulate_batch(lp.make_cfg(sf, flags=flags | lp.F_DECHIRP, window=1), iq_raw + noise)
        ok = bool((out["status"] >= 0).all())
        phy.estimate_offsets_batch(cfg, iq_raw)
        phy.compensate_offsets_batch(cfg, iq_raw.clone(), torch.full((P,), 0.2, device=dev), torch.full((P,), 3.4, device=dev))
        phy.detect_batch(sf, noise.view(-1, N)[:7].contiguous())
        # osr 2 (two-kernel generation)
        cfg2 = lp.make_cfg(sf, osr=2)
        iq2 = phy.modulate_batch(cfg2, sym[:2])
        phy.dechirp_batch(cfg2, iq2)
        phy.demodulate_batch(cfg2, iq2)
        # stream detector: frames in noise
        cc = lp.make_chain_cfg(sf, cr=4)
        pay = torch.randint(0, 256, (2, 6), dtype=torch.uint8, generator=g).to(dev)
        csym = phy.encode_chain_batch(cc, pay)
        fr = phy.modulate_frames_batch(lp.make_cfg(sf, amplitude=1.0), csym, padding=2)
        stream = torch.cat([torch.zeros(N, dtype=torch.complex64, device=dev), fr.reshape(-1),
                            torch.zeros(2 * N, dtype=torch.complex64, device=dev)])
        stream = (stream + 0.05 * torch.randn(stream.shape, dtype=torch.complex64, device=dev))[None, :].contiguous()
        for serial in ("0", "1"):
            os.environ["LORA_PHY_STREAM_SERIAL"] = serial
            o = phy.detect_stream(sf, stream, thresh_db=-8.0, mtu=csym.shape[1], max_frames=4)
        os.environ.pop("LORA_PHY_STREAM_SERIAL", None)
        got = o["symbols"][0, :2, :].to(torch.uint16).contiguous()
        phy.decode_chain_batch(cc, got)
        phy.awgn_roundtrip(cc, -5.0, 6, 9, cell_id=sf)
        phy.synchronize()
        assert ok
        n_kernels = phy.launch_count()
    for sf in range(2, 7):
        N = 1 << sf
        x = torch.randn((3, 6 * N), dtype=torch.complex64, device=dev)
        phy.demodulate_batch(lp.make_cfg(sf), x)
        phy.detect_batch(sf, x.view(-1, N)[:5].contiguous())
    # shipped coder
    pay = torch.randint(0, 256, (9, 12), dtype=torch.uint8, generator=g).to(dev)
    s = phy.encode_batch(pay)
    _, back, _ = phy.decode_batch(s)
    phy.synchronize()
    assert torch.equal(back, pay)
    # host entry points
    rc, _ = phy.demodulate_batch_host(lp.make_cfg(7), np.zeros((3, 128 * 5), np.complex64) + 0.5)
    assert rc == 0
    phy.synchronize()
    print("sanitize_driver ok: %d launches" % phy.launch_count())
    phy.close()


if __name__ == "__main__":
    main()
