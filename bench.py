#!/usr/bin/env python
"""bench.py -- demodulated symbols/s of the batched LoRa PHY hot path on B200 (driver contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload sf7..sf12] [--impl reference]

A "step" is one pass of the demodulator (lora_phy_cuda_demodulate_batch: two launches -- demod_lean_kernel, one warp per
packet: exact estimator symbols, then every window through the streaming pruned transform and the Parseval test; and
demod_cert_kernel over the packets the first one flags: full margin-certified transform, exact replay of the windows that
cannot be proven) over one batch of synthetic, clean, pre-dechirped IQ already resident in HBM.
The default run (cfg 2, one GPU) also measures, in an `also` block of the same JSON line: a batch with per-packet random
gain, the raw-IQ variant with the on-chip downchirp, F_EXACT, SF8-SF11, the cfg-3 SF12 chunk, the stream detector and
one AWGN Monte-Carlo cell.
Workloads (BASELINE.json configs[1] / configs[2]):
  sf7   SF7 BW125k, 32-byte payloads (66 symbols = 8448 samples = 67.6 KB per packet), 1 Mi packets
  sf12  SF12 BW125k, 64-byte payloads (130 symbols = 4.26 MB per packet), 16 Ki packets resident
Under torchrun every rank owns its own packet range (weak scaling, no collective on the data path).
`--impl reference` times the reference's own CPU code (oracle/_ref, else the oracle port) on a
bounded sample with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    "sf7": dict(sf=7, payload=32, packets=1 << 20, name="SF7 BW125k, 1Mi packets x 32-byte payload, clean pre-dechirped IQ, demodulate-only"),
    "sf12": dict(sf=12, payload=64, packets=1 << 14, name="SF12 BW125k, 16Ki-packet resident chunk x 64-byte payload (of the 64Ki batch), clean pre-dechirped IQ, demodulate-only"),
    # not BASELINE configs: the spreading factors in between, same 32-byte framing, ~70 GB resident each (north_star: SF7-SF12)
    "sf8": dict(sf=8, payload=32, packets=1 << 19, name="SF8 BW125k, 512Ki packets x 32-byte payload, clean pre-dechirped IQ, demodulate-only"),
    "sf9": dict(sf=9, payload=32, packets=1 << 18, name="SF9 BW125k, 256Ki packets x 32-byte payload, clean pre-dechirped IQ, demodulate-only"),
    "sf10": dict(sf=10, payload=32, packets=1 << 17, name="SF10 BW125k, 128Ki packets x 32-byte payload, clean pre-dechirped IQ, demodulate-only"),
    "sf11": dict(sf=11, payload=32, packets=1 << 16, name="SF11 BW125k, 64Ki packets x 32-byte payload, clean pre-dechirped IQ, demodulate-only"),
}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                for line in out.strip().splitlines():
                    self.rows.append([c.strip() for c in line.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(int(r[1]) for r in self.rows if len(r) > 8 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) > 8 and r[2].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) > 8 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def bind_near_gpu(index):
    """Multi-rank runs: keep this rank's threads (and so the pinned host buffers it allocates from here on) on the CPUs
    NVML names as local to its GPU.  Eight ranks staging 2.2 GB each through one socket's memory is what limited the
    end-to-end leg at 4 and 8 GPUs.  Best effort: any failure leaves the affinity alone."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = None
        try:                                   # the CUDA device's own UUID: right whatever CUDA_VISIBLE_DEVICES says
            import torch
            h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(index).uuid))
        except Exception:
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",")
            tok = vis[index].strip() if index < len(vis) else ""
            h = pynvml.nvmlDeviceGetHandleByIndex(int(tok) if tok.isdigit() else index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return {"cpus": len(cpus), "first": min(cpus), "last": max(cpus)}
    except Exception:
        pass
    return None


def cpu_arm(wl, seconds_target=12.0):
    """Reference CPU implementation (lora_demodulate, one packet per thread) on a bounded sample."""
    import numpy as np
    import cpu_libs
    lib = cpu_libs.load_ref()
    kind = "reference"
    if lib is None:
        lib, kind = cpu_libs.load_oracle(), "port"
    oracle = cpu_libs.load_oracle()
    sf, L = wl["sf"], wl["payload"]
    N = 1 << sf
    nsym = 2 * L + 2
    cores = os.cpu_count() or 1
    # ~38 Msamples/s/core measured in SURVEY section 6; size the sample for ~seconds_target of work
    per_pkt = nsym * N
    n_pkt = max(cores, int(seconds_target * 30e6 * cores / per_pkt))
    n_pkt = min(n_pkt, max(cores, int(2e9 // (per_pkt * 8))))          # <= 2 GB of IQ
    rng = np.random.default_rng(0x10AA51F7)
    base = []
    down = oracle.downchirp(sf)
    for _ in range(min(n_pkt, 16)):                                       # 16 distinct packets, tiled
        pay = rng.integers(0, 256, L, dtype=np.uint8)
        iq = oracle.lora_modulate(oracle.lora_encode(pay), sf)
        base.append((iq.reshape(-1, N) * down[None, :]).astype(np.complex64).ravel())
    base = np.stack(base)
    iq = np.ascontiguousarray(np.tile(base, ((n_pkt + len(base) - 1) // len(base), 1))[:n_pkt])
    return lib, kind, cores, n_pkt, nsym, per_pkt, iq


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lib, kind, cores, n_pkt, nsym, per_pkt, iq = cpu_arm(wl, seconds_target=max(2.0, 40.0 / max(args.steps + args.warmup, 1)))
    secs = []
    for i in range(args.warmup + args.steps):
        s, _, _ = lib.demod_batch_mt(iq, n_pkt, per_pkt, wl["sf"], 1, 0, cores)
        if i >= args.warmup:
            secs.append(s)
    t = sum(secs) / len(secs)
    value = n_pkt * nsym / t
    sample = f"{n_pkt} packets ({n_pkt * per_pkt * 8 / 1e6:.0f} MB IQ) per step, {cores} threads, one packet per thread"
    print(json.dumps({
        "impl": "reference", "metric": "demod symbols/s", "value": value, "unit": "symbols/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msamples_per_s": n_pkt * per_pkt / t / 1e6,
        "config": {"workload": wl["name"], "sample": sample},
        "cpu_baseline": {"value": value, "unit": "symbols/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "symbols/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def make_batch(phy, lp, torch, sf, L, P, seed, raw=False, gains=False):
    """P packets of L random bytes: encode -> modulate (-> caller-side dechirp unless raw) on the device, in chunks.
    gains: every packet scaled by its own random factor in [0.5, 2) -- the max-abs prediction then misses every packet."""
    dev = phy.device
    N = 1 << sf
    nsym = 2 * L + 2
    spp = nsym * N
    cfg = lp.make_cfg(sf)
    g = torch.Generator(device="cpu").manual_seed(seed)
    iq = torch.empty((P, spp), dtype=torch.complex64, device=dev)
    chunk = max(1, min(P, (1 << 28) // spp))
    payload = torch.randint(0, 256, (P, L), dtype=torch.uint8, generator=g)
    for lo in range(0, P, chunk):
        hi = min(P, lo + chunk)
        sym = phy.encode_batch(payload[lo:hi].to(dev))
        phy.modulate_batch(cfg, sym, iq[lo:hi])
        if not raw:
            phy.dechirp_batch(cfg, iq[lo:hi])
        if gains:
            gv = (0.5 + 1.5 * torch.rand((hi - lo, 1), generator=g)).to(dev)
            torch.view_as_real(iq[lo:hi]).mul_(gv.unsqueeze(-1))
    torch.cuda.synchronize()
    return iq, payload


def time_demod(phy, lp, torch, cfg, iq, out, steps, warmup):
    """CUDA-event time of `steps` demodulate_batch launches on torch's current stream (the context enqueues there)."""
    for _ in range(warmup):
        phy.demodulate_batch(cfg, iq, out=out)
    torch.cuda.synchronize()
    c0 = phy.cert_stats()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in evs:
        a.record()
        phy.demodulate_batch(cfg, iq, out=out)
        b.record()
    torch.cuda.synchronize()
    c1 = phy.cert_stats()
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    return ms, {"windows": (c1[0] - c0[0]) // steps, "fallback_windows": (c1[1] - c0[1]) // steps,
                "second_estimates": (c1[2] - c0[2]) // steps, "rewalked_packets": (c1[3] - c0[3]) // steps,
                "pruned_windows": (c1[4] - c0[4]) // steps, "packets_handed_to_certified_kernel": (c1[5] - c0[5]) // steps}


def also_demod(phy, lp, torch, peak, name, sf, L, P, raw=False, gains=False, exact=False, steps=3):
    """One extra demodulator workload for the `also` block: ms per launch, fraction of the measured HBM peak."""
    import numpy as np
    N = 1 << sf
    nsym = 2 * L + 2
    iq, payload = make_batch(phy, lp, torch, sf, L, P, 0x10AA51F7 ^ (sf << 8) ^ (raw << 4) ^ gains, raw=raw, gains=gains)
    flags = (lp.F_DECHIRP if raw else 0) | (lp.F_EXACT if exact else 0)
    cfg = lp.make_cfg(sf, flags=flags)
    out = phy.alloc_demod_outputs(P, nsym - 2)
    sampler = ClockSampler(phy.device.index or 0)
    sampler.start()
    ms, cert = time_demod(phy, lp, torch, cfg, iq, out, steps, 2)
    sampler.stop_flag = True
    sampler.join(timeout=3)
    ok = bool((out["status"] == nsym - 2).all())
    if not raw:                                            # the legacy path round-trips: the batch must decode to its payload
        _, pay, _ = phy.decode_batch(out["symbols"])
        ok = ok and torch.equal(pay.cpu(), payload)
    byts = P * nsym * (8 * N + 2)
    rec = {"workload": name, "sf": sf, "packets": P, "ms": ms, "symbols_per_s": P * nsym / (ms * 1e-3),
           "msamples_per_s": P * nsym * N / (ms * 1e-3) / 1e6, "achieved_gbs": byts / (ms * 1e-3) / 1e9,
           "frac": byts / (ms * 1e-3) / 1e9 / peak, "validated": ok, "certified": None if exact else cert,
           "clocks": sampler.summary()}
    del iq, out
    torch.cuda.empty_cache()
    return rec


def also_stream(phy, lp, torch, peak, sf, gbytes=4.0, symbols=32, steps=3):
    """Streaming preamble / sync-word detector (cfg 4 shape, one SF): frames back to back in noise, every frame found."""
    N = 1 << sf
    cfg = lp.make_cfg(sf, amplitude=1.0)
    per_frame = (14 + symbols + 2) * N + N // 4
    F = max(1, (1 << 17) // per_frame)              # many short streams: the detector's parallelism is streams x N/16 threads
    S = max(1, int(gbytes * 1e9 / 8 / (F * per_frame)))
    g = torch.Generator(device=phy.device).manual_seed(sf)
    sym = torch.randint(0, N, (S * F, symbols), device=phy.device, generator=g).to(torch.uint16)
    iq = phy.modulate_frames_batch(cfg, sym, padding=2).reshape(S, F * per_frame)
    iq = torch.cat([iq, torch.zeros((S, 3 * N), dtype=iq.dtype, device=iq.device)], dim=1).contiguous()
    iq = (iq + torch.view_as_complex(torch.randn((S, iq.shape[1], 2), device=phy.device, generator=g) * 0.05)).contiguous()
    times = []
    out = None
    sampler = ClockSampler(phy.device.index or 0)
    sampler.start()
    for it in range(steps + 2):                    # two warm-ups: the result tensors alternate between two allocations
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = phy.detect_stream(sf, iq, thresh_db=-8.0, mtu=symbols, max_frames=F + 1)
        b.record()
        torch.cuda.synchronize()
        if it >= 2:
            times.append(a.elapsed_time(b))
    sampler.stop_flag = True
    sampler.join(timeout=3)
    ok = bool((out["n_frames"] == F).all())
    got = out["symbols"][:, :F, :].reshape(S * F, symbols).to(torch.int32)
    ok = ok and bool((got == sym.to(torch.int32)).all())
    ms = sum(times) / len(times)
    n = S * iq.shape[1]
    rec = {"workload": "stream detector SF%d, %d streams x %d frames back to back in noise" % (sf, S, F), "sf": sf, "ms": ms,
           "msamples_per_s": n / ms / 1e3, "achieved_gbs": n * 8 / ms / 1e6, "frac": n * 8 / ms / 1e6 / peak, "validated": ok,
           "ms_each": [round(t, 3) for t in times], "clocks": sampler.summary()}
    del iq, out, sym, got
    torch.cuda.empty_cache()
    return rec


def also_stream_mixed(phy, lp, torch, peak, gbytes=6.0, symbols=24, steps=3):
    """BASELINE configs[3] shape: ONE batch of continuous streams that differ in SF (7..12), bandwidth (the reference's
    bw_scale semantics) and oversampling, frames in the upstream format back to back with a per-stream carrier offset,
    through lora_phy_cuda_detect_streams_ex (one launch per distinct (sf, bw, osr))."""
    import math
    import numpy as np
    groups = [(sf, k, osr) for sf in range(7, 13) for (k, osr) in ((1, 1), (2, 1), (1, 2), (1, 4))]
    per_group = gbytes * 1e9 / 8 / len(groups)
    n_row = max((14 + symbols + 2 + 3) * (1 << sf) * osr + (1 << sf) * osr // 4 for sf, k, osr in groups)   # the longest frame fits a row
    g = torch.Generator(device=phy.device).manual_seed(0xC0FFEE)
    rows, sfs, bws, osrs, counts, want = [], [], [], [], [], []
    for sf, k, osr in groups:
        NN = (1 << sf) * osr
        per_frame = (14 + symbols + 2) * NN + NN // 4
        F = max(1, (n_row - 3 * NN) // per_frame)
        S = max(1, int(per_group / n_row))
        sym = torch.randint(0, 1 << sf, (S * F, symbols), device=phy.device, generator=g).to(torch.uint16)
        fr = phy.modulate_frames_batch(lp.make_cfg(sf, bw_hz=125000 * k, osr=osr, amplitude=1.0), sym, padding=2).reshape(S, F * per_frame)
        # per-stream carrier offset, phase restarting every symbol length (scripts/generate_lora_phy_vectors.py:113-117)
        cfo = (torch.rand((S, 1), device=phy.device, generator=g) - 0.5) * 0.5
        ph = 2 * math.pi * cfo * (torch.arange(fr.shape[1], device=phy.device) % NN).to(torch.float32)[None, :] / NN
        fr = fr * torch.polar(torch.ones_like(ph), ph)
        row = torch.zeros((S, n_row), dtype=torch.complex64, device=phy.device)
        row[:, :fr.shape[1]] = fr
        rows.append(row)
        sfs += [sf] * S; bws += [k] * S; osrs += [osr] * S
        counts += [fr.shape[1] + 3 * NN] * S
        want += [F] * S
    iq = torch.cat(rows)
    del rows
    iq = (iq + torch.view_as_complex(torch.randn((iq.shape[0], n_row, 2), device=phy.device, generator=g) * 0.05)).contiguous()
    d_counts = torch.tensor(counts, dtype=torch.int64).to(torch.uint32).to(phy.device)
    Fmax = max(want) + 1
    times, out = [], None
    for it in range(steps + 2):                    # two warm-ups: the result tensors alternate between two allocations
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = phy.detect_stream(sfs, iq, thresh_db=-8.0, mtu=symbols, max_frames=Fmax, sample_counts=d_counts, bw_scale=bws, osr=osrs)
        b.record()
        torch.cuda.synchronize()
        if it >= 2:
            times.append(a.elapsed_time(b))
    nf = out["n_frames"].cpu().numpy()
    want = np.array(want)
    home = (np.array(bws) == 1) & (np.array(osrs) == 1)
    ms = sum(times) / len(times)
    n = int(sum(counts))
    rec = {"workload": "cfg-4 shape: %d streams, SF7-12 x (BW125, BW250, osr 2, osr 4), frames back to back with per-stream CFO in "
                       "noise, one detect_streams_ex call (%d launches)" % (iq.shape[0], len(groups)),
           "ms": ms, "msamples_per_s": n / ms / 1e3, "achieved_gbs": n * 8 / ms / 1e6, "frac": n * 8 / ms / 1e6 / peak,
           "frames_found_frac": float(np.minimum(nf, want).sum() / want.sum()),
           "validated": bool((nf[home] == want[home]).all()), "ms_each": [round(t, 3) for t in times]}
    del iq, out
    torch.cuda.empty_cache()
    return rec


def also_awgn(phy, lp, torch, sf=7, snr_db=-6.0, packets=200000, payload=16):
    """AWGN Monte-Carlo driver (cfg 5 shape, one cell): modulate + noise + demodulate + decode + score on the device."""
    cc = lp.make_chain_cfg(sf, cr=4)
    phy.awgn_roundtrip(cc, snr_db, packets, payload, cell_id=1)      # untimed: the context's scratch grows to the cell's size here
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    counts, _ = phy.awgn_roundtrip(cc, snr_db, packets, payload, cell_id=2)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    c = counts.cpu().tolist()
    return {"workload": "AWGN Monte-Carlo cell SF%d %.0f dB, %d packets x %d B, CR4/8" % (sf, snr_db, packets, payload), "sf": sf,
            "ms": ms, "packets_per_s": packets / (ms * 1e-3), "packet_error_rate": c[1] / max(c[0], 1), "validated": c[0] == packets}


def run_ours(args, wl):
    import numpy as np
    import torch
    import torch.distributed as dist
    import lora_phy_b200 as lp

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_near_gpu(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    phy = lp.LoraPhyCuda(local_rank)

    sf, L = wl["sf"], wl["payload"]
    P = args.packets or wl["packets"]
    N = 1 << sf
    nsym = 2 * L + 2
    spp = nsym * N
    cfg = lp.make_cfg(sf)

    # ---- synthetic batch, generated on the device by our own modulator (per-rank packet range) ----
    iq, payload_all = make_batch(phy, lp, torch, sf, L, P, 0x10AA51F7 + rank)
    out = phy.alloc_demod_outputs(P, nsym - 2)

    two = args.two_kernel
    cfg_pre = lp.make_cfg(sf, flags=lp.F_ONLY_PREPASS)
    cfg_sym = lp.make_cfg(sf, flags=lp.F_ONLY_SYMBOLS)
    cfg_one = lp.make_cfg(sf, flags=lp.F_EXACT) if args.exact else cfg

    def step(ev=None):
        if not two:                            # default: TWO launches per batch (demod_lean_kernel + demod_cert_kernel over its flags)
            if ev is not None:
                ev[0].record()
            phy.demodulate_batch(cfg_one, iq, out=out)
            if ev is not None:
                ev[1].record()
        elif ev is None:
            phy.demodulate_batch(lp.make_cfg(sf, flags=lp.F_TWO_KERNEL), iq, out=out)
        else:                                  # A/B: the older two launches, events around the symbols kernel
            phy.demodulate_batch(cfg_pre, iq, out=out)
            ev[0].record()
            phy.demodulate_batch(cfg_sym, iq, out=out)
            ev[1].record()

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    # correctness guard: the batch must decode back to the payload it was made from
    _, pay, _ = phy.decode_batch(out["symbols"])
    assert torch.equal(pay.cpu(), payload_all), "bench batch did not demodulate to its payload"
    assert bool((out["status"] == nsym - 2).all())

    sampler = ClockSampler(local_rank)
    sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = phy.launch_count()
    cert0 = phy.cert_stats()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0.record()
    for i in range(args.steps):
        step(evs[i])
    t1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches = phy.launch_count() - launches0
    cert1 = phy.cert_stats()
    sampler.stop_flag = True
    sampler.join(timeout=3)
    ms_total = t0.elapsed_time(t1)
    if world > 1:
        tt = torch.tensor([ms_total], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms_total = float(tt.item())
    ms_step = ms_total / args.steps
    sym_kernel_ms = sum(a.elapsed_time(b) for a, b in evs) / args.steps

    # ---- end-to-end through the host-buffer C-ABI call (pinned host IQ in, host symbols out) ----
    # host staging is pinned memory on every rank: bound it to ~2.2 GB of IQ per rank whatever the packet size
    Pe = max(1, min(P, args.e2e_packets, int(2.3e9 // (spp * 8))))
    h_iq = torch.empty((Pe, spp), dtype=torch.complex64).pin_memory()
    h_iq.copy_(iq[:Pe])
    h_iq_np = h_iq.numpy()
    e2e_steps = max(2, min(args.steps, 5))
    rc, o = phy.demodulate_batch_host(cfg, h_iq_np)      # warm-up (allocates staging)
    assert rc == 0
    if world > 1:
        dist.barrier()
    te = time.perf_counter()
    for _ in range(e2e_steps):
        rc, o = phy.demodulate_batch_host(cfg, h_iq_np)
    te = (time.perf_counter() - te) / e2e_steps
    assert rc == 0 and np.array_equal(o["symbols"], out["symbols"][:Pe].cpu().numpy())
    te_rank = te
    if world > 1:
        tt = torch.tensor([te], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        te = float(tt.item())
    e2e_value = world * Pe * nsym / te
    h2d = Pe * spp * 8
    d2h = Pe * ((nsym - 2) * 2 + 4 + 4 + 4 + 1)
    h2d_gbs = h2d / te_rank / 1e9                           # this rank's own host -> device rate inside the call
    if world > 1:
        tt = torch.tensor([h2d_gbs], device=dev)
        allg = [torch.zeros_like(tt) for _ in range(world)]
        dist.all_gather(allg, tt)
        h2d_all = [float(x.item()) for x in allg]
    else:
        h2d_all = [h2d_gbs]

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks, peak_kind = load_peaks()
    peak = float(peaks["hbm_gbs"])
    b_sym = 8 * N + 2                                   # algorithmic bytes per demodulated symbol
    if two:
        kname = "symbols_kernel<SF%d>" % sf
        sym_bytes = P * (nsym - 2) * b_sym              # symbols kernel: data symbols of every packet, once
    else:
        kname = ("demod_cert_kernel<SF%d>, every window replayed exactly" if args.exact else "demod_lean_kernel<SF%d> (+ demod_cert_kernel over the packets it flags)") % sf
        sym_bytes = P * nsym * b_sym                    # single launch: every symbol of every packet, once
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if not two and not args.exact and os.path.exists(tpath):     # profiled, not measured in this run: per-packet DRAM bytes x packets
        t = json.load(open(tpath)).get(args.workload) or {}
        if t:
            traffic, traffic_src = t["bytes_per_packet"] * P, "profiled: " + t["source"]
    achieved = sym_bytes / (sym_kernel_ms * 1e-3) / 1e9
    step_bytes = P * nsym * b_sym
    line = {
        "metric": "demod symbols/s", "value": world * P * nsym / (ms_step * 1e-3), "unit": "symbols/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msamples_per_s": world * P * spp / (ms_step * 1e-3) / 1e6,
        "step_hbm_frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak,
        "config": {"workload": wl["name"], "packets_per_gpu": P, "symbols_per_packet": nsym, "sf": sf,
                   "batch": "homogeneous amplitude (every packet's max-abs is the same: the max prediction always hits; "
                            "also.sf7_random_gain is the batch where it never does)",
                   "l2_policy": "inputs larger than L2 (%.1f GB per GPU per step)" % (P * spp * 8 / 1e9)},
        "roofline": {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_kind": peak_kind, "kernel_ms": sym_kernel_ms, "algorithmic_bytes_per_launch": sym_bytes},
        "e2e": {"value": e2e_value, "unit": "symbols/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "packets_per_step": Pe, "h2d_gbs_per_rank": h2d_all, "rank0_cpu_affinity": numa},
        "gpu_launches": launches,
        "certified": {"windows": cert1[0] - cert0[0], "fallback_windows": cert1[1] - cert0[1],
                      "second_estimates": cert1[2] - cert0[2], "rewalked_packets": cert1[3] - cert0[3],
                      "pruned_windows": cert1[4] - cert0[4], "packets_handed_to_certified_kernel": cert1[5] - cert0[5]},
        "clocks": sampler.summary(),
    }
    if not args.no_cpu and world == 1:
        lib, kind, cores, n_pkt, nsym_c, per_pkt, ciq = cpu_arm(wl)
        lib.demod_batch_mt(ciq[:max(cores, n_pkt // 8)], max(cores, n_pkt // 8), per_pkt, sf, 1, 0, cores)   # page in / warm the threads
        reps = 3                                            # ~2 GB of IQ x 3 passes: about 20 CPU-seconds on 16 cores
        s = sum(lib.demod_batch_mt(ciq, n_pkt, per_pkt, sf, 1, 0, cores)[0] for _ in range(reps))
        line["cpu_baseline"] = {"value": reps * n_pkt * nsym_c / s, "unit": "symbols/s", "cores": cores, "kind": kind,
                                "sample": f"{reps} passes over {n_pkt} packets ({n_pkt * per_pkt * 8 / 1e6:.0f} MB IQ), one packet "
                                          f"per thread, {s:.1f} s wall = {s * cores:.0f} CPU-seconds"}
        del ciq
    else:
        line["cpu_baseline"] = None
    # ---- the other workloads north_star names, measured in the same run (single GPU only; ~16 GB resident each) ----
    if world == 1 and not args.no_also and not two and not args.exact and args.workload == "sf7" and not args.packets:
        del iq, out, h_iq
        torch.cuda.empty_cache()
        also = {}
        also["sf7_random_gain"] = also_demod(phy, lp, torch, peak, "SF7, per-packet random gain in [0.5, 2): the max-abs prediction misses "
                                             "every packet (estimator symbols redone, walk stands)", 7, 32, 1 << 18, gains=True)
        also["sf7_raw_iq"] = also_demod(phy, lp, torch, peak, "SF7 raw (un-dechirped) IQ, on-chip conjugate downchirp (F_DECHIRP, "
                                        "phy.cpp:182-243)", 7, 32, 1 << 18, raw=True)
        also["sf7_exact"] = also_demod(phy, lp, torch, peak, "SF7, F_EXACT: every window through the operation-for-operation replay "
                                       "(round-1 kernel)", 7, 32, 1 << 18, exact=True)
        for s8, p8 in ((8, 1 << 17), (9, 1 << 16), (10, 1 << 15), (11, 1 << 14)):
            also["sf%d" % s8] = also_demod(phy, lp, torch, peak, "SF%d BW125k, 32-byte payload, clean pre-dechirped IQ" % s8, s8, 32, p8)
        also["sf12_cfg3"] = also_demod(phy, lp, torch, peak, WORKLOADS["sf12"]["name"], 12, 64, 1 << 14)
        also["stream_sf7"] = also_stream(phy, lp, torch, peak, 7)
        also["stream_sf12"] = also_stream(phy, lp, torch, peak, 12)
        also["stream_cfg4_mixed"] = also_stream_mixed(phy, lp, torch, peak)
        also["awgn_sf7"] = also_awgn(phy, lp, torch)
        also["awgn_sf10"] = also_awgn(phy, lp, torch, sf=10, packets=40000)
        line["also"] = also
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="sf7", choices=sorted(WORKLOADS))
    ap.add_argument("--packets", type=int, default=0, help="packets per GPU (default: the workload's)")
    ap.add_argument("--e2e-packets", type=int, default=1 << 15)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the `also` block (SF8-12, raw IQ, random gain, stream detector, AWGN)")
    ap.add_argument("--two-kernel", action="store_true", help="A/B: time the older prepass + symbols kernel pair")
    ap.add_argument("--exact", action="store_true", help="A/B: F_EXACT, every window through the operation-for-operation replay (round-1 kernel)")
    args = ap.parse_args()
    # stdout carries the ONE JSON line and nothing else: libraries that write to fd 1 (NCCL prints its version
    # banner there under some NCCL_DEBUG settings) are sent to stderr for the rest of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(json_fd, "w", buffering=1)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
