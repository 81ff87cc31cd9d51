"""BASELINE.json configs[3] ("cfg 4"): continuous IQ streams, each with its own spreading factor,
bandwidth (as the reference's bw_scale: slope and offset, so symbols read back as bw_scale*s mod N) and oversampling,
carrying frames separated by random gaps, every frame with its own carrier and timing offset.

Impairments are applied exactly as the reference's own vector script does (scripts/generate_lora_phy_vectors.py:113-124):
rotation by 2*pi*cfo_bins*(n mod N*osr)/(N*osr) -- the phase restarts every symbol length -- then an integer shift with
zero fill.  Seed 0xC0FFEE, gaps U[0.5, 4] symbol times, CFO U[-0.25, 0.25] bins, STO U[0, N*osr) samples.

Two framings: "upstream" (10 up-chirps, sync word, 2 1/4 down-chirps, data: what the streaming receiver locks on) and
"shipped" (the reference's own lora_modulate frame: 2 sync chirps + data, no preamble -- the reference has no detector
for it, so its frames are scored at their true positions through estimate_offsets / demodulate).
TEST / BENCH INPUT ONLY: expectations always come from the oracle run on the same samples."""
import numpy as np

SEED = 0xC0FFEE


def impair(frame, N_osr, cfo_bins, shift):
    n = np.arange(frame.size)
    out = frame * np.exp(2j * np.pi * cfo_bins * (n % N_osr) / N_osr)
    if shift > 0:
        out = np.concatenate([out[shift:], np.zeros(shift, out.dtype)])
    elif shift < 0:
        out = np.concatenate([np.zeros(-shift, out.dtype), out[:shift]])
    return out.astype(np.complex64)


def make_streams(oracle, configs, frames_per_stream=3, n_symbols=12, framing="upstream", noise=0.02, seed=SEED,
                 sto=True, sync=0x12):
    """configs: list of (sf, bw_k, osr), one per stream.  Returns (iq [S, n] complex64 zero padded, counts, truth) with
    truth[s] = list of dict(start, symbols, expect, cfo, sto) per frame; start = sample index of the frame's first
    sample BEFORE the shift."""
    rng = np.random.default_rng(seed)
    rows, truth = [], []
    for sf, k, osr in configs:
        N = 1 << sf
        NN = N * osr
        parts = [np.zeros(int(rng.uniform(0.5, 4.0) * NN), np.complex64)]
        pos = parts[0].size
        frames = []
        for _ in range(frames_per_stream):
            sym = rng.integers(0, N, n_symbols).astype(np.uint16)
            if framing == "upstream":
                fr = oracle.frame_modulate(sym, sf, osr, sync, 1.0, 1, k)
            else:
                fr = oracle.lora_modulate(sym, sf, osr, 125000 * k, 1.0, sync)
            cfo = float(rng.uniform(-0.25, 0.25))
            shift = int(rng.integers(0, NN)) if sto else 0
            fr = impair(fr, NN, cfo, shift)
            frames.append(dict(start=pos, symbols=sym, expect=((sym.astype(np.int64) * k) % N).astype(np.int16), cfo=cfo, sto=shift,
                               samples=fr.size))
            parts.append(fr)
            pos += fr.size
            gap = np.zeros(int(rng.uniform(0.5, 4.0) * NN), np.complex64)
            parts.append(gap)
            pos += gap.size
        parts.append(np.zeros(3 * NN, np.complex64))
        x = np.concatenate(parts)
        x = (x + noise * (rng.standard_normal(x.size) + 1j * rng.standard_normal(x.size))).astype(np.complex64)
        rows.append(x)
        truth.append(frames)
    n = max(r.size for r in rows)
    iq = np.zeros((len(rows), n), np.complex64)
    counts = np.zeros(len(rows), np.uint32)
    for s, r in enumerate(rows):
        iq[s, :r.size] = r
        counts[s] = r.size
    return iq, counts, truth


def score_upstream(configs, truth, n_frames, frame_start, frame_err, symbols, n_symbols):
    """Frames found / symbols right / carrier estimate against the generator's ground truth, per stream."""
    rows = []
    for s, (sf, k, osr) in enumerate(configs):
        N = 1 << sf
        NN = N * osr
        found = sym_ok = sym_n = 0
        err_cfo = []
        for fr in truth[s]:
            # sync word0 sits 10 symbols into the frame, the shift moves it left
            want = fr["start"] + 10 * NN - fr["sto"]
            hit = [f for f in range(int(n_frames[s])) if abs(int(frame_start[s, f]) - want) <= NN]
            if not hit:
                continue
            f = hit[0]
            found += 1
            got = symbols[s, f, :n_symbols].astype(np.int64)
            sym_ok += int((got == fr["expect"].astype(np.int64)).sum())
            sym_n += n_symbols
            err_cfo.append(abs(frame_err[s, f] / 2.0 - fr["cfo"]))
        rows.append(dict(sf=sf, bw_k=k, osr=osr, frames=len(truth[s]), found=found, symbols_right=sym_ok, symbols=sym_n,
                         mean_abs_cfo_err_bins=float(np.mean(err_cfo)) if err_cfo else None))
    return rows
