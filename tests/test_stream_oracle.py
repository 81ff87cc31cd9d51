"""Upstream streaming receiver (external/LoRa-SDR/LoRaDemod.cpp:145-327) as restated in
oracle/lora_stream_oracle.c -- CPU only.  The Pothos block cannot be compiled here and no in-tree test runs
it, so its behaviour is pinned by the restatement; these tests pin what the upstream design
promises: frames in the upstream format (LoRaMod.cpp:141-215) are found at the right place with the right
symbols, with CFO and arbitrary start offsets, and noise alone yields nothing."""
import numpy as np
import pytest


def make_stream(oracle, rng, sf, frames, noise=0.02, lead=None):
    N = 1 << sf
    parts, truth = [], []
    pos = 0
    for k, (syms, cfo) in enumerate(frames):
        gap = int(rng.integers(N // 2, 4 * N)) if lead is None or k else lead
        parts.append(np.zeros(gap, np.complex64))
        pos += gap
        fr = oracle.stream_frame(syms, sf, cfo_bins=cfo)
        truth.append((pos + 10 * N, syms))                  # sync word0 starts 10 symbols into the frame
        parts.append(fr)
        pos += fr.size
    parts.append(np.zeros(3 * N, np.complex64))
    iq = np.concatenate(parts)
    iq = iq + noise * (rng.standard_normal(iq.size) + 1j * rng.standard_normal(iq.size))
    return iq.astype(np.complex64), truth


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_frames_are_found(oracle, sf):
    rng = np.random.default_rng(sf)
    N = 1 << sf
    frames = [(rng.integers(0, N, 12), 0.0), (rng.integers(0, N, 20), 0.25), (rng.integers(0, N, 5), -0.3)]
    iq, truth = make_stream(oracle, rng, sf, frames)
    n, start, err, nsym, sym, end = oracle.stream_demod(iq, sf, thresh=-8.0, mtu=64, max_frames=8)
    assert n == 3
    for f, (want_start, syms) in enumerate(truth):
        assert abs(int(start[f]) - want_start) <= 2, (f, start[f], want_start)     # STO tracked to the sample grid
        assert err[f] == 0
        assert nsym[f] == len(syms) + 1                                           # + the squelched window that ends it
        assert np.array_equal(sym[f, :len(syms)], syms)


def test_noise_only_and_mtu(oracle):
    rng = np.random.default_rng(99)
    noise = (0.1 * (rng.standard_normal(40000) + 1j * rng.standard_normal(40000))).astype(np.complex64)
    assert oracle.stream_demod(noise, 7, thresh=-8.0)[0] == 0
    assert oracle.stream_demod(np.zeros(5000, np.complex64), 8)[0] == 0
    # mtu ends a frame without squelch (upstream default threshold -30 dB practically never squelches)
    syms = rng.integers(0, 128, 40)
    iq, _ = make_stream(oracle, rng, 7, [(syms, 0.0)])
    n, start, err, nsym, sym, end = oracle.stream_demod(iq, 7, thresh=-30.0, mtu=16, max_frames=4)
    assert n >= 1 and nsym[0] == 16 and np.array_equal(sym[0, :16], syms[:16])


@pytest.mark.parametrize("sf", [7, 9, 12])
def test_upstream_transmitter_into_upstream_receiver(oracle, sf):
    """external/LoRa-SDR/LoRaMod.cpp:109-238 frames (restated) through the restated receiver: the preamble tracking
    locks one sample early (the transmitter's ramp leads the receiver's table by one step) and every symbol survives."""
    rng = np.random.default_rng(sf)
    N = 1 << sf
    syms = rng.integers(0, N, 24)
    fr = oracle.frame_modulate(syms, sf, ampl=0.3, padding=2)
    assert fr.size == (14 + 24 + 2) * N + N // 4
    assert np.all(fr[-2 * N:] == 0) and np.allclose(np.abs(fr[:-2 * N]), 0.3, atol=1e-6)
    lead = 3 * N + 17
    iq = np.concatenate([np.zeros(lead, np.complex64), fr, np.zeros(3 * N, np.complex64)])
    iq = (iq + 0.005 * (rng.standard_normal(iq.size) + 1j * rng.standard_normal(iq.size))).astype(np.complex64)
    n, start, err, nsym, sym, end = oracle.stream_demod(iq, sf, thresh=-8.0, mtu=64, max_frames=4)
    assert n == 1 and abs(int(start[0]) - (lead + 10 * N)) <= 1 and err[0] == 0
    assert np.array_equal(sym[0, :24], syms)
