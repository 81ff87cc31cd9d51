"""Host replay of the stream detector's fine-tune index walk (csrc/stream_sync.cu, sync_window): the reference's
recurrence (external/LoRa-SDR/LoRaDemod.cpp:160-163 -- an int decremented by a float per sample, truncated, wrapped
once) against the kernel's parallel form: an arithmetic progression mod N*128 checked sample by sample, restarted
from the literal next value at the first sample where the check fails (up to four pieces), the literal loop otherwise.
Pure integer / float32 arithmetic, no GPU: this pins the ALGORITHM; tests/test_gpu_parity.py and test_gpu_cfg4.py pin
the kernel against the oracle."""
import numpy as np
import pytest

FINE_STEPS = 128
SEGMENTS = 4
f32 = np.float32


def ft_step(idx, d, wrap):
    """one literal step: (int)((float)idx - d), wrapped once"""
    r = int(f32(idx) - d)                      # float32 subtraction, truncation toward zero
    if r < 0:
        r += wrap
    elif r >= wrap:
        r -= wrap
    return r


def literal(idx0, d, n, wrap):
    out, idx = [], idx0
    for _ in range(n):
        out.append(idx)
        idx = ft_step(idx, d, wrap)
    return out, idx


def kernel_walk(idx0, fine_err, n):
    """what sync_window computes: (indices, index after the window, how it got there)"""
    wrap = n * FINE_STEPS
    mask = wrap - 1
    d = f32(fine_err) * f32(FINE_STEPS)
    if d == 0:
        return [idx0] * n, idx0, "constant"
    if not abs(d) <= f32(wrap // 2):
        return (*literal(idx0, d, n, wrap), "serial")

    def step_of(e):
        return e - int(f32(e) - d)

    def check(j0, e0, k):                      # first failing sample at or after j0, n when none
        for i in range(j0, n):
            ci = (e0 - (i - j0) * k) & mask
            if ((int(f32(ci) - d) - ci + k) & mask) != 0:
                return i
        return n

    idx = [0] * n
    j0, e0, k = 0, idx0, step_of(idx0)
    pieces = 1
    while True:
        bad = check(j0, e0, k)
        hi = min(bad, n - 1)
        for i in range(j0, hi + 1):
            idx[i] = (e0 - (i - j0) * k) & mask
        if bad == n:
            return idx, (e0 - (n - j0) * k) & mask, "progression" if pieces == 1 else "pieces%d" % pieces
        if pieces == SEGMENTS:
            return (*literal(idx0, d, n, wrap), "serial")
        e0 = ft_step(idx[bad], d, wrap)
        j0 = bad + 1
        k = step_of(e0)
        pieces += 1


@pytest.mark.parametrize("sf", [7, 9, 12])
def test_piecewise_progression_is_the_literal_recurrence(sf):
    rng = np.random.default_rng(7700 + sf)
    n = 1 << sf
    wrap = n * FINE_STEPS
    how = {}
    cases = 400 if sf < 12 else 60
    for case in range(cases):
        kind = case % 4
        if kind == 0:                           # tracking: a fraction of a bin
            fe = f32(rng.uniform(-0.5, 0.5))
        elif kind == 1:                         # after the quarter-chirp step: whole bins plus a fraction
            fe = f32(rng.integers(-n // 4, n // 4) + rng.uniform(-0.5, 0.5))
        elif kind == 2:                         # integers and halves (exact in float32: no truncation asymmetry)
            fe = f32(rng.integers(-64, 64) / 2.0)
        else:                                   # tiny: the step is 0 or 1 and changes with the float binade
            fe = f32(rng.uniform(-1, 1) * 2.0 ** -rng.integers(6, 14))
        idx0 = int(rng.integers(0, wrap))
        if case % 7 == 0:
            idx0 = int(rng.integers(0, 64))     # about to wrap below zero
        got, end, tag = kernel_walk(idx0, fe, n)
        want, want_end = literal(idx0, f32(fe) * f32(FINE_STEPS), n, wrap)
        assert got == want and end == want_end, (sf, case, float(fe), idx0, tag)
        how[tag.rstrip("0123456789")] = how.get(tag.rstrip("0123456789"), 0) + 1
    print(sf, how)
    assert how.get("progression", 0) > 0 and how.get("pieces", 0) > 0


def test_a_quarter_bin_offset_wraps_every_eighth_window_and_needs_two_pieces():
    """the case the pieces were built for: d = 16.5 steps per sample at SF12"""
    n = 4096
    wrap = n * FINE_STEPS
    fe = f32(0.12890625)                        # 16.5 / 128
    idx, tags = 12345, []
    for _ in range(40):
        got, end, tag = kernel_walk(idx, fe, n)
        want, want_end = literal(idx, f32(fe) * f32(FINE_STEPS), n, wrap)
        assert got == want and end == want_end
        tags.append(tag)
        idx = end
    assert "serial" not in tags
    assert 3 <= sum(t.startswith("pieces") for t in tags) <= 8, tags


def test_lean_pass_window_runs_equal_the_reference_window_bases():
    """csrc/demod_lean.cu streams a packet with a time offset as ONE contiguous run of total-1 shifted windows plus one
    unshifted window; LoRaDemod.cpp:143-150 (stream_common.cuh: shifted_base) states the bases window by window.  Pure
    integer replay of both, every |t_off| < N and the packet lengths the route accepts (count == total * N)."""
    def shifted_base(s, n, t_off, count):                   # the reference's rule
        base = s * n
        if t_off > 0:
            if base + t_off + n <= count:
                base += t_off
        elif t_off < 0:
            if -t_off <= base:
                base += t_off
        return base

    for n in (128, 512):
        for total in (2, 3, 7, 66):
            count = total * n
            for t_off in list(range(-n + 1, n, 7)) + [-1, 1, n // 2, -n // 2]:
                want = {s: shifted_base(s, n, t_off, count) for s in range(total)}
                got = {}
                if t_off == 0:
                    got = {s: s * n for s in range(total)}
                else:
                    run0 = t_off if t_off > 0 else n + t_off          # first sample of the run
                    sym0 = 1 if t_off < 0 else 0                      # symbol of its first window
                    for j in range(total - 1):
                        got[sym0 + j] = run0 + j * n
                    s_plain = total - 1 if t_off > 0 else 0           # the one unshifted window
                    got[s_plain] = s_plain * n
                assert got == want, (n, total, t_off)
                assert all(0 <= b and b + n <= count for b in got.values())
