"""BASELINE.json configs[3] (SURVEY.md 8d "cfg 4") on the GPU: mixed SF x BW x osr continuous streams with per-frame
carrier / timing offsets (generator: tests/cfg4_streams.py, impairments as scripts/generate_lora_phy_vectors.py:113-124),
through lora_phy_cuda_detect_streams_ex.  Parity = the restated upstream receiver (oracle/lora_stream_oracle.c, extended
to osr and the reference's bw_scale semantics -- this row stays "pinned by restatement": Pothos is not compilable) and
the reference's estimate_offsets / compensate_offsets / demodulate on the same windows."""
import numpy as np
import pytest

import cfg4_streams as c4
from test_gpu_parity import bits, to_dev

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def phy():
    import lora_phy_b200 as lp
    p = lp.LoraPhyCuda(0)
    yield p
    p.close()


@pytest.fixture(scope="module")
def lp():
    import lora_phy_b200
    return lora_phy_b200


CONFIGS = [(7, 1, 1), (8, 1, 2), (9, 2, 1), (7, 2, 2), (8, 4, 1), (10, 1, 4), (7, 1, 4), (9, 1, 1), (11, 1, 1), (12, 1, 1),
           (12, 2, 2), (10, 4, 2), (11, 2, 1), (7, 4, 4)]


def run_oracle(oracle, iq, counts, configs, mtu, F):
    S = len(configs)
    res = dict(n=np.zeros(S, np.int64), start=np.zeros((S, F), np.uint32), err=np.zeros((S, F), np.int32),
               nsym=np.zeros((S, F), np.int32), sym=np.zeros((S, F, mtu), np.int16), end=np.zeros(S, np.int64))
    for s, (sf, k, osr) in enumerate(configs):
        n, a, b, c, d, e = oracle.stream_demod(iq[s, :counts[s]], sf, 0x12, -8.0, mtu, F, osr, k)
        res["n"][s], res["start"][s], res["err"][s], res["nsym"][s], res["sym"][s], res["end"][s] = n, a, b, c, d, e
    return res


def compare(out, res, F):
    o = {k: v.cpu().numpy() for k, v in out.items()}
    S = res["n"].size
    for s in range(S):
        assert int(o["n_frames"][s]) == res["n"][s], (s, o["n_frames"][s], res["n"][s])
        assert int(o["end_pos"][s]) == res["end"][s], s
        f = min(int(res["n"][s]), F)
        assert np.array_equal(o["frame_start"][s, :f].astype(np.uint32), res["start"][s, :f]), s
        assert np.array_equal(o["frame_err"][s, :f], res["err"][s, :f]), s
        assert np.array_equal(o["frame_nsym"][s, :f], res["nsym"][s, :f]), s
        for k in range(f):
            m = int(res["nsym"][s, k])
            assert np.array_equal(o["symbols"][s, k, :m], res["sym"][s, k, :m]), (s, k)
    return o


@pytest.mark.parametrize("sto", [False, True])
def test_cfg4_mixed_streams_parity_and_score(phy, lp, oracle, sto):
    mtu, F = 12, 6
    iq, counts, truth = c4.make_streams(oracle, CONFIGS, frames_per_stream=3, n_symbols=mtu, sto=sto)
    sf = [c[0] for c in CONFIGS]
    bw = [c[1] for c in CONFIGS]
    osr = [c[2] for c in CONFIGS]
    out = phy.detect_stream(sf, to_dev(phy, iq), thresh_db=-8.0, mtu=mtu, max_frames=F, sample_counts=to_dev(phy, counts),
                            bw_scale=bw, osr=osr)
    phy.synchronize()
    o = compare(out, run_oracle(oracle, iq, counts, CONFIGS, mtu, F), F)
    rows = c4.score_upstream(CONFIGS, truth, o["n_frames"], o["frame_start"], o["frame_err"], o["symbols"], mtu)
    for r in rows:
        print(r)
        if r["bw_k"] == 1 and r["osr"] == 1:                      # the receiver's home ground: everything found and right
            assert r["found"] == r["frames"] and r["symbols_right"] == r["symbols"]
    assert sum(r["found"] for r in rows) >= 0.7 * sum(r["frames"] for r in rows)


@pytest.mark.parametrize("sf,osr", [(8, 2), (7, 1), (10, 1)])
def test_cfg4_fused_compensation(phy, lp, oracle, sf, osr):
    """compensate_offsets (phy.cpp:150-180) fused into the detector's window load == compensate_offsets_batch on a copy
    followed by the plain detector == the oracle's compensate + receiver, bit for bit."""
    import torch
    mtu, F = 10, 5
    configs = [(sf, 1, osr)] * 4
    iq, counts, _ = c4.make_streams(oracle, configs, frames_per_stream=2, n_symbols=mtu, sto=False, seed=c4.SEED + sf)
    n = iq.shape[1]
    N = 1 << sf
    cfo = np.array([0.2, -0.15, 0.0, 0.31], np.float32)
    to = np.array([7.4, -(N * osr) / 2.0 - 3.2, 0.0, float(n + 5)], np.float32)     # right, left, none, |offset| >= n (no shift)
    d = to_dev(phy, iq)
    fused = phy.detect_stream([sf] * 4, d, thresh_db=-8.0, mtu=mtu, max_frames=F, osr=[osr] * 4, bw_scale=[1] * 4,
                              cfo=to_dev(phy, cfo), time_offset=to_dev(phy, to))
    phy.synchronize()
    copy = d.clone()
    phy.compensate_offsets_batch(lp.make_cfg(sf, osr=osr), copy, to_dev(phy, cfo), to_dev(phy, to))
    plain = phy.detect_stream([sf] * 4, copy, thresh_db=-8.0, mtu=mtu, max_frames=F, osr=[osr] * 4, bw_scale=[1] * 4)
    phy.synchronize()
    for key in fused:
        assert torch.equal(fused[key], plain[key]), key
    comp = np.stack([oracle.hl_compensate_offsets(iq[s], sf, osr, float(cfo[s]), float(to[s])) for s in range(4)])
    assert np.array_equal(bits(copy.cpu().numpy()), bits(comp))
    compare(fused, run_oracle(oracle, comp, np.full(4, n, np.uint32), configs, mtu, F), F)


def test_cfg4_shipped_framing_offsets_on_true_windows(phy, lp, oracle):
    """Shipped framing (2 sync chirps + data): the reference has no detector for it, so every frame is cut at its true
    position and scored through estimate_offsets (phy.cpp:81-148) and the high-level demodulate (phy.cpp:182-243)."""
    configs = [(7, 1, 1), (8, 2, 2), (9, 1, 4), (10, 4, 1)]
    nsym = 6
    iq, counts, truth = c4.make_streams(oracle, configs, frames_per_stream=3, n_symbols=nsym, framing="shipped", noise=0.05)
    for s, (sf, k, osr) in enumerate(configs):
        N = 1 << sf
        spp = (nsym + 2) * N * osr
        wins = np.stack([iq[s, fr["start"]:fr["start"] + spp] for fr in truth[s]])
        cfg = lp.make_cfg(sf, bw_hz=125000 * k, osr=osr)
        cfo, to = phy.estimate_offsets_batch(cfg, to_dev(phy, wins))
        out = phy.demodulate_batch(lp.make_cfg(sf, bw_hz=125000 * k, osr=osr, flags=lp.F_DECHIRP), to_dev(phy, wins))
        phy.synchronize()
        for f in range(wins.shape[0]):
            c, t = oracle.hl_estimate_offsets(wins[f], sf, osr, 0)
            assert cfo[f].cpu().numpy().tobytes() == c.tobytes() and to[f].cpu().numpy().tobytes() == t.tobytes()
            r, sy, sync, c2, t2 = oracle.hl_demodulate(wins[f], sf, 125000 * k, osr, 0)
            assert int(out["status"][f]) == r and np.array_equal(out["symbols"][f, :r].cpu().numpy(), sy)
            assert out["cfo"][f].cpu().numpy().tobytes() == c2.tobytes()
