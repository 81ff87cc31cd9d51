"""Round-2 GPU parity additions (run with -m gpu on a B200), all through the C ABI:

* the reference-held `detection` inputs (vectors/golden/detection_tests.bin: uniform-random IQ, noise-only windows
  with close top-two bins -- SURVEY.md section 4 calls it the adversarial case) through demodulate_batch on every
  demodulator path;
* high-level demodulate() at SF10 / SF12;
* tied estimator candidates per SF: LoRaDemod.cpp:102 breaks an equal-power tie towards the lower bin,
  phy.cpp:116 keeps the first candidate -- the two rules give different cfo / time_offset on the same input;
* the upstream property tests (external/LoRa-SDR/TestCodesSx.cpp:8-159, TestLoopback.cpp:14-64) asserted on the
  CUDA coding chain itself rather than on its oracle;
* the round-1 advisor cases: a per-packet count above samples_per_packet, a batch that starts 8 bytes into an
  allocation.
"""
import ctypes
import errno
import os

import numpy as np
import pytest

from test_gpu_parity import bits, check_legacy, make_packets, to_dev

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


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


def demod_paths(lp):
    return [("default", 0), ("exact", lp.F_EXACT), ("two_kernel", lp.F_TWO_KERNEL)]


# ---- 1. golden detection inputs ---------------------------------------------------------------------------
@pytest.mark.parametrize("case", range(6))
def test_golden_detection_inputs_on_gpu(phy, lp, oracle, case):
    z = np.load(os.path.join(GOLD, "detection_inputs.npz"))
    sf = int(z[f"case{case}_meta"][0])
    iq = z[f"case{case}_iq"]
    N = 1 << sf
    # the file's window run as one packet, and every 3-symbol slice of it as its own packet (more estimator inputs)
    rows = [iq]
    for s in range(0, iq.size // N - 2):
        rows.append(np.concatenate([iq[s * N:(s + 3) * N], np.zeros(iq.size - 3 * N, np.complex64)]))
    batch = np.stack(rows)
    counts = np.array([iq.size] + [3 * N] * (len(rows) - 1), np.uint32)
    for name, flags in demod_paths(lp):
        check_legacy(phy, lp, oracle, batch, sf, 1, 0, flags=flags, counts=counts)
    # stored "expected" symbols are placeholders (0,1,2,4,...): the reference itself does not produce them
    r, s, _, _, _ = oracle.lora_demodulate(iq, sf)
    assert r == iq.size // N - 2 and not np.array_equal(s, z[f"case{case}_symbols"][:r])


def test_golden_detection_inputs_scaled_and_hann(phy, lp, oracle):
    """Same inputs above 1.0 (normalised by 1/max, LoRaDemod.cpp:69-78) and under the Hann window."""
    z = np.load(os.path.join(GOLD, "detection_inputs.npz"))
    for case in (0, 2):
        sf = int(z[f"case{case}_meta"][0])
        iq = z[f"case{case}_iq"]
        batch = np.stack([iq, (iq * np.float32(3.7)).astype(np.complex64), (iq * np.float32(0.01)).astype(np.complex64)])
        for name, flags in demod_paths(lp):
            check_legacy(phy, lp, oracle, batch, sf, 1, 0, flags=flags)
            check_legacy(phy, lp, oracle, batch, sf, 1, 1, flags=flags)
            check_legacy(phy, lp, oracle, batch, sf, 1, 0, flags=flags | lp.F_NO_SCRATCH)


# ---- 2. high-level demodulate at the large SFs ---------------------------------------------------------------
@pytest.mark.parametrize("sf,osr,bw,window", [(10, 1, 125000, 0), (10, 2, 125000, 1), (12, 1, 125000, 0), (12, 1, 500000, 1)])
def test_highlevel_demodulate_large_sf(phy, lp, oracle, sf, osr, bw, window):
    rng = np.random.default_rng(sf * 11 + osr)
    P, S = 3, 4
    rows = []
    for p in range(P):
        iq = oracle.lora_modulate(rng.integers(0, 1 << sf, S).astype(np.uint16), sf, osr, bw)
        if p % 2:
            iq = (iq + 0.3 * (rng.standard_normal(iq.size) + 1j * rng.standard_normal(iq.size))).astype(np.complex64)
        rows.append(iq)
    iq = np.stack(rows)
    for flags in (lp.F_DECHIRP, lp.F_DECHIRP | lp.F_EXACT):
        cfg = lp.make_cfg(sf, bw_hz=bw, osr=osr, window=window, flags=flags)
        out = phy.demodulate_batch(cfg, to_dev(phy, iq))
        phy.synchronize()
        o = {k: v.cpu().numpy() for k, v in out.items()}
        for p in range(P):
            r, s, sync, cfo, to = oracle.hl_demodulate(iq[p], sf, bw, osr, window)
            assert o["status"][p] == r and np.array_equal(o["symbols"][p, :r], s) and o["sync"][p] == sync
            assert o["cfo"][p].tobytes() == cfo.tobytes() and o["time_offset"][p].tobytes() == to.tobytes()


# ---- 3. tied estimator candidates ----------------------------------------------------------------------------
def tied_packet(sf, n_sym=3, osr=2):
    """Sample phase t = 0 carries an exact tone at bin 3N/4, phase t = 1 one at bin N/4 (samples in {1, j, -1, -j}):
    both phases reach the same float power, so the candidate rule decides."""
    N = 1 << sf
    i = np.arange(N)
    a = np.exp(2j * np.pi * (3 * N // 4) * i / N)
    b = np.exp(2j * np.pi * (N // 4) * i / N)
    a = (np.round(a.real) + 1j * np.round(a.imag)).astype(np.complex64)
    b = (np.round(b.real) + 1j * np.round(b.imag)).astype(np.complex64)
    pk = np.zeros(n_sym * N * osr, np.complex64)
    for s in range(n_sym):
        pk[s * N * osr + 0:(s + 1) * N * osr:osr] = a
        pk[s * N * osr + 1:(s + 1) * N * osr:osr] = b
    return pk


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_estimator_tied_candidates(phy, lp, oracle, sf):
    pk = tied_packet(sf)
    r, s, sync, cfo_l, to_l = oracle.lora_demodulate(pk, sf, 2, 0)
    cfo_h, to_h = oracle.hl_estimate_offsets(pk, sf, 2, 0)
    # the tie is real: the two rules land on different sample phases (LoRaDemod.cpp:102 vs phy.cpp:116)
    assert (float(cfo_l), float(to_l)) == (0.25, 1.0) and (float(cfo_h), float(to_h)) == (0.75, 0.0)
    batch = np.stack([pk, pk * np.float32(0.5)]).astype(np.complex64)
    check_legacy(phy, lp, oracle, batch, sf, 2, 0)
    cfg = lp.make_cfg(sf, osr=2)
    cfo, to = phy.estimate_offsets_batch(cfg, to_dev(phy, batch))
    phy.synchronize()
    for p in range(2):
        c, t = oracle.hl_estimate_offsets(batch[p], sf, 2, 0)
        assert cfo[p].cpu().numpy().tobytes() == c.tobytes() and to[p].cpu().numpy().tobytes() == t.tobytes()
    # the same through the high-level demodulator (estimate_offsets feeds its rotation and shift)
    cfg = lp.make_cfg(sf, osr=2, flags=lp.F_DECHIRP)
    out = phy.demodulate_batch(cfg, to_dev(phy, batch))
    phy.synchronize()
    for p in range(2):
        r, s, sync, c, t = oracle.hl_demodulate(batch[p], sf, 125000, 2, 0)
        assert int(out["status"][p]) == r and np.array_equal(out["symbols"][p, :r].cpu().numpy(), s)
        assert out["cfo"][p].cpu().numpy().tobytes() == c.tobytes()


@pytest.mark.parametrize("sf,osr,window", [(7, 1, 0), (7, 4, 1), (9, 2, 0), (10, 1, 1), (11, 2, 1), (12, 1, 0)])
def test_estimate_and_compensate_more_configs(phy, lp, oracle, sf, osr, window):
    rng = np.random.default_rng(sf * 13 + osr)
    N = 1 << sf
    P = 3
    iq = np.stack([(oracle.lora_modulate(rng.integers(0, N, 3).astype(np.uint16), sf, osr) +
                    0.2 * (rng.standard_normal(5 * N * osr) + 1j * rng.standard_normal(5 * N * osr))).astype(np.complex64)
                   for _ in range(P)])
    cfg = lp.make_cfg(sf, osr=osr, window=window)
    d = to_dev(phy, iq)
    cfo, to = phy.estimate_offsets_batch(cfg, d)
    phy.synchronize()
    for p in range(P):
        c, t = oracle.hl_estimate_offsets(iq[p], sf, osr, window)
        assert cfo[p].cpu().numpy().tobytes() == c.tobytes() and to[p].cpu().numpy().tobytes() == t.tobytes()
    cf_np = np.array([0.45, -0.31, 0.0], np.float32)
    to_np = np.array([-(N * osr) / 3.0, 17.5, -0.49], np.float32)
    phy.compensate_offsets_batch(cfg, d, to_dev(phy, cf_np), to_dev(phy, to_np))
    phy.synchronize()
    got = d.cpu().numpy()
    for p in range(P):
        ref = oracle.hl_compensate_offsets(iq[p], sf, osr, float(cf_np[p]), float(to_np[p]))
        assert np.array_equal(bits(got[p]), bits(ref))


# ---- 4. upstream property tests on the CUDA chain ----------------------------------------------------------------
@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_upstream_encoder_to_decoder_property(phy, lp, sf):
    """TestLoopback.cpp:14-64 (test_encoder_to_decoder): random packets through encoder -> decoder come back, for
    SF7-12 x CR 4/4..4/8.  Asserted on the CUDA kernels alone -- no oracle in the loop."""
    import torch
    rng = np.random.default_rng(sf)
    for cr in range(5):
        for L in (1, 8, 47, 128, 255):
            pay = rng.integers(0, 256, (17, L), dtype=np.uint8)
            cc = lp.make_chain_cfg(sf, cr=cr)
            sym = phy.encode_chain_batch(cc, to_dev(phy, pay))
            assert int(sym.to(torch.int32).max()) < (1 << sf)
            out, ln, fl = phy.decode_chain_batch(cc, sym)
            phy.synchronize()
            assert np.array_equal(ln.cpu().numpy(), np.full(17, L))
            assert np.array_equal(out.cpu().numpy()[:, :L], pay)


@pytest.mark.parametrize("sf", [7, 10, 12])
def test_upstream_fec_properties_through_cuda_chain(phy, lp, sf):
    """TestCodesSx.cpp:8-139 seen through the whole CUDA chain.  One flipped bit in one payload-block symbol reaches
    at most one bit of each codeword of that block: Hamming(8,4) and (7,4) correct it (:27-37, :77-84), parity (6,4)
    and (5,4) only detect it (:104-110, :131-137) -- with error_check the frame is dropped; two flipped bits in one
    Hamming(8,4) codeword are detected, not corrected (:39-50)."""
    rng = np.random.default_rng(100 + sf)
    L = 24
    pay = rng.integers(0, 256, (1, L), dtype=np.uint8)
    for cr in (4, 3, 2, 1):
        cc = lp.make_chain_cfg(sf, cr=cr)
        sym = phy.encode_chain_batch(cc, to_dev(phy, pay)).cpu().numpy()[0]
        ppm = sf
        rows = []
        for k in range(8, sym.size):
            for bit in range(ppm):
                bad = sym.copy()
                bad[k] ^= np.uint16(1 << bit)
                rows.append(bad)
        rows = np.stack(rows)
        out, ln, fl = phy.decode_chain_batch(cc, to_dev(phy, rows))
        strict = lp.make_chain_cfg(sf, cr=cr, error_check=1)
        out_s, ln_s, fl_s = phy.decode_chain_batch(strict, to_dev(phy, rows))
        phy.synchronize()
        out, ln, ln_s = out.cpu().numpy(), ln.cpu().numpy(), ln_s.cpu().numpy()
        ln_strict = ln_s
        if cr >= 3:                     # correcting codes: every single-bit error is repaired
            assert np.all(ln == L) and np.all(out[:, :L] == pay[0])
        else:                           # detecting codes: a wrong payload needs the 16-bit checksum to collide as well
            good = ln == L                  # (sx1272DataChecksum is not a true CRC: a handful of single-bit flips near the
            wrong = np.any(out[good, :L] != pay[0], axis=1)    # end of an SF7 frame do collide, in the oracle too)
            assert wrong.sum() <= 4
        # error_check drops every frame in which a codeword raised its error flag; what survives must be right
        assert np.all((ln_strict == -1) | (ln_strict == L))
        assert np.mean(ln_strict == -1) > 0.5
        assert np.all(out_s.cpu().numpy()[ln_strict == L, :L] == pay[0])
    # two flipped bits (TestCodesSx.cpp:39-50: detected, not corrected): never a wrong payload that passes the CRC
    cc = lp.make_chain_cfg(sf, cr=4)
    sym = phy.encode_chain_batch(cc, to_dev(phy, pay)).cpu().numpy()[0]
    rows = []
    for _ in range(400):
        bad = sym.copy()
        k = int(rng.integers(8, sym.size))
        b0, b1 = rng.choice(sf, 2, replace=False)
        bad[k] ^= np.uint16((1 << int(b0)) | (1 << int(b1)))
        k2 = int(rng.integers(8, sym.size))
        bad[k2] ^= np.uint16(1 << int(rng.integers(0, sf)))
        rows.append(bad)
    out, ln, fl = phy.decode_chain_batch(cc, to_dev(phy, np.stack(rows)))
    phy.synchronize()
    out, ln = out.cpu().numpy(), ln.cpu().numpy()
    assert np.all((ln == -1) | (ln == L)) and np.any(out[ln == L, :L] != pay[0], axis=1).sum() <= 2 and np.any(ln == -1)


def test_upstream_interleaver_identity_through_cuda_chain(phy, lp):
    """TestCodesSx.cpp:141-159: interleave -> deinterleave is the identity for PPM 7..12 x RDD 0..4.  Through the
    chain: every (sf, ppm, cr) encodes and decodes back (ppm < sf exercises the reduced-rate symbol shift)."""
    rng = np.random.default_rng(3)
    for sf in range(7, 13):
        for ppm in range(7, sf + 1):
            for cr in range(5):
                pay = rng.integers(0, 256, (5, 19), dtype=np.uint8)
                cc = lp.make_chain_cfg(sf, cr=cr, ppm=ppm)
                sym = phy.encode_chain_batch(cc, to_dev(phy, pay))
                out, ln, fl = phy.decode_chain_batch(cc, sym)
                phy.synchronize()
                assert np.all(ln.cpu().numpy() == 19) and np.array_equal(out.cpu().numpy()[:, :19], pay), (sf, ppm, cr)


@pytest.mark.parametrize("cr", [3, 4])
def test_upstream_loopback_property(phy, lp, cr):
    """TestLoopback.cpp:66-140 (test_loopback): SF10, CR 4/7 and 4/8, encoder -> modulator (frame format, amplitude
    1.0, padding) -> + Gaussian noise -> stream demodulator (MTU 512) -> decoder: the packets come back.  (Upstream
    adds noise of amplitude 4.0 from a Pothos source of unspecified scaling; here sigma = 1.0 per component, i.e.
    -3 dB SNR, well inside SF10's processing gain.)"""
    import torch
    sf, N = 10, 1024
    rng = np.random.default_rng(cr)
    F, L = 5, 8
    pay = rng.integers(0, 256, (F, L), dtype=np.uint8)
    cc = lp.make_chain_cfg(sf, cr=cr)
    sym = phy.encode_chain_batch(cc, to_dev(phy, pay))
    nsym = sym.shape[1]
    frames = phy.modulate_frames_batch(lp.make_cfg(sf, amplitude=1.0), sym, padding=2).cpu().numpy()
    x = np.concatenate([np.zeros(N, np.complex64)] + [f for f in frames] + [np.zeros(2 * N, np.complex64)])
    x = (x + (rng.standard_normal(x.size) + 1j * rng.standard_normal(x.size))).astype(np.complex64)
    out = phy.detect_stream(sf, to_dev(phy, x[None, :]), thresh_db=-10.0, mtu=nsym, max_frames=F + 2)
    phy.synchronize()
    assert int(out["n_frames"][0]) == F
    got = out["symbols"][0, :F, :].to(torch.uint16).contiguous()
    dec, ln, fl = phy.decode_chain_batch(cc, got)
    phy.synchronize()
    assert np.all(ln.cpu().numpy() == L) and np.array_equal(dec.cpu().numpy()[:, :L], pay)


# ---- 5. advisor cases ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("sf", [7, 10, 12])
def test_oversized_sample_count_is_clamped(phy, lp, oracle, sf):
    """A per-packet count above samples_per_packet is clamped on the device (it used to read past the row and write
    past symbol_stride; tools/sanitize.sh runs this test under memcheck); the host entry point refuses it with
    -EINVAL before copying anything."""
    rng = np.random.default_rng(sf)
    N = 1 << sf
    iq = make_packets(oracle, rng, sf, 1, 3, 4, kind="noisy")
    spp = iq.shape[1]
    counts = np.array([spp, spp + 5 * N, 0xffffffff], np.uint32)
    for name, flags in demod_paths(lp):
        cfg = lp.make_cfg(sf, flags=flags)
        res = phy.demodulate_batch(cfg, to_dev(phy, iq), sample_counts=to_dev(phy, counts))
        phy.synchronize()
        for p in range(3):
            r, s, sync, cfo, to = oracle.lora_demodulate(iq[p], sf)
            assert int(res["status"][p]) == r and np.array_equal(res["symbols"][p, :r].cpu().numpy(), s), (name, p)
    rc, _ = phy.demodulate_batch_host(lp.make_cfg(sf), iq, sample_counts=counts)
    assert rc == -errno.EINVAL


@pytest.mark.parametrize("sf", [7, 9, 11, 12])
def test_batch_starting_8_bytes_into_an_allocation(phy, lp, oracle, sf):
    """d_iq only 8-byte aligned and at the very start of readable memory's interior: odd windows are fetched without
    touching anything outside the window (the aligned bulk copy used to start 8 bytes before the first sample)."""
    import torch
    rng = np.random.default_rng(40 + sf)
    iq = make_packets(oracle, rng, sf, 1, 3, 4, kind="noisy")
    P, spp = iq.shape
    flat = torch.zeros(P * spp + 1, dtype=torch.complex64, device=phy.device)
    view = flat[1:].view(P, spp)
    view.copy_(to_dev(phy, iq))
    assert view.data_ptr() % 16 == 8
    for name, flags in demod_paths(lp):
        cfg = lp.make_cfg(sf, flags=flags)
        out = phy.demodulate_batch(cfg, view)
        phy.synchronize()
        for p in range(P):
            r, s, sync, cfo, to = oracle.lora_demodulate(iq[p], sf)
            assert int(out["status"][p]) == r and np.array_equal(out["symbols"][p, :r].cpu().numpy(), s), (name, p)
            assert out["cfo"][p].cpu().numpy().tobytes() == cfo.tobytes()
    bad = lp.make_cfg(sf)
    # a pointer that is not even 8-byte aligned cannot hold complex64 samples: refused
    rc = phy.lib.lora_phy_cuda_demodulate_batch(phy.h, __import__("ctypes").byref(bad), flat.data_ptr() + 4, spp, None, spp, 1,
                                                out["symbols"].data_ptr(), out["symbols"].shape[1], None, None, None, None)
    assert rc == -errno.EINVAL


# ---- 6. the certified path: its replay branch really runs, and every kernel family on awkward small cases ---------------
def test_certified_fallbacks_fire(phy, lp, oracle):
    """Windows whose argmax cannot be proven are replayed through the exact chain: noise-only packets (close top bins),
    exact ties ([1,0,1,0]...), all-zero and NaN windows must show up in the replay counter, clean chirps must not --
    and every packet still equals the oracle."""
    rng = np.random.default_rng(77)
    for sf in (7, 9, 12):
        N = 1 << sf
        S = 6 if sf < 12 else 4
        clean = make_packets(oracle, rng, sf, 1, 4, S - 2, kind="clean")
        c0 = phy.cert_stats()
        check_legacy(phy, lp, oracle, clean, sf, 1, 0)
        c1 = phy.cert_stats()
        assert c1[0] - c0[0] >= 4 * S and c1[1] == c0[1], "clean chirps must certify without replay"
        hard = np.zeros((5, S * N), np.complex64)
        hard[0] = np.tile(np.array([1, 0, 1, 0], np.complex64), S * N // 4)          # bins 0 and N/2 tie exactly
        hard[1] = 0                                                               # all-zero: q1 == 0
        hard[2] = clean[0]
        hard[2, N + 3] = np.complex64(complex(np.nan, 1.0))                        # one NaN window
        hard[3] = (1e-30 * clean[1]).astype(np.complex64)                         # |X|^2 underflows in the reference
        hard[4] = clean[2]
        hard[4, 2 * N:3 * N] = np.exp(2j * np.pi * (5.5 - 0.09) * np.arange(N) / N).astype(np.complex64)   # half-bin tone: two close bins
        with np.errstate(all="ignore"):
            check_legacy(phy, lp, oracle, hard, sf, 1, 0)
        c2 = phy.cert_stats()
        assert c2[1] - c1[1] >= 3 * S, (sf, c2[1] - c1[1])
    z = np.load(os.path.join(GOLD, "detection_inputs.npz"))
    c0 = phy.cert_stats()
    for case in range(6):
        sf = int(z[f"case{case}_meta"][0])
        check_legacy(phy, lp, oracle, z[f"case{case}_iq"][None, :].copy(), sf, 1, 0)
    c1 = phy.cert_stats()
    print("golden detection inputs: %d windows, %d replayed" % (c1[0] - c0[0], c1[1] - c0[1]))


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_pruned_transform_policy_over_snr_and_sync_words(phy, lp, oracle, sf):
    """The pruned transform (one bin class, accepted by the Parseval test) and its hand-over to the full transform:
    packets whose per-sample SNR steps from clean to -6 dB and back, sync words that put the reference's carrier
    correction anywhere between 0 and 1 bin (0x88: half a bin, where the top bin holds less than 9/16 of the energy and
    the pruned test must refuse), two tones in different bin classes.  Every packet equals the oracle; the counters
    show both outcomes."""
    rng = np.random.default_rng(900 + sf)
    N = 1 << sf
    S = 10 if sf <= 9 else (6 if sf < 12 else 4)
    down = oracle.gen_chirp(N, 1, N, 0.0, True, 1.0, 0.0, 1.0)[0]
    rows = []
    snrs = [None, 20.0, 10.0, 6.0, 3.0, 0.0, -6.0, None, 6.0, None]
    for snr in snrs:
        for sync in (0x12, 0x34, 0x88, 0xF1):
            syms = rng.integers(0, N, S - 2).astype(np.uint16)
            d = (oracle.lora_modulate(syms, sf, 1, 125000, 1.0, sync).reshape(-1, N) * down[None, :]).astype(np.complex64).ravel()
            if snr is not None:
                sg = np.float32(10 ** (-snr / 20) / np.sqrt(2))
                d = (d + sg * (rng.standard_normal(d.size) + 1j * rng.standard_normal(d.size))).astype(np.complex64)
            rows.append(d)
    two = rows[0].copy()                                          # a second tone of 0.8x in another class on top of window 3
    two[3 * N:4 * N] += (0.8 * np.exp(2j * np.pi * ((int(rng.integers(0, N)) | 1) ^ 6) * np.arange(N) / N)).astype(np.complex64)
    rows.append(two)
    iq = np.stack(rows)
    c0 = phy.cert_stats()
    check_legacy(phy, lp, oracle, iq, sf, 1, 0)
    c1 = phy.cert_stats()
    windows, replayed, pruned = c1[0] - c0[0], c1[1] - c0[1], c1[4] - c0[4]
    assert windows >= iq.shape[0] * S
    assert 0 < pruned < windows, (windows, pruned)
    print("sf%d: %d windows, %d settled by the pruned transform, %d replayed exactly" % (sf, windows, pruned, replayed))
    # the same batch, clean packets only: everything but the half-bin sync word is settled by the pruned transform
    clean = np.stack([r for r, k in zip(rows, [sn for sn in snrs for _ in range(4)]) if k is None])
    c0 = phy.cert_stats()
    check_legacy(phy, lp, oracle, clean, sf, 1, 0)
    c1 = phy.cert_stats()
    assert c1[4] - c0[4] >= (c1[0] - c0[0]) // 2
    # single-warp windows hand a refused window to the full transform in the same step; multi-warp windows replay it
    # exactly and back off (a bounded number of retries)
    assert c1[1] == c0[1] if sf <= 9 else c1[1] - c0[1] <= (c1[0] - c0[0]) // 4


@pytest.mark.parametrize("sf", [10, 11, 12])
def test_lean_route_hands_over_what_it_cannot_settle(phy, lp, oracle, sf):
    """SF10-12, pre-dechirped IQ: estimator pass, streaming pruned pass, certified pass over the flagged packets.  Quiet and
    loud clean packets (max-abs above 1: the estimator is redone with the true scale, the symbols stand), noisy ones
    (handed over whole), NaN / Inf samples, a packet that ends in a partial symbol, one that starts 8 bytes into a
    16-byte unit: every output equals the oracle, and the counters show all three outcomes."""
    rng = np.random.default_rng(1200 + sf)
    N = 1 << sf
    S = 5
    down = oracle.gen_chirp(N, 1, N, 0.0, True, 1.0, 0.0, 1.0)[0]

    def pkt(amp, noise=0.0, sync=0x12):
        syms = rng.integers(0, N, S - 2).astype(np.uint16)
        d = (oracle.lora_modulate(syms, sf, 1, 125000, amp, sync).reshape(-1, N) * down[None, :]).astype(np.complex64).ravel()
        if noise:
            d = (d + np.float32(noise) * (rng.standard_normal(d.size) + 1j * rng.standard_normal(d.size))).astype(np.complex64)
        return d

    rows = [pkt(0.5), pkt(1.0), pkt(1.7), pkt(33.0), pkt(0.9, 0.02), pkt(1.0, 0.9), pkt(0.7, 2.0), pkt(0.25, sync=0x88),
            pkt(1.3), pkt(0.8)]
    rows[8][2 * N + 17] = np.complex64(complex(np.nan, 0.0))
    rows[9][3 * N + 5] = np.complex64(complex(np.inf, 1.0))
    iq = np.stack(rows)
    counts = np.full(len(rows), S * N, np.uint32)
    counts[0] = S * N - 3                                          # partial last symbol: not a regular packet
    c0 = phy.cert_stats()
    with np.errstate(all="ignore"):
        check_legacy(phy, lp, oracle, iq, sf, 1, 0, counts=counts)
    c1 = phy.cert_stats()
    assert c1[4] - c0[4] >= 3 * S and c1[5] - c0[5] >= 4, (c0, c1)
    # the same rows one sample (8 bytes) into the allocation: no 16-byte alignment, the certified kernel takes them all
    big = np.zeros((len(rows), S * N + 2), np.complex64)
    flat = big.ravel()
    dev = to_dev(phy, flat)
    view = dev[1:1 + len(rows) * (S * N + 2) - 1]
    host = np.zeros(len(rows) * (S * N + 2) - 1, np.complex64)
    for p, r in enumerate(rows):
        host[p * (S * N + 2):p * (S * N + 2) + S * N] = r
    view.copy_(to_dev(phy, host))
    cfg = lp.make_cfg(sf)
    out = phy.alloc_demod_outputs(len(rows), S - 2)
    rc = phy.lib.lora_phy_cuda_demodulate_batch(phy.h, ctypes.byref(cfg), view.data_ptr(), S * N + 2, None, S * N, len(rows),
                                                out["symbols"].data_ptr(), out["symbols"].stride(0), out["sync"].data_ptr(),
                                                out["cfo"].data_ptr(), out["time_offset"].data_ptr(), out["status"].data_ptr())
    assert rc == 0
    phy.synchronize()
    with np.errstate(all="ignore"):
        for p, r in enumerate(rows):
            n, s_ref, sync, cfo, to = oracle.lora_demodulate(r, sf, 1, 0, 1)
            assert int(out["status"][p]) == n
            assert np.array_equal(out["symbols"][p, :n].cpu().numpy(), s_ref) and int(out["sync"][p]) == sync


@pytest.mark.parametrize("sf", [7, 9, 10, 12])
def test_lean_route_many_rounds_gain_changes_and_long_packets(phy, lp, oracle, sf):
    """More packets than the streaming kernel has warps, so every warp goes through several rounds and its max-abs
    prediction meets every transition (quiet -> loud -> louder -> quiet: re-estimates with the true scale); tiny batches
    (fewer packets than warps of one CTA); and a packet of 600 symbols (no staging limit on that route).  The whole batch
    equals the exact-replay arm on the GPU, a sample of it equals the oracle."""
    import torch
    rng = np.random.default_rng(3100 + sf)
    N = 1 << sf
    S = 4
    down = oracle.gen_chirp(N, 1, N, 0.0, True, 1.0, 0.0, 1.0)[0]
    props = torch.cuda.get_device_properties(0)
    n_warps = props.multi_processor_count * 3 * 8
    P = (2 * n_warps + 17) if sf <= 9 else 1200
    base = []
    for _ in range(8):
        syms = rng.integers(0, N, S - 2).astype(np.uint16)
        base.append((oracle.lora_modulate(syms, sf, 1, 125000, 1.0, 0x12).reshape(-1, N) * down[None, :]).astype(np.complex64).ravel())
    base = np.stack(base)
    gains = np.array([0.5, 3.0, 7.25, 0.9, 1.0, 2.0, 2.0, 0.25], np.float32)
    # gain class changes from round to round for every warp: packet p of round r gets gains[(p + r) % 8]
    pick = rng.integers(0, 8, P)
    g = gains[(np.arange(P) + np.arange(P) // max(1, n_warps)) % 8]
    iq = (base[pick] * g[:, None]).astype(np.complex64)
    d = to_dev(phy, iq)
    cfg, cfg_x = lp.make_cfg(sf), lp.make_cfg(sf, flags=lp.F_EXACT)
    c0 = phy.cert_stats()
    out = {k: v.clone() for k, v in phy.demodulate_batch(cfg, d).items()}
    c1 = phy.cert_stats()
    ref = phy.demodulate_batch(cfg_x, d)
    phy.synchronize()
    for k in ("symbols", "sync", "status"):
        assert torch.equal(out[k], ref[k]), k
    for k in ("cfo", "time_offset"):
        assert torch.equal(out[k].view(torch.int32), ref[k].view(torch.int32)), k
    assert c1[4] - c0[4] >= P * S * 9 // 10 and c1[2] - c0[2] > 0, (c0, c1)      # pruned windows; re-estimates happened
    for p in list(range(0, P, max(1, P // 40)))[:40]:
        n, s_ref, sync, cfo, to = oracle.lora_demodulate(iq[p], sf, 1, 0, 1)
        assert int(out["status"][p]) == n and np.array_equal(out["symbols"][p, :n].cpu().numpy(), s_ref)
        assert bits(out["cfo"][p:p + 1].cpu().numpy())[0] == bits(np.array([cfo], np.float32))[0]
    for small in (1, 3, 7):
        check_legacy(phy, lp, oracle, iq[:small].copy(), sf, 1, 0)
    if sf in (7, 10):
        L = 600
        syms = rng.integers(0, N, L - 2).astype(np.uint16)
        long_pkt = (oracle.lora_modulate(syms, sf, 1, 125000, 0.8, 0x12).reshape(-1, N) * down[None, :]).astype(np.complex64).ravel()
        check_legacy(phy, lp, oracle, np.stack([long_pkt, long_pkt[::-1].copy() * np.complex64(0.5)]), sf, 1, 0)


@pytest.mark.parametrize("sf", [7, 8, 9, 10, 11, 12])
def test_lean_route_shifted_windows(phy, lp, oracle, sf):
    """The reference's estimator turns a fractional peak position of the two sync symbols into an integer time offset and
    every window but one then starts t_off samples later (or earlier), LoRaDemod.cpp:138-150.  Packets whose sync tones
    sit a few N-ths of a bin off the grid, both signs, odd and even offsets, with and without noise: every output equals
    the oracle, and the streaming pass keeps them (shifted run through the ring, the unshifted window from global
    memory)."""
    rng = np.random.default_rng(4100 + sf)
    N = 1 << sf
    S = 9
    i = np.arange(N)
    rows, want_toff = [], []
    # the estimator's fIndex grows with the square of a small offset: 0.1 .. 0.3 bin give |t_off| = 1 .. 9 at SF7, 6 .. 76 at SF10
    for k, delta in enumerate([0.1, -0.1, 0.15, -0.15, 0.2, -0.2, 0.25, -0.25, 0.0, 0.3, -0.3, 0.12]):
        syms = rng.integers(0, N, S - 2)
        b0, b1 = 8 * (N // 128) + 3, 16 * (N // 128) + 3
        wins = [np.exp(2j * np.pi * (b0 + delta) * i / N), np.exp(2j * np.pi * (b1 + delta) * i / N)]
        wins += [np.exp(2j * np.pi * int(b) * i / N) for b in syms]
        d = np.concatenate(wins).astype(np.complex64) * np.float32(0.9 if k % 3 else 1.6)
        if k % 2:
            d = (d + np.float32(0.03) * (rng.standard_normal(d.size) + 1j * rng.standard_normal(d.size))).astype(np.complex64)
        rows.append(d)
    iq = np.stack(rows)
    got_toff = [int(round(oracle.lora_demodulate(r, sf, 1, 0, 1)[4])) for r in rows]
    assert sum(1 for t in got_toff if t != 0) >= 8, got_toff         # the construction really shifts the windows
    c0 = phy.cert_stats()
    check_legacy(phy, lp, oracle, iq, sf, 1, 0)
    c1 = phy.cert_stats()
    assert c1[5] - c0[5] <= 6, (got_toff, c0, c1)                   # (most of them) kept by the streaming pass


@pytest.mark.parametrize("sf", [7, 8, 9, 11])
def test_noisy_batches_switch_to_the_certified_kernel_without_pruning(phy, lp, oracle, sf):
    """The adaptive route: a batch in which the streaming pass hands over more than a third of its packets makes the next
    calls skip that pass -- the certified kernel alone, for SF7-9 its copy without the pruned transform.  96 noisy packets
    (with a few clean and loud ones among them), five calls in a row: every call equals the oracle, the later calls hand
    nothing over (the streaming pass did not run) and, at SF7-9, settle nothing by the pruned transform although a third
    of the batch is clean."""
    rng = np.random.default_rng(4100 + sf)
    N = 1 << sf
    S = 6 if sf <= 9 else 4
    down = oracle.gen_chirp(N, 1, N, 0.0, True, 1.0, 0.0, 1.0)[0]
    rows = []
    for p in range(96):
        syms = rng.integers(0, N, S - 2).astype(np.uint16)
        amp = 1.0 if p % 3 else float(rng.uniform(0.4, 3.0))
        d = (oracle.lora_modulate(syms, sf, 1, 125000, amp, 0x12).reshape(-1, N) * down[None, :]).astype(np.complex64).ravel()
        if p % 3 != 1:                                            # two thirds noisy: -6 .. +2 dB per sample
            sg = np.float32(amp * 10 ** (-rng.uniform(-6.0, 2.0) / 20) / np.sqrt(2))
            d = (d + sg * (rng.standard_normal(d.size) + 1j * rng.standard_normal(d.size))).astype(np.complex64)
        rows.append(d)
    iq = np.stack(rows)
    seen_alone = 0
    for call in range(5):
        c0 = phy.cert_stats()
        check_legacy(phy, lp, oracle, iq, sf, 1, 0)
        c1 = phy.cert_stats()
        windows, pruned, handed = c1[0] - c0[0], c1[4] - c0[4], c1[5] - c0[5]
        if call >= 2 and handed == 0:                             # the certified kernel alone (the verdict has arrived)
            seen_alone += 1
            assert windows >= 96 * S, (call, windows)
            if sf <= 9:
                assert pruned == 0, (call, pruned)
            else:
                assert pruned > 0, (call, pruned)
    assert seen_alone >= 2, seen_alone
    # a clean batch right after is still right (and, once the skip has run out, goes back to the streaming pass)
    clean = np.stack([r for p, r in enumerate(rows) if p % 3 == 1])
    for _ in range(8):
        check_legacy(phy, lp, oracle, clean, sf, 1, 0)


def test_every_kernel_family_small_cases():
    """tools/sanitize_driver.py (written for compute-sanitizer, which is closed on this pool) as a plain run: every kernel
    family on ragged / oversized / odd-aligned / loud / noise-only inputs must complete without a CUDA error."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "sanitize_driver.py")], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "sanitize_driver ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
