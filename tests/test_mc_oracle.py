"""AWGN Monte-Carlo channel of the spot check (oracle/lora_mc_oracle.c) -- CPU only.  The reference ships no
AWGN code, so these tests pin the published pieces: Philox4x32-10 against the Random123 known answers, the
noise statistics, and the packet error behaviour at the two ends of the sweep."""
import numpy as np


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32 10 rounds
    assert oracle.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert oracle.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert oracle.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_noise_is_gaussian_and_streams_are_independent(oracle):
    n = oracle.mc_noise(0xA76E, 7 << 8 | 3, 5, 20000, 0.5)
    assert abs(n.real.std() - 0.5) < 0.01 and abs(n.imag.std() - 0.5) < 0.01
    assert abs(n.real.mean()) < 0.01 and abs(n.imag.mean()) < 0.01
    assert abs(np.mean(n.real * n.imag)) < 0.01
    kurt = np.mean(n.real ** 4) / np.mean(n.real ** 2) ** 2
    assert abs(kurt - 3.0) < 0.15
    m = oracle.mc_noise(0xA76E, 7 << 8 | 3, 6, 2000, 0.5)
    assert abs(np.corrcoef(n.real[:2000], m.real)[0, 1]) < 0.08
    a, b = oracle.mc_payload(1, 2, 3, 40), oracle.mc_payload(1, 2, 4, 40)
    assert not np.array_equal(a, b) and np.array_equal(a, oracle.mc_payload(1, 2, 3, 40))


def test_roundtrip_ends_of_the_sweep(oracle):
    cfg = oracle.chain_cfg(7, cr=4)
    ok = bad = 0
    for p in range(12):
        r = oracle.awgn_roundtrip(cfg, float(np.sqrt(10 ** (0 / 10) / 2)), 0xA76E, 7 << 8 | 10, p, 16)   # SNR 0 dB
        ok += r["length"] == 16 and np.array_equal(r["rx_payload"], r["tx_payload"])
        r = oracle.awgn_roundtrip(cfg, float(np.sqrt(10 ** (20 / 10) / 2)), 0xA76E, 7 << 8 | 0, p, 16)   # SNR -20 dB
        bad += not (r["length"] == 16 and np.array_equal(r["rx_payload"], r["tx_payload"]))
    assert ok == 12 and bad == 12
