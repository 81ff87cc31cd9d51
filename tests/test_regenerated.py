"""tests/golden/regenerated.npz (generator: tests/golden/make_regenerated.py) -- the frozen outputs of the oracle
for the vector classes the reference names without shipping usable data (modulation, awgn, hamming, interleaver,
crc) and for the upstream chain / frame / stream rows.  CPU: the oracle still reproduces them and the code tables
equal the known answers of SURVEY.md 8c.  GPU (test_gpu_parity.py::test_regenerated_golden_set): CUDA == file."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_oracle_reproduces_regenerated_set(oracle):
    import sys
    sys.path.insert(0, GOLD)
    import make_regenerated
    z = np.load(os.path.join(GOLD, "regenerated.npz"))
    fresh = make_regenerated.build(oracle)
    assert list(z["manifest"]) == list(fresh["manifest"])
    for k in z["manifest"]:
        a, b = z[k], fresh[k]
        assert a.dtype == b.dtype and a.shape == b.shape, k
        assert a.tobytes() == b.tobytes(), k


def test_code_tables_match_known_answers():
    z = np.load(os.path.join(GOLD, "regenerated.npz"))
    assert z["hamming_84_table"].tobytes().hex() == "00d172a3b465c617e8399a4b5c8d2eff"
    assert z["hamming_74_table"].tobytes().hex() == "00517223346546176839" + "1a4b5c0d2e7f"
    assert z["parity_64_table"].tobytes().hex() == "0011322334250617" + "28391a0b1c0d2e3f"
    assert z["parity_54_table"].tobytes().hex() == "0011120314050617" + "18090a1b0c1d1e0f"
    want = [0x4377, 0x301A, 0x6CFC, 0x6679, 0xC90F, 0x170B, 0x7F2B]
    assert [int(z[f"crc_{k}_value"][0]) for k in range(7)] == want
