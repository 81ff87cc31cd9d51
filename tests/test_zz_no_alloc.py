"""lora_phy_cuda_modulate_batch_static: the allocation-free modulator entry point (no counterpart in the reference;
added for callers that cannot call cudaMalloc).  Its per-chirp phase scratch is a
static device array inside the library, so every call shares it and it keeps the previous call's contents -- hence two
calls in a row, the second one smaller and with other symbols, each compared BIT FOR BIT (the tolerance of every
modulator test here: the modulator replays the reference's float operations) with lora_phy_cuda_modulate_batch.

Collected last on purpose (zz)."""
import ctypes as C
import errno

import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lp():
    import lora_phy_b200
    return lora_phy_b200


@pytest.fixture(scope="module")
def phy(lp):
    p = lp.LoraPhyCuda(0)
    yield p
    p.close()


def same_bits(torch, a, b):
    return torch.equal(torch.view_as_real(a).view(torch.int32), torch.view_as_real(b).view(torch.int32))


@pytest.mark.parametrize("sf,osr,bw", [(7, 1, 125000), (9, 2, 250000), (12, 1, 125000)])
def test_static_scratch_modulator_matches_the_allocating_one_twice_in_a_row(phy, lp, sf, osr, bw):
    import torch
    N = 1 << sf
    cfg = lp.make_cfg(sf, osr=osr, bw_hz=bw)
    g = torch.Generator(device="cpu").manual_seed(1000 + sf)
    shapes = [(257 if sf < 12 else 9, 21), (31 if sf < 12 else 3, 5)]          # second call smaller: stale phases behind it
    sym = [torch.randint(0, N, s, dtype=torch.int32, generator=g).to(torch.uint16).to(phy.device) for s in shapes]
    spp = [(s[1] + 2) * N * osr for s in shapes]
    # outputs are the caller's: allocated before the calls under test, with a sentinel so that unwritten samples show
    got = [torch.full((s[0], n + 3), complex(7.0, -7.0), dtype=torch.complex64, device=phy.device) for s, n in zip(shapes, spp)]
    phy.synchronize()
    torch.cuda.synchronize()

    phy.modulate_batch_static(cfg, sym[0], got[0])              # first call on this device: symbol address looked up
    phy.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    phy.modulate_batch_static(cfg, sym[1], got[1])              # second call: nothing may be allocated
    phy.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0

    for k in range(2):
        want = torch.full_like(got[k], complex(7.0, -7.0))
        phy.modulate_batch(cfg, sym[k], want)                   # the existing entry point (context-owned scratch)
        phy.synchronize()
        assert same_bits(torch, got[k], want)
        assert bool((got[k][:, spp[k]:] == complex(7.0, -7.0)).all())          # nothing past the packet
        assert bool((got[k][:, :spp[k]].abs() > 0.5).all())                    # and every sample of it written


def test_static_scratch_modulator_rejects_oversize_and_bad_arguments(phy, lp):
    import torch
    cfg = lp.make_cfg(7)
    fn = phy.lib.lora_phy_cuda_modulate_batch_static
    S = 30
    P = lp.STATIC_MAX_CHIRPS // (S + 2) + 1                    # one packet beyond the limit
    sym = torch.zeros((4, S), dtype=torch.uint16, device=phy.device)
    iq = torch.zeros((4, (S + 2) * 128), dtype=torch.complex64, device=phy.device)
    before = phy.launch_count()
    # row stride 0 on both sides keeps the buffers tiny: the size check must refuse before anything is launched
    rc = fn(phy.h, C.byref(cfg), sym.data_ptr(), S, S, P, iq.data_ptr(), iq.stride(0))
    assert rc == -errno.E2BIG
    assert b"LORA_PHY_CUDA_STATIC_MAX_CHIRPS" in phy.lib.lora_phy_cuda_last_error(phy.h)
    assert phy.launch_count() == before
    assert fn(phy.h, C.byref(cfg), sym.data_ptr(), S, S, 4, iq.data_ptr(), 100) == -errno.ERANGE
    assert fn(phy.h, C.byref(cfg), None, S, S, 4, iq.data_ptr(), iq.stride(0)) == -errno.EINVAL
    # exactly at the limit is accepted (SF5 keeps the output small: 32768 packets x 32 chirps x 32 samples)
    cfg5 = lp.make_cfg(5)
    P5 = lp.STATIC_MAX_CHIRPS // 32
    sym5 = torch.randint(0, 32, (P5, 30), dtype=torch.int32).to(torch.uint16).to(phy.device)
    got = torch.empty((P5, 32 * 32), dtype=torch.complex64, device=phy.device)
    phy.modulate_batch_static(cfg5, sym5, got)
    want = phy.modulate_batch(cfg5, sym5)
    phy.synchronize()
    assert same_bits(torch, got, want)
