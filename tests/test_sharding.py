"""Multi-GPU host logic on CPU: packet-range sharding + result gather over torch.distributed with
the gloo backend, world_size 2.  The per-rank compute stand-in is the CPU oracle
(tests may use it); on the GPU box the same code path runs with NCCL and the CUDA library."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_packet_ranges_partition():
    import lora_phy_b200 as lp
    for P in (0, 1, 7, 8, 1000, 1 << 20):
        for G in (1, 2, 3, 4, 8):
            rs = lp.all_ranges(P, G)
            assert rs[0][0] == 0 and rs[-1][1] == P
            assert all(rs[i][1] == rs[i + 1][0] for i in range(G - 1))
            sizes = [hi - lo for lo, hi in rs]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        lp.packet_range(10, 2, 2)


def _worker(rank, world, port, n_packets, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import cpu_libs
    import lora_phy_b200 as lp
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle = cpu_libs.load_oracle()
        sf, n_sym = 7, 6
        lo, hi = lp.packet_range(n_packets, rank, world)
        sym = np.zeros((hi - lo, n_sym), np.uint16)
        sync = np.zeros(hi - lo, np.uint8)
        cfo = np.zeros(hi - lo, np.float32)
        status = np.zeros(hi - lo, np.int32)
        for p in range(lo, hi):                      # every rank synthesises ITS packets only
            iq = make_packet(oracle, p, sf, n_sym)
            r, s, sw, c, _ = oracle.lora_demodulate(iq, sf)
            sym[p - lo, :r], sync[p - lo], cfo[p - lo], status[p - lo] = s, sw, c, r
        local = dict(symbols=torch.from_numpy(sym), sync=torch.from_numpy(sync), cfo=torch.from_numpy(cfo),
                     status=torch.from_numpy(status))
        full = lp.gather_results(local, n_packets)
        if rank == 0:
            q.put({k: v.numpy() for k, v in full.items()})
        else:
            assert full is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def make_packet(oracle, p, sf, n_sym):
    rng = np.random.default_rng(1000 + p)
    N = 1 << sf
    syms = rng.integers(0, N, n_sym).astype(np.uint16)
    iq = oracle.lora_modulate(syms, sf)
    d = (iq.reshape(-1, N) * oracle.downchirp(sf)[None, :]).astype(np.complex64).ravel()
    return (d + 0.3 * (rng.standard_normal(d.size) + 1j * rng.standard_normal(d.size))).astype(np.complex64)


@pytest.mark.parametrize("n_packets", [9, 2])
def test_gloo_world2_shard_and_gather(oracle, n_packets):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_packets, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # unsharded run of the same packets
    for p in range(n_packets):
        r, s_, sw, c, _ = oracle.lora_demodulate(make_packet(oracle, p, 7, 6), 7)
        assert got["status"][p] == r and got["sync"][p] == sw
        assert np.array_equal(got["symbols"][p, :r], s_)
        assert got["cfo"][p].tobytes() == np.float32(c).tobytes()
