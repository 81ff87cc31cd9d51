"""Packet-range sharding of a batch over the GPUs of one box.

Packets are independent, so rank g of G owns the contiguous range [g*P/G, (g+1)*P/G) and the
data path has NO collective: IQ never crosses NVLink.  Only the small per-packet results
(symbols, sync, cfo, time_offset, status: ~0.2 % of the input bytes) are gathered at the end,
with torch.distributed (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""


def packet_range(n_packets, rank, world):
    """Contiguous, balanced range of rank `rank`: sizes differ by at most one packet."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    lo = (n_packets * rank) // world
    hi = (n_packets * (rank + 1)) // world
    return lo, hi


def all_ranges(n_packets, world):
    return [packet_range(n_packets, r, world) for r in range(world)]


def gather_results(local, n_packets, group=None, dst=0):
    """Gather per-packet result tensors (dict name -> tensor with leading dim = local packets) onto
    rank `dst`, in global packet order.  Returns the dict of full tensors on dst, None elsewhere.
    Ragged shard sizes are padded to the largest shard for the collective and trimmed after."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return local
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    ranges = all_ranges(n_packets, world)
    biggest = max(hi - lo for lo, hi in ranges)
    out = {} if rank == dst else None
    for name in sorted(local):
        t = local[name]
        lo, hi = ranges[rank]
        if t.shape[0] != hi - lo:
            raise ValueError(f"{name}: {t.shape[0]} rows, shard owns {hi - lo}")
        pad = torch.zeros((biggest,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: hi - lo] = t
        # uint16 has no NCCL/gloo reduction type but moves fine as raw bytes
        raw = pad.contiguous().view(torch.uint8)
        bufs = [torch.empty_like(raw) for _ in range(world)]
        dist.all_gather(bufs, raw, group=group)
        if rank == dst:
            parts = []
            for r, (rlo, rhi) in enumerate(ranges):
                full = bufs[r].view(t.dtype).reshape((biggest,) + tuple(t.shape[1:]))
                parts.append(full[: rhi - rlo])
            out[name] = torch.cat(parts, dim=0)
    return out
