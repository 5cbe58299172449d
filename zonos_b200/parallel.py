"""Batch sharding of independent utterances across the GPUs of one box.

The hot path has no cross-utterance op, so ranks run complete replicas on contiguous blocks of the global utterance
list and nothing is exchanged while generating.  Only the results are gathered (lengths first, then padded codes as
int32), over whatever torch.distributed backend is initialised (NCCL on the box, gloo in CPU tests).
"""
import torch
import torch.distributed as dist


def shard_bounds(n_items: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`; the first n_items % world ranks get one extra item."""
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_codes(local_codes: list, n_items: int, dst: int = 0, device=None):
    """local_codes: this rank's list of int64 [Q, T_i] (its shard, in global order).  Returns the full list of n_items
    tensors (on CPU) on rank `dst`, None elsewhere.  Two collectives: all_gather of lengths, gather of padded int32 codes."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if device is None:
        device = local_codes[0].device if local_codes else torch.device("cpu")
    Q = local_codes[0].shape[0] if local_codes else 9
    per_rank = max(shard_bounds(n_items, world, r)[1] - shard_bounds(n_items, world, r)[0] for r in range(world))
    lens = torch.zeros(per_rank, dtype=torch.int32, device=device)
    for i, c in enumerate(local_codes):
        lens[i] = c.shape[1]
    all_lens = [torch.zeros_like(lens) for _ in range(world)]
    dist.all_gather(all_lens, lens)
    t_max = max(1, int(torch.stack(all_lens).max()))
    pad = torch.zeros(per_rank, Q, t_max, dtype=torch.int32, device=device)
    for i, c in enumerate(local_codes):
        pad[i, :, : c.shape[1]] = c.to(torch.int32)
    bufs = [torch.zeros_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, bufs, dst=dst)
    if rank != dst:
        return None
    out = []
    for r in range(world):
        lo, hi = shard_bounds(n_items, world, r)
        for i in range(hi - lo):
            out.append(bufs[r][i, :, : int(all_lens[r][i])].to(torch.int64).cpu())
    return out
