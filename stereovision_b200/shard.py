"""Multi-GPU sharding of a batch of stereo pairs: one process per GPU, rank r owns a contiguous
slice of the pairs, no collective on the data path (pairs are independent -- BASELINE.json's
sharded configuration).  torch.distributed is plumbing: NCCL on the GPU box, gloo in CPU tests.
"""


def shard_range(n_items, rank=None, world=None):
    """[start, stop) of the items rank `rank` owns; the first n_items % world ranks get one extra."""
    if rank is None or world is None:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            rank, world = dist.get_rank(), dist.get_world_size()
        else:
            rank, world = 0, 1
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(int(n_items), world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_sizes(n_items, world):
    return [shard_range(n_items, r, world)[1] - shard_range(n_items, r, world)[0] for r in range(world)]


def gather_outputs(local_out, n_items, group=None):
    """All-gather the per-rank output slices (tensor [n_local, ...]) into [n_items, ...] on every
    rank.  Only for consumers that want every disparity everywhere; the data path itself needs no
    exchange.  Ragged shards are padded to the largest shard for the collective and trimmed after."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local_out
    world = dist.get_world_size(group)
    sizes = shard_sizes(n_items, world)
    if int(local_out.shape[0]) != sizes[dist.get_rank(group)]:
        raise ValueError("local_out does not match this rank's shard")
    mx = max(sizes)
    pad = local_out
    if int(local_out.shape[0]) < mx:
        pad = torch.cat([local_out, local_out.new_zeros((mx - local_out.shape[0],) + tuple(local_out.shape[1:]))])
    # neither NCCL nor gloo moves int16: ship bytes
    raw = pad.contiguous().view(torch.uint8)
    parts = [torch.empty_like(raw) for _ in range(world)]
    dist.all_gather(parts, raw, group=group)
    return torch.cat([p.view(local_out.dtype)[:s] for p, s in zip(parts, sizes)])


def compute_sharded(matcher, left, right, gather=True):
    """matcher.compute over this rank's slice of a [B,H,W] batch that every rank holds (or can
    index); returns the local slice, or the gathered [B,H,W] result when gather=True."""
    import torch
    B = int(left.shape[0])
    a, b = shard_range(B)
    out = matcher.compute(left[a:b], right[a:b]) if b > a else None
    if out is None:
        ref = left if hasattr(left, "new_zeros") else torch.as_tensor(left)
        out = torch.zeros((0,) + tuple(left.shape[1:]), dtype=torch.int16, device=ref.device)
    elif not hasattr(out, "new_zeros"):
        out = torch.from_numpy(out)
    if gather:
        import torch.distributed as dist
        # NCCL moves device memory only: a NumPy-path result goes to this rank's GPU for the collective
        if dist.is_available() and dist.is_initialized() and dist.get_backend() == "nccl" and not out.is_cuda:
            out = out.cuda()
    return gather_outputs(out, B) if gather else out
