"""Multi-GPU use of the enhance() path: independent utterances are dealt to ranks (one
process per GPU), each rank enhances its shard in exact-length batches, and the results are
gathered on the host of rank 0.  There is no exchange step in the algorithm, hence no
data-path collective; the reference itself is single-device
(open_universe/bin/enhance.py:117-139)."""
from collections import defaultdict

import torch


def shard_bounds(n_items, rank, world):
    """Contiguous, balanced [lo, hi) slice of `n_items` for `rank`."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def length_buckets(lengths, indices, batch_size):
    """Batches of utterances with IDENTICAL length (padding and normalisation statistics depend
    on the length, reference universe.py:219-223,272, so mixed-length batches would change results)."""
    by_len = defaultdict(list)
    for i in indices:
        by_len[int(lengths[i])].append(i)
    batches = []
    for length in sorted(by_len):
        idx = by_len[length]
        for k in range(0, len(idx), batch_size):
            batches.append(idx[k : k + batch_size])
    return batches


def ragged_batches(lengths, indices, batch_size, quantum=8000, max_waste=0.25):
    """Batches of utterances of DIFFERENT lengths for enhance(lengths=...): the kernels give every row its own
    padding / statistics / frames (csrc RowGeo), so mixing lengths does not change results; what it costs is the
    work spent on the rows' tails.  Utterances are taken in order of length; a batch is closed when it is full or
    when its shortest member would waste more than `max_waste` of the batch length.  Returns (indices, T) pairs with
    T = the longest member rounded up to `quantum` samples (so that a corpus maps to a few plan shapes)."""
    order = sorted(indices, key=lambda i: (int(lengths[i]), i), reverse=True)
    batches, cur = [], []
    for i in order:
        if cur and (len(cur) >= batch_size or int(lengths[i]) < (1.0 - max_waste) * int(lengths[cur[0]])):
            batches.append(cur)
            cur = []
        cur.append(i)
    if cur:
        batches.append(cur)
    out = []
    for b in batches:
        longest = int(lengths[b[0]])
        out.append((b, -(-longest // quantum) * quantum))
    return out


def enhance_sharded(enhance_fn, clips, batch_size, rank=0, world=1, group=None, **kwargs):
    """clips: list of 1-D tensors. Returns the list of enhanced clips (original order) on rank 0,
    None elsewhere.  `enhance_fn(batch [B, T], **kwargs) -> [B, T]` is normally `model.enhance`."""
    lo, hi = shard_bounds(len(clips), rank, world)
    mine = list(range(lo, hi))
    results = {}
    for batch in length_buckets([c.shape[-1] for c in clips], mine, batch_size):
        x = torch.stack([clips[i] for i in batch])
        y = enhance_fn(x, **kwargs)
        for i, yi in zip(batch, y):
            results[i] = yi.detach().cpu()
    if world == 1:
        return [results[i] for i in range(len(clips))]
    import torch.distributed as dist

    gathered = [None] * world if rank == 0 else None
    dist.gather_object(results, gathered, dst=0, group=group)  # final host gather
    if rank != 0:
        return None
    merged = {}
    for r in gathered:
        merged.update(r)
    return [merged[i] for i in range(len(clips))]
