"""Document sharding across the GPUs of one box and the host-side gather of the per-rank results
(documents are independent units, no collective on the data path).

shard_bounds    contiguous shards of about equal bytes (or tokens) by document
gather_batches  concatenation of per-rank outputs in document order with rebased offsets; uses
                torch.distributed.all_gather_object when a process group is given (gloo on CPU, nccl on GPU boxes)"""
import numpy as np


def shard_bounds(offsets, world):
    """offsets: uint64[ndocs+1] byte (or token) offsets. -> list of (first_doc, end_doc) per rank, contiguous,
    every document in exactly one shard, shard sizes balanced by bytes."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    ndocs = len(offsets) - 1
    total = int(offsets[-1]) if ndocs >= 0 else 0
    cuts = [0]
    for r in range(1, world):
        target = total * r // world
        d = int(np.searchsorted(offsets, np.uint64(target), side="left"))
        cuts.append(min(max(d, cuts[-1]), ndocs))
    cuts.append(ndocs)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def take_shard(data, offsets, bounds):
    """-> (data slice, offsets rebased to 0) of the documents [bounds[0], bounds[1])."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    b, e = bounds
    lo, hi = int(offsets[b]), int(offsets[e])
    return data[lo:hi], (offsets[b:e + 1] - offsets[b]).astype(np.uint64)


def concat_outputs(parts, record_keys):
    """parts: per-rank dicts in rank order. record_keys: {records key: offsets key} for every variable length array
    (e.g. {"lexems": "offsets"} or {"results": "result_offsets", "items": "item_offsets"}); all other keys are per
    document arrays and are concatenated as they are."""
    out = {}
    offset_keys = set(record_keys.values())
    for rk, ok in record_keys.items():
        out[rk] = np.concatenate([p[rk] for p in parts]) if parts else np.zeros(0)
        offs, base = [np.zeros(1, dtype=np.uint64)], 0
        for p in parts:
            o = np.asarray(p[ok], dtype=np.uint64)
            offs.append(o[1:] + np.uint64(base))
            base += int(o[-1])
        out[ok] = np.concatenate(offs)
    for k in (parts[0].keys() if parts else ()):
        if k in record_keys or k in offset_keys or np.isscalar(parts[0][k]) or not hasattr(parts[0][k], "__len__"):
            continue
        out[k] = np.concatenate([p[k] for p in parts])
    return out


def gather_batches(local, record_keys, group=None):
    """All ranks call this with their shard's output; every rank gets the whole batch in document order."""
    if group is None:
        return concat_outputs([local], record_keys)
    import torch.distributed as dist
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, local, group=group)
    return concat_outputs(parts, record_keys)
