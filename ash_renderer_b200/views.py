"""View sharding across GPUs.

Camera views are independent: each reads the shared, read-only scene (instances, draw templates,
lights — replicated once per GPU) plus its own camera / depth / colour / TAA history.  A view is
never split across GPUs.  The only inter-GPU exchange is one all_gather of the 32-byte per-view
`AshViewStats` records (`types.VIEW_STATS`) at the end of a batch: KBs, latency-bound, NCCL over
NVLink on GPUs, gloo in the CPU tests.  No image, mask or list ever crosses a link.
"""
import numpy as np

from . import types as T


def shard_views(n_views, world_size, rank, mode="block"):
    """View ids owned by `rank`.  "block": contiguous ranges whose sizes differ by at most one
    (keeps a view's TAA sequence and its neighbours' inputs on one GPU); "cyclic": v mod world."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    if mode == "cyclic":
        return list(range(rank, n_views, world_size))
    base, extra = divmod(n_views, world_size)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


def owner_of(view_id, n_views, world_size, mode="block"):
    if mode == "cyclic":
        return view_id % world_size
    base, extra = divmod(n_views, world_size)
    edge = extra * (base + 1)
    return view_id // (base + 1) if view_id < edge else extra + (view_id - edge) // max(base, 1)


def stats_to_words(stats):
    """list / array of VIEW_STATS records -> (n, 8) int32 array (the wire format of the all_gather)."""
    a = np.ascontiguousarray(np.asarray(stats, T.VIEW_STATS).reshape(-1))
    return a.view(np.uint32).reshape(-1, 8).astype(np.int32)


def words_to_stats(words):
    w = np.ascontiguousarray(np.asarray(words, np.int32).reshape(-1, 8)).astype(np.uint32)
    return w.view(T.VIEW_STATS).reshape(-1)


def all_gather_stats(local_stats, n_views, world_size, rank, mode="block", device=None):
    """Gather every rank's per-view stats; returns a VIEW_STATS array ordered by view_id.

    Uses torch.distributed's default process group (nccl: pass device="cuda"; gloo: device=None).
    Ranks may own different numbers of views, so records are padded to the largest shard and
    padding rows (view_id == 0xFFFFFFFF) are dropped after the gather."""
    import torch
    import torch.distributed as dist

    per_rank = max(len(shard_views(n_views, world_size, r, mode)) for r in range(world_size))
    mine = np.full((per_rank, 8), -1, np.int32)
    w = stats_to_words(local_stats)
    mine[:len(w)] = w
    t = torch.from_numpy(mine)
    if device is not None:
        t = t.to(device)
    if world_size == 1:
        out = t
    else:
        out = torch.empty((world_size * per_rank, 8), dtype=torch.int32, device=t.device)
        dist.all_gather_into_tensor(out, t)
    allw = out.cpu().numpy()
    allw = allw[allw[:, 0] != -1]
    st = words_to_stats(allw)
    return st[np.argsort(st["view_id"], kind="stable")]
