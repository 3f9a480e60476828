"""Document sharding across ranks: one process per GPU, no data-path collective.

Documents are independent units (every reference call is per string, ac.cr:280-295), so the corpus is cut
into contiguous document ranges balanced by bytes: rank g scans docs [D_g, D_{g+1}) and rank order is
document order.  Two collectives only, both over `torch.distributed` (NCCL on GPUs, gloo in the CPU tests):

  * `broadcast_blob`  -- the flattened DFA blob built on rank 0, once per dictionary (ncclBroadcast);
  * `gather_layout`   -- per-rank hit totals -> where each rank's ordered hit block sits in the
                         concatenated result (all-gather of one int64 per rank + exclusive scan).

`gather_hits` additionally ships the hit blocks themselves to rank 0 for callers that want one array.
Nothing here touches the scan itself: a rank scans its shard with `AC.scan_batch` / `AC.scan_device`.
"""
import numpy as np
import warnings


def partition_docs(doc_sizes, world):
    """Contiguous document ranges [lo, hi) per rank, balanced by bytes.

    Rank r gets the documents whose cumulative start offset falls in [r, r+1) * total / world; ranges are
    contiguous, cover every document exactly once (empty documents included) and may be empty.
    """
    sizes = np.asarray(doc_sizes, dtype=np.int64)
    n = int(sizes.size)
    if world <= 0:
        raise ValueError("world must be positive")
    if n == 0:
        return [(0, 0)] * world
    starts = np.concatenate([[0], np.cumsum(sizes)[:-1]])
    total = int(sizes.sum())
    if total == 0:
        cuts = [(n * r) // world for r in range(world + 1)]
    else:
        bounds = [(total * r + world - 1) // world for r in range(world + 1)]  # ceil(total * r / world)
        cuts = [int(np.searchsorted(starts, b, side="left")) for b in bounds]
        cuts[0], cuts[-1] = 0, n
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def _dist():
    import torch.distributed as dist
    return dist


def broadcast_blob(blob, src=0, device=None):
    """Sends the DFA blob (bytes on `src`, ignored elsewhere) to every rank; returns a uint8 tensor on `device`."""
    import torch
    dist = _dist()
    rank = dist.get_rank()
    dev = device if device is not None else "cpu"
    n = torch.tensor([len(blob) if rank == src else 0], dtype=torch.int64, device=dev)
    dist.broadcast(n, src)
    buf = torch.empty(int(n.item()), dtype=torch.uint8, device=dev)
    if rank == src:
        src_arr = np.frombuffer(blob, dtype=np.uint8) if isinstance(blob, (bytes, bytearray, memoryview)) else np.asarray(blob)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")  # read-only view of the library's blob: only read from
            buf.copy_(torch.from_numpy(src_arr))
    dist.broadcast(buf, src)
    return buf


def gather_layout(n_hits_local, device=None):
    """All-gathers the per-rank hit totals.  Returns (offsets, total): offsets[r] = first index of rank r's
    block in the concatenated hit stream (document order = rank order)."""
    import torch
    dist = _dist()
    world = dist.get_world_size()
    dev = device if device is not None else "cpu"
    mine = torch.tensor([int(n_hits_local)], dtype=torch.int64, device=dev)
    parts = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    counts = [int(x.item()) for x in parts]
    offsets = [0] * world
    for r in range(1, world):
        offsets[r] = offsets[r - 1] + counts[r - 1]
    return offsets, sum(counts)


def gather_hits(hits_local, doc_hit_off_local, dst=0, device=None):
    """Concatenates the ranks' hit blocks on `dst` in document order.

    hits_local: structured array (start, end, value) of this rank's shard; doc_hit_off_local: int64
    [n_docs_local + 1] offsets into it.  Returns (hits, doc_hit_off) on `dst`, (None, None) elsewhere.
    """
    import torch
    dist = _dist()
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = device if device is not None else "cpu"
    flat = np.ascontiguousarray(hits_local).view(np.int32).reshape(-1)
    offsets, total = gather_layout(len(hits_local), dev)
    n_docs = torch.tensor([len(doc_hit_off_local) - 1], dtype=torch.int64, device=dev)
    doc_counts = [torch.zeros_like(n_docs) for _ in range(world)]
    dist.all_gather(doc_counts, n_docs)
    doc_counts = [int(x.item()) for x in doc_counts]
    t_hits = torch.from_numpy(flat.copy()).to(dev)
    t_off = torch.from_numpy(np.ascontiguousarray(doc_hit_off_local, dtype=np.int64).copy()).to(dev)
    if rank == dst:
        out = np.zeros(total * 3, dtype=np.int32)
        doc_off = np.zeros(sum(doc_counts) + 1, dtype=np.int64)
        d0 = 0
        for r in range(world):
            n_r = (offsets[r + 1] if r + 1 < world else total) - offsets[r]
            if r == rank:
                h, o = t_hits, t_off
            else:
                h = torch.empty(n_r * 3, dtype=torch.int32, device=dev)
                o = torch.empty(doc_counts[r] + 1, dtype=torch.int64, device=dev)
                if n_r:
                    dist.recv(h, src=r)
                dist.recv(o, src=r)
            out[offsets[r] * 3:(offsets[r] + n_r) * 3] = h.cpu().numpy()
            doc_off[d0:d0 + doc_counts[r] + 1] = o.cpu().numpy() + offsets[r]
            d0 += doc_counts[r]
        from .ac import HIT_DTYPE
        return out.view(HIT_DTYPE), doc_off
    if len(hits_local):
        dist.send(t_hits, dst=dst)
    dist.send(t_off, dst=dst)
    return None, None
