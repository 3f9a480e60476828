"""Multi-rank host logic on CPU: world_size 2, gloo.  The per-rank scan is played by the
oracle here (test infrastructure; the product scans on the GPU) -- what is under test is the sharding:
byte-balanced contiguous document ranges, the blob broadcast, and the hit-offset layout / concatenation
in document order."""
import os
import random
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from aha_b200 import shard  # noqa: E402


def test_partition_docs_properties():
    rng = random.Random(5)
    for _ in range(200):
        n = rng.choice([0, 1, 2, 7, 100])
        sizes = [rng.choice([0, 0, 1, 5, 1000, rng.randint(0, 5000)]) for _ in range(n)]
        world = rng.choice([1, 2, 3, 8])
        parts = shard.partition_docs(sizes, world)
        assert len(parts) == world
        assert parts[0][0] == 0 and parts[-1][1] == n
        for (a, b), (c, d) in zip(parts, parts[1:]):
            assert a <= b == c <= d  # contiguous, ordered
        total = sum(sizes)
        if total and n:
            per = [sum(sizes[a:b]) for a, b in parts]
            assert max(per) <= total / world + max(sizes) + 1  # balanced up to one document


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from oracle import oracle as O
    from aha_b200.ac import HIT_DTYPE
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        rng = random.Random(11)  # same corpus on every rank
        alpha = "abc "
        keys = list(dict.fromkeys("".join(rng.choice(alpha) for _ in range(rng.randint(1, 4))) for _ in range(60)))
        docs = ["".join(rng.choice(alpha) for _ in range(rng.choice([0, 3, 40, 400]))).encode() for _ in range(37)]
        # "blob": rank 0 owns the dictionary, the others receive it
        payload = "\n".join(keys).encode() if rank == 0 else b""
        got = bytes(shard.broadcast_blob(payload, 0).numpy().tobytes())
        keys_here = got.decode().split("\n")
        assert keys_here == keys
        o = O.AC.compile(keys_here)
        lo, hi = shard.partition_docs([len(d) for d in docs], world)[rank]
        mine = [o.scan(d) for d in docs[lo:hi]]
        hits = np.concatenate(mine) if mine else np.zeros(0, HIT_DTYPE)
        off = np.zeros(hi - lo + 1, np.int64)
        np.cumsum([len(m) for m in mine], out=off[1:])
        offsets, total = shard.gather_layout(len(hits))
        all_hits, all_off = shard.gather_hits(hits, off, 0)
        if rank == 0:
            want = [o.scan(d) for d in docs]
            want_hits = np.concatenate(want)
            want_off = np.zeros(len(docs) + 1, np.int64)
            np.cumsum([len(w) for w in want], out=want_off[1:])
            assert total == len(want_hits)
            assert offsets[0] == 0
            assert all_hits.tobytes() == want_hits.tobytes()
            assert np.array_equal(all_off, want_off)
            open(os.path.join(out_dir, "ok"), "w").write("%d" % total)
        else:
            assert all_hits is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_shard_concat(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(os.path.join(str(tmp_path), "ok"))
