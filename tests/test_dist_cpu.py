"""world_size-2 gloo test of the sharding + logit gather used by bench.py --gpus N."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from honk_b200.parallel import LogitGatherer, gather_logits, shard_range


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 65, 16384):
        for w in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(n_total * 12, dtype=torch.float32).reshape(n_total, 12)
    lo, hi = shard_range(n_total, rank, world)
    out = gather_logits(full[lo:hi].clone(), n_total)
    ok = bool(torch.equal(out, full))
    # the pipelined gatherer: the result of step i is handed out by step i+1's submit (or flush)
    g = LogitGatherer(n_total, 12, "cpu")
    assert g.submit(full[lo:hi] * 1.0) is None
    for step in (2.0, 3.0):
        prev = g.submit(full[lo:hi] * step)
        ok = ok and bool(torch.equal(prev, full * (step - 1.0)))
    ok = ok and bool(torch.equal(g.flush(), full * 3.0)) and g.flush() is None
    q.put((rank, ok))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [37, 64])          # ragged and equal shards
def test_gather_logits_gloo_world2(n_total):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    [p.start() for p in procs]
    res = [q.get(timeout=120) for _ in procs]
    [p.join(60) for p in procs]
    assert sorted(res) == [(0, True), (1, True)]
