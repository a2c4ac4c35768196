"""Multi-GPU plumbing: one process per GPU, contiguous clip shards, one gather of logits.

Clips are independent (BatchNorm uses running statistics in eval mode), so the hot path has
no data-path collective; the only exchange is the final all_gather of ``[B/R, n_labels]``
logits (48 B/clip) over NCCL/NVLink - or gloo in the CPU tests.

``gather_logits`` is the simple synchronous form.  ``LogitGatherer`` keeps the collective off
the compute stream: the gather of step i runs on a side stream into a preallocated ``[B, L]``
buffer while step i+1's kernels already run, so the ranks are not re-synchronised by the
collective every step (the slowest power-capped board would otherwise set every step's time).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_clips, rank, world_size):
    """Contiguous ``[lo, hi)`` of rank's clips; the first ``n % R`` ranks take one extra."""
    base, rem = divmod(int(n_clips), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _world(group=None):
    if not dist.is_available() or not dist.is_initialized():
        return 1
    return dist.get_world_size(group)


def gather_logits(local, n_total, group=None):
    """all_gather of ragged ``[n_r, L]`` shards into ``[n_total, L]`` on every rank, in shard order."""
    world = _world(group)
    if world == 1:
        return local
    L = local.shape[1]
    base, rem = divmod(int(n_total), world)
    if rem == 0:                                   # equal shards: straight into the result, no padding / slicing / cat
        out = local.new_empty((n_total, L))
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    cap = base + 1
    buf = local.new_zeros((cap, L))
    buf[: local.shape[0]] = local
    out = local.new_empty((world * cap, L))
    dist.all_gather_into_tensor(out, buf, group=group)
    pieces = []
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        pieces.append(out[r * cap: r * cap + (hi - lo)])
    return torch.cat(pieces, 0)


class LogitGatherer:
    """Gathers each step's ``[n_r, L]`` logits on a side stream into one of two preallocated ``[n_total, L]`` buffers.

    ``submit(local)`` returns immediately (the compute stream is never made to wait for the collective); the gathered
    tensor of that step is returned by the NEXT ``submit`` (or by ``flush()``), i.e. consumed one step later.  On CPU
    tensors (gloo tests) the gather is synchronous and the same one-step-late contract holds.
    """

    def __init__(self, n_total, n_labels, device, dtype=torch.float32, group=None):
        self.group = group
        self.world = _world(group)
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.n_total, self.L = int(n_total), int(n_labels)
        self.device = torch.device(device)
        base, rem = divmod(self.n_total, self.world)
        self.equal = rem == 0
        self.cap = base if self.equal else base + 1
        self.cuda = self.device.type == "cuda"
        self.side = torch.cuda.Stream(device=self.device) if (self.cuda and self.world > 1) else None
        self.out = [torch.empty((self.world * self.cap, self.L), dtype=dtype, device=self.device) for _ in range(2)]
        self.pad = None if self.equal else [torch.zeros((self.cap, self.L), dtype=dtype, device=self.device) for _ in range(2)]
        self.done = [None, None]            # events: gather into out[k] finished
        self.k = 0
        self.pending = None                 # index of the buffer holding the previous step's gather
        self.seconds_in_submit = 0.0        # host time spent enqueueing (not device time)

    def _result(self, k):
        if self.done[k] is not None:
            torch.cuda.current_stream(self.device).wait_event(self.done[k])
        o = self.out[k]
        if self.equal:
            return o
        pieces = []
        for r in range(self.world):
            lo, hi = shard_range(self.n_total, r, self.world)
            pieces.append(o[r * self.cap: r * self.cap + (hi - lo)])
        return torch.cat(pieces, 0)

    def submit(self, local):
        """Enqueues this step's gather; returns the PREVIOUS step's gathered ``[n_total, L]`` (None on the first call)."""
        import time
        t0 = time.perf_counter()
        prev = self._result(self.pending) if self.pending is not None else None
        k = self.k
        if self.world == 1:
            self.out[k] = local
            self.done[k] = None
        elif self.side is None:                       # CPU / gloo
            src = local.contiguous()
            if not self.equal:
                self.pad[k][: local.shape[0]] = local
                src = self.pad[k]
            dist.all_gather_into_tensor(self.out[k], src, group=self.group)
        else:
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(self.device))
            local.record_stream(self.side)
            with torch.cuda.stream(self.side):
                self.side.wait_event(ready)
                src = local.contiguous()
                if not self.equal:
                    self.pad[k][: local.shape[0]].copy_(local)
                    src = self.pad[k]
                dist.all_gather_into_tensor(self.out[k], src, group=self.group)
                ev = torch.cuda.Event()
                ev.record(self.side)
                self.done[k] = ev
        self.pending = k
        self.k ^= 1
        self.seconds_in_submit += time.perf_counter() - t0
        return prev

    def flush(self):
        """The last submitted step's gathered logits (waits for its gather on the current stream)."""
        if self.pending is None:
            return None
        r = self._result(self.pending)
        self.pending = None
        return r
