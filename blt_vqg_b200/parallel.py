"""Data parallelism: one process per GPU, gradient all-reduce (sum) over NCCL on NVLink 5 / NVSwitch.

The reference has no in-tree distributed code; Lightning's `gpus=N` (train_iq.py:349,373) gives DDP: per-rank batch
shards, per-rank BatchNorm statistics, bucketed gradient all-reduce overlapped with backward, mean over ranks.
Here the backward schedule is ours, so there are no autograd hooks: the schedule reports "everything at flat offset
>= o is final" (`mark_ready`) as it walks the layers in reverse, and the reducer launches one NCCL all-reduce per
ready range (bucketed to `bucket_bytes`) on a dedicated communication stream, ordered after the producing kernels by
an event.  The 1/world_size factor is folded into the clip+Adam kernel (hyper[3]); the global grad norm is computed on
the already-reduced arena, so clipping needs no extra collective.
"""
import os

import torch
import torch.distributed as dist


class GradReducer:
    def __init__(self, group=None, bucket_bytes=None, tail_bucket_bytes=None, tail_fraction=0.5):
        """bucket_bytes: all-reduce granularity while most of the backward is still ahead (default B200IQ_BUCKET_MB = 48);
        tail_bucket_bytes: granularity once the unsent part is the lower `tail_fraction` of the arena (the encoders and the
        embedding, the end of the backward) — whatever is still pending when the backward ends cannot overlap anything, so
        the pieces get smaller there (default B200IQ_TAIL_BUCKET_MB = 16; 0 or an explicit bucket_bytes without it: off)."""
        if bucket_bytes is None:
            bucket_bytes = int(os.environ.get("B200IQ_BUCKET_MB", "48")) << 20
            if tail_bucket_bytes is None:
                tail_bucket_bytes = int(os.environ.get("B200IQ_TAIL_BUCKET_MB", "16")) << 20
        assert dist.is_initialized()
        self.group = group
        self.world_size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.bucket_elems = bucket_bytes // 4
        self.tail_elems = (tail_bucket_bytes // 4) if tail_bucket_bytes else None
        self.tail_fraction = tail_fraction
        self.comm_stream = None
        self.hi = None           # everything in grad[hi:] has been handed to NCCL
        self.pending_lo = None   # lowest offset reported ready and not yet sent
        self.collectives = 0
        self._grad = None
        self.sent = []           # [lo, hi) pieces reduced out of order through mark_range (disjoint)
        self.range_pending = None

    def _stream(self, device):
        if self.comm_stream is None and device.type == "cuda":
            # same priority as the step stream (train.py): 13.63 vs 13.74 ms per step at 8 GPUs, no difference at 2
            self.comm_stream = torch.cuda.Stream(device=device, priority=int(os.environ.get("B200IQ_COMM_PRIORITY", "-1")))
        return self.comm_stream

    def begin_step(self, model):
        self._grad = model.grad_arena
        self.hi = self._grad.numel()
        self.pending_lo = self.hi
        self.sent, self.range_pending = [], None
        model.on_grad_ready = self.mark_ready
        model.on_grad_range = self.mark_range
        if hasattr(model, "seed_rank"):
            model.seed_rank = self.rank   # replicas draw different dropout masks / eps, like DDP processes do
        model.seed_rank = self.rank   # every rank draws its own dropout masks / latent eps, like DDP replicas

    def _send(self, lo, hi):
        if hi <= lo:
            return
        g = self._grad
        st = self._stream(g.device)
        if st is None:  # CPU tensors (gloo tests)
            dist.all_reduce(g[lo:hi], group=self.group)
            self.collectives += 1
            return
        ev = torch.cuda.Event()
        ev.record()  # after every kernel enqueued so far on the compute stream
        st.wait_event(ev)
        with torch.cuda.stream(st):
            dist.all_reduce(g[lo:hi], group=self.group)
        self.collectives += 1

    def _bucket(self, top):
        if self.tail_elems is not None and top <= self.tail_fraction * self._grad.numel():
            return min(self.tail_elems, self.bucket_elems)
        return self.bucket_elems

    def _send_unsent(self, lo, hi):
        """[lo, hi) minus the pieces mark_range already reduced"""
        for a, b in sorted(self.sent):
            if b <= lo or a >= hi:
                continue
            self._send(lo, min(a, hi))
            lo = max(lo, b)
        self._send(lo, hi)

    def _flush_range(self):
        if self.range_pending is not None:
            lo, hi = self.range_pending
            self._send(lo, hi)
            self.sent.append((lo, hi))
            self.range_pending = None

    def mark_range(self, lo, hi):
        """grad[lo:hi] is final although parts of grad[hi:] may not be: a stack that runs on a side stream next to the main
        backward (the context encoder beside the response encoder) reports its layers this way, from ITS stream, and the
        suffix sweep of mark_ready later skips them.  Adjacent pieces (layers, top down) are merged up to one bucket."""
        lo, hi = max(0, lo), min(hi, self._grad.numel())
        if hi <= lo:
            return
        if self.range_pending is not None and self.range_pending[0] == hi:
            self.range_pending = (lo, self.range_pending[1])
        else:
            self._flush_range()
            self.range_pending = (lo, hi)
        if self.range_pending[1] - self.range_pending[0] >= self._bucket(self.range_pending[1]):
            self._flush_range()

    def mark_ready(self, lo, force=False):
        """the backward schedule guarantees grad[lo:] will not be written again in this step"""
        lo = max(0, min(lo, self.hi))
        if lo < self.pending_lo:
            self.pending_lo = lo
        while self.hi - self.pending_lo >= self._bucket(self.hi) or (force and self.hi > self.pending_lo):
            lo_b = max(self.pending_lo, self.hi - self._bucket(self.hi))
            if self.range_pending is not None and self.range_pending[1] > lo_b:
                self._flush_range()   # the sweep has reached an out-of-order piece that is still waiting for neighbours
            self._send_unsent(lo_b, self.hi)
            self.hi = lo_b

    def finish_step(self, model):
        """flush what is left and make the compute stream wait for the communication stream"""
        self.mark_ready(0, force=True)
        model.on_grad_ready = None
        model.on_grad_range = None
        st = self._stream(self._grad.device)
        if st is not None:
            ev = torch.cuda.Event()
            ev.record(st)
            torch.cuda.current_stream().wait_event(ev)


def shard_batch(batch, rank, world_size):
    """contiguous B/N split of every tensor in a batch dict"""
    out = {}
    for k, v in batch.items():
        n = v.shape[0]
        assert n % world_size == 0, (k, n, world_size)
        per = n // world_size
        out[k] = v[rank * per:(rank + 1) * per]
    return out
