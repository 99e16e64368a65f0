"""CPU (gloo, world_size 2): the bucketed gradient reducer that the backward schedule drives (parallel.GradReducer)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _FakeModel:
    def __init__(self, n, rank):
        g = torch.Generator().manual_seed(100 + rank)
        self.grad_arena = torch.randn(n, generator=g)
        self.on_grad_ready = None


def _worker(rank, world, port, n, bucket_bytes, marks, out, tail_bytes=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from blt_vqg_b200.parallel import GradReducer, shard_batch
        m = _FakeModel(n, rank)
        local = m.grad_arena.clone()
        red = GradReducer(bucket_bytes=bucket_bytes, tail_bucket_bytes=tail_bytes)
        red.begin_step(m)
        assert m.on_grad_ready is not None
        sent_before_finish = []
        for off in marks:                      # the backward schedule reports ready suffixes in decreasing offset order
            if isinstance(off, tuple):         # ... and (lo, hi) pieces of a stack that runs beside it, out of arena order
                m.on_grad_range(*off)
            else:
                m.on_grad_ready(off)
            sent_before_finish.append(red.collectives)
        red.finish_step(m)
        assert m.on_grad_ready is None
        gathered = [torch.empty(n) for _ in range(world)]
        dist.all_gather(gathered, local)
        want = sum(gathered)
        ok = torch.allclose(m.grad_arena, want, atol=1e-6)
        b = shard_batch({"x": torch.arange(8).view(8, 1), "y": torch.arange(16).view(8, 2)}, rank, world)
        shard_ok = b["x"].shape[0] == 4 and int(b["x"][0]) == 4 * rank
        if rank == 0:
            torch.save({"ok": ok, "shard_ok": shard_ok, "collectives": red.collectives, "sent": sent_before_finish}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,bucket_bytes,marks", [
    (10000, 4 * 3000, [9000, 6500, 3000, 100]),   # several buckets, some flushed before finish (overlap with "backward")
    (1000, 1 << 20, [500]),                        # everything fits one bucket: a single collective at finish
])
def test_grad_reducer_sums_over_ranks_in_buckets(tmp_path, n, bucket_bytes, marks):
    out = str(tmp_path / "res.pt")
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n, bucket_bytes, marks, out), nprocs=2, join=True)
    res = torch.load(out)
    assert res["ok"] and res["shard_ok"]
    elems = bucket_bytes // 4
    assert res["collectives"] == -(-n // elems)          # ceil(n / bucket): every element reduced exactly once
    if n > elems:
        assert res["sent"][-1] >= 1                      # buckets were launched before the end of "backward"
        assert res["sent"] == sorted(res["sent"])


@pytest.mark.parametrize("marks", [
    # suffix marks down to 7000; a side stack owns [2500, 5000) and reports it top-down as ranges while the main backward
    # goes on with suffix marks over [5000, 7000); then the join mark at 2500 and the rest at finish
    [9000, 7000, (4000, 5000), (3000, 4000), 6000, (2500, 3000), 5000, 2500],
    # ranges larger than a bucket, ranges that are not adjacent, a suffix sweep that runs into a piece still pending
    [(1000, 6000), (500, 800), 9500, 300],
])
def test_grad_reducer_out_of_order_ranges_are_reduced_exactly_once(tmp_path, marks):
    out = str(tmp_path / "res.pt")
    port = _free_port()
    mp.spawn(_worker, args=(2, port, 10000, 4 * 3000, marks, out), nprocs=2, join=True)
    res = torch.load(out)
    assert res["ok"]                                    # a piece sent twice would be summed twice: caught by the sum check
    assert res["sent"][-1] >= 2


def test_grad_reducer_smaller_buckets_at_the_end_of_the_backward(tmp_path):
    """below half of the arena the pieces shrink to tail_bucket_bytes: less is left pending when the backward ends"""
    out = str(tmp_path / "res.pt")
    port = _free_port()
    marks = [9000, 6000, 4500, 4000, 3000, (1000, 2000), 2000, 500]
    mp.spawn(_worker, args=(2, port, 10000, 4 * 3000, marks, out, 4 * 800), nprocs=2, join=True)
    res = torch.load(out)
    assert res["ok"]
    # 9000: nothing yet (1000 < 3000); 6000: one 3000 bucket; 4500: top is 7000 > 5000 -> still large buckets, 2500 pending;
    # 4000: 3000 sent, top 4000 is in the tail; 3000: one 800 piece; the range (1000, 2000) is a piece of its own
    assert res["sent"][:5] == [0, 1, 1, 2, 3] and res["sent"][5] == 4
    assert res["collectives"] > -(-10000 // 3000)
