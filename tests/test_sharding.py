"""Host-side sharding logic on CPU, including a world_size-2 gloo run: blocks are contiguous, cover
every utterance exactly once, and each rank calls inference() only on its own rows."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from transformertts_b200.sharding import micro_batches, shard_bounds, sharded_inference


@pytest.mark.parametrize("n,world", [(512, 1), (512, 2), (512, 8), (13, 4), (3, 8), (0, 2)])
def test_shard_bounds_partition(n, world):
    spans = [shard_bounds(n, world, r) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for (a, b), (c, d) in zip(spans, spans[1:]):
        assert b == c and a <= b
    sizes = [b - a for a, b in spans]
    assert max(sizes) - min(sizes) <= 1


def test_micro_batches_cover_block():
    mbs = micro_batches(10, 75, 32)
    assert mbs == [(10, 42), (42, 74), (74, 75)]
    with pytest.raises(ValueError):
        micro_batches(0, 4, 0)


class _FakeModel:
    """Stands in for the CUDA model: records which rows it was asked to decode."""
    def inference(self, phoneme, phoneme_lens, max_len=1500, stop_threshold=0.5):
        assert phoneme.shape[1] == int(phoneme_lens.max())
        return {"pred_melspec": phoneme[:, :1].float().unsqueeze(-1).repeat(1, 2, 80), "rows": phoneme[:, 0].clone()}


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 21
    ph = torch.arange(n).view(n, 1).repeat(1, 6) + 1           # row id encoded in the data
    lens = torch.full((n,), 6)
    lens[5:] = 4
    parts = sharded_inference(_FakeModel(), ph, lens, micro_batch=4, max_len=10, gather=True)
    rows = torch.cat([r["rows"] for _, _, r in parts]).tolist()
    mine = sharded_inference(_FakeModel(), ph, lens, micro_batch=4, max_len=10, gather=False)
    q.put((rank, rows, [(s, e) for s, e, _ in mine]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_run_covers_every_utterance_once():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, rows, spans in got:
        assert rows == [i + 1 for i in range(21)]               # gathered view: every utterance exactly once
        lo, hi = shard_bounds(21, 2, rank)
        assert spans[0][0] == lo and spans[-1][1] == hi        # local view: only this rank's block
