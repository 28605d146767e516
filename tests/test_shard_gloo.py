"""World-size-2 run of the multi-GPU host logic on CPU (gloo): utterance sharding and the timing reduction.

The data path has no collective (independent utterances per rank), so this is all there is
to the N > 1 path besides N copies of the single-GPU path.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from wavernn_b200.shard import shard_utterances, whole_job_rate


def test_shard_utterances_partitions_exactly():
    for n in (0, 1, 2, 7, 8, 9, 64):
        for world in (1, 2, 3, 8):
            parts = [shard_utterances(n, world, r) for r in range(world)]
            flat = [i for p in parts for i in p]
            assert flat == list(range(n))                                  # disjoint, complete, ordered
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        shard_utterances(4, 2, 2)
    with pytest.raises(ValueError):
        shard_utterances(-1, 2, 0)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        mine = shard_utterances(5, world, rank)                                # 3 + 2 utterances
        samples = 220000 * len(mine)
        ms = [100.0 * len(mine) + rank, 130.0 * len(mine)]                     # rank 0 is the slow one
        rate, ms_max = whole_job_rate(samples, ms)
        got = torch.zeros(world, dtype=torch.int64)
        got[rank] = len(mine)
        dist.all_reduce(got)
        q.put((rank, mine, rate, ms_max, got.tolist()))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_rate_is_sum_over_max():
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, m0, rate0, ms0, cnt0), (r1, m1, rate1, ms1, cnt1) = out
    assert m0 == [0, 1, 2] and m1 == [3, 4] and cnt0 == cnt1 == [3, 2]
    assert ms0 == ms1 == [300.0, 390.0]                                       # MAX over ranks of each leg
    assert rate0 == rate1 == pytest.approx(5 * 220000 / 0.300)                 # SUM of samples / slowest rank


def test_single_process_needs_no_group():
    rate, ms = whole_job_rate(1000, [500.0])
    assert rate == pytest.approx(2000.0) and ms == [500.0]
    assert not dist.is_initialized()
