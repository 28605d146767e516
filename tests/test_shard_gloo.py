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


# ------------------------------------------------------------------ fold-range sharding of one utterance (section 8e)
def test_shard_folds_partitions_exactly():
    from wavernn_b200.shard import shard_folds
    for n in (1, 5, 20, 160):
        for world in (1, 2, 3, 8):
            rng = [shard_folds(n, world, r) for r in range(world)]
            assert [i for lo, hi in rng for i in range(lo, hi)] == list(range(n))
            assert max(hi - lo for lo, hi in rng) - min(hi - lo for lo, hi in rng) <= 1


def _sharded_case(mode):
    import numpy as np
    from oracle import fatchord as ofc
    from wavernn_b200 import synth
    from wavernn_b200.config import VocoderDims
    from common import TINY
    dims = VocoderDims(mode=mode, **TINY)
    sd = synth.synth_state_dict(dims, seed=0, style='trained')
    mel = synth.synth_mel(61, seed=1234, feat_dims=dims.feat_dims)
    target, overlap, mu_law = 100, 20, mode == 'RAW'
    W = ofc.to_torch(sd)
    with torch.no_grad():
        m, aux, wave_len = ofc.conditioning(W, dims, mel, True, target, overlap)
        B, S = int(m.shape[0]), int(m.shape[1])
        noise = synth.synth_noise_mol(S, B, 4321) if mode == 'MOL' else synth.synth_noise_exp(S, B, dims.n_classes, 4321)
        samples = ofc.run_loop(W, dims, m, aux, noise)['samples']
    want = ofc.finish(dims, samples, wave_len, True, target, overlap, mu_law)
    return dims, mel, target, overlap, mu_law, samples.numpy(), want


class _NoGpuVocoder:
    """binding.Vocoder stand-in: dims and host-only helpers; the per-shard GPU call is replaced by run_shard."""

    def __init__(self, dims):
        self.dims = dims

    @staticmethod
    def num_folds(total_len, target, overlap):
        from oracle.fatchord import num_folds_for
        return num_folds_for(total_len, target, overlap)


def _fold_worker(rank, world, port, mode, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        import numpy as np
        from wavernn_b200.shard import generate_sharded
        dims, mel, target, overlap, mu_law, samples, want = _sharded_case(mode)
        # the kernel is stubbed by the oracle's samples of the rank's own fold range
        out = generate_sharded(_NoGpuVocoder(dims), mel[0], target, overlap, mu_law,
                               run_shard=lambda lo, hi: samples[lo:hi])
        if rank == 0:
            q.put((out.dtype == np.float64, out.shape == want.shape, bool(np.array_equal(out, want))))
        else:
            q.put((out is None,))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('mode', ['MOL', 'RAW'])
def test_world_size_2_gloo_fold_sharded_utterance_is_bit_exact(mode):
    """Two ranks each hold half of the folds (the oracle's samples stand in for the kernel); rank 0 gathers and applies
    the float64 cross-fade / unfold / fade-out, incl. the <= overlap samples that straddle the shard edge: equal, bit
    for bit, to the reference-order tail of the unsharded utterance (oracle.finish, pinned against the reference)."""
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_fold_worker, args=(r, world, port, mode, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(out, key=len) == [(True,), (True, True, True)]


def test_combine_fold_samples_single_process_equals_oracle_tail():
    import numpy as np
    from wavernn_b200.shard import combine_fold_samples
    for mode in ('MOL', 'RAW'):
        dims, mel, target, overlap, mu_law, samples, want = _sharded_case(mode)
        got = combine_fold_samples(samples, want.shape[0], target, overlap, dims.hop_length, dims.n_classes if mu_law else None)
        assert np.array_equal(got, want)


# ---------------------------------------------------------------------------------------------------------------
# Tacotron decoder loop: sentences are independent, so N GPUs = N replicas, each decoding the
# sentences `shard_utterances` gives it; nothing crosses ranks but the timing reduction.  The decoder here is the
# CPU oracle standing in for the kernel (the host logic is what is under test).
def _taco_worker(rank, world, port, q):
    import numpy as np
    from oracle import tacotron as otc
    from wavernn_b200 import synth
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        sd = synth.synth_taco_state_dict(seed=0)
        lengths = [12, 30, 21]
        mine = shard_utterances(len(lengths), world, rank)
        out = {}
        for i in mine:
            mel, attn = otc.decode(sd, synth.synth_encoder_seq(lengths[i], seed=100 + i), steps=12, r=2)
            out[i] = (mel.shape, float(np.abs(mel).sum()), attn.shape)
        frames = sum(v[0][1] for v in out.values())
        rate, ms_max = whole_job_rate(frames, [10.0 * len(mine)])
        q.put((rank, out, rate, ms_max))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_sentences_shard_as_replicas():
    import numpy as np
    from oracle import tacotron as otc
    from wavernn_b200 import synth
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_taco_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    merged = {}
    for _, out, rate, ms_max in got:
        merged.update(out)
        assert ms_max == [20.0]                                             # the slower rank decodes two sentences
        assert rate == pytest.approx(3 * 12 / 0.020)                        # all frames of the job over that time
    sd = synth.synth_taco_state_dict(seed=0)
    for i, L in enumerate([12, 30, 21]):
        mel, attn = otc.decode(sd, synth.synth_encoder_seq(L, seed=100 + i), steps=12, r=2)
        assert merged[i] == (mel.shape, float(np.abs(mel).sum()), attn.shape)
