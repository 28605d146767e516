"""GPU: fold-range sharding on one GPU: the ranks' fold ranges run one after the other on the same
device, the host combines them (wavernn_b200/shard.py), and the result is compared with the unsharded call.

The same loop kernel is forced for both (the automatic choice depends on the fold count), so what is tested is that a
fold range reproduces its folds: noise and Philox addressed by the global fold index, conditioning restricted to the
frames the range covers.  RAW outputs are discrete: bit-exact.  MoL: a fold's position inside a 20-fold launch group
changes the order of its reduce-scatter sums, so the tolerance is the one of the other MoL tests (2e-4).
"""
import numpy as np
import pytest

from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
from wavernn_b200.shard import combine_fold_samples, shard_folds

pytestmark = pytest.mark.gpu


def _run_sharded(v, dims, mel, target, overlap, mu_law, world, noise=None, seed=0):
    T = mel.shape[-1]
    B = binding.num_folds(T * dims.hop_length, target, overlap)
    parts = []
    for r in range(world):
        lo, hi = shard_folds(B, world, r)
        if hi == lo:
            continue
        v.set_fold_range(lo, hi)
        v.stage([mel], True, target, overlap, mu_law, noise=noise)
        v.run_staged(seed=seed)
        s = v.fetch_samples()
        assert s.shape == (hi - lo, target + 2 * overlap) and v.last_timing()['folds'] == hi - lo
        with pytest.raises(RuntimeError):                    # no waveform from a shard: the tail needs every shard
            v.fetch()
        parts.append(s)
    v.set_fold_range(0, 0)
    samples = np.concatenate(parts, axis=0)
    classes = dims.n_classes if (dims.mode == 'RAW' and mu_law) else None
    return combine_fold_samples(samples, (T - 1) * dims.hop_length, target, overlap, dims.hop_length, classes), samples


@pytest.mark.parametrize('mode,world', [('MOL', 3), ('MOL', 8), ('RAW', 3)])
def test_fold_sharded_equals_unsharded_default_dims(mode, world):
    dims = VocoderDims(mode=mode) if mode == 'MOL' else VocoderDims(mode='RAW', bits=9)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    v.set_loop_kernel(3)                                     # the register-resident kernel for shards and whole alike
    target, overlap, mu_law = 300, 50, mode == 'RAW'
    mel = synth.synth_mel(31, seed=77)[0]                    # 25 folds x 400 steps
    B, S = binding.num_folds(31 * dims.hop_length, target, overlap), target + 2 * overlap
    assert B == 25
    noise = synth.synth_noise_mol(S, B, 9) if mode == 'MOL' else synth.synth_noise_exp(S, B, dims.n_classes, 9)
    whole = v.generate([mel], True, target, overlap, mu_law, noise=noise)[0]
    whole_samples = v.fetch_samples()
    got, samples = _run_sharded(v, dims, mel, target, overlap, mu_law, world, noise=noise)
    assert got.shape == whole.shape and got.dtype == np.float64
    if mode == 'RAW':
        assert np.array_equal(samples, whole_samples) and np.array_equal(got, whole)
    else:
        assert np.abs(samples - whole_samples).max() <= 2e-4 and np.abs(got - whole).max() <= 2e-4
    # Philox noise: keyed (seed, global fold, step), so a shard draws what the whole call draws for its folds
    whole_p = v.generate([mel], True, target, overlap, mu_law, seed=5)[0]
    got_p, _ = _run_sharded(v, dims, mel, target, overlap, mu_law, world, seed=5)
    if mode == 'RAW':
        assert np.array_equal(got_p, whole_p)
    else:
        assert np.abs(got_p - whole_p).max() <= 2e-4


def test_fold_range_tiny_dims_bit_exact_and_errors():
    """Tiny dims run on the grid-barrier kernel, whose per-fold arithmetic does not depend on the other folds: bit-exact."""
    from common import TINY
    dims = VocoderDims(mode='MOL', **TINY)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='trained'))
    mel = synth.synth_mel(61, seed=1234, feat_dims=dims.feat_dims)[0]
    target, overlap = 100, 20
    B = binding.num_folds(61 * dims.hop_length, target, overlap)
    noise = synth.synth_noise_mol(target + 2 * overlap, B, 4321)
    whole = v.generate([mel], True, target, overlap, False, noise=noise)[0]
    got, _ = _run_sharded(v, dims, mel, target, overlap, False, 2, noise=noise)
    assert np.array_equal(got, whole)
    v.set_fold_range(B + 3, B + 5)
    with pytest.raises(ValueError):                          # empty range
        v.stage([mel], True, target, overlap, False)
    v.set_fold_range(0, 2)
    with pytest.raises(ValueError):                          # a range needs one utterance, batched
        v.stage([mel, mel], True, target, overlap, False)
    with pytest.raises(ValueError):
        v.stage([mel], False, target, overlap, False)
    v.set_fold_range(0, 0)
    assert np.array_equal(v.generate([mel], True, target, overlap, False, noise=noise)[0], whole)
