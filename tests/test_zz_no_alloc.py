"""GPU: `wrnn_unfold_noalloc` (no counterpart in the reference).

The entry point allocates nothing: its device scratch is a file-scope static __device__ array of
WRNN_NOALLOC_MAX_DOUBLES doubles (csrc/noalloc.cu).  It must give what the existing entry point `wrnn_unfold` gives,
bit for bit (the tolerance of test_unfold_bit_exact: equality), on two calls in a row that share the array: the second
call has another shape and other data, so stale contents of the first would show.  Sizes beyond the array are rejected
with a clear error.  This file is named so that pytest collects it last.
"""
import numpy as np
import pytest

from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims

pytestmark = pytest.mark.gpu

TINY = dict(rnn_dims=64, fc_dims=64, bits=6, pad=2, upsample_factors=(2, 2, 3), feat_dims=16,
            compute_dims=32, res_out_dims=32, res_blocks=2, hop_length=12, sample_rate=22050)
MAX_DOUBLES = 524288          # WRNN_NOALLOC_MAX_DOUBLES (include/wavernn_b200.h)


def _vocoder():
    dims = VocoderDims(mode='MOL', **TINY)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, 0, 'init'))
    return v


def test_unfold_noalloc_equals_unfold_on_two_calls_in_a_row():
    v = _vocoder()
    rng = np.random.default_rng(7)
    # config 2's shape (20 folds x 12 100 samples: 474 650 of the 524 288 doubles), then a smaller, different call
    for B, tgt, ov in [(20, 11000, 550), (7, 300, 41)]:
        y = rng.uniform(-1, 1, size=(B, tgt + 2 * ov))
        want = v.unfold(y, tgt, ov)                           # existing entry point (cudaMalloc'ed scratch)
        out = np.full(B * (tgt + ov) + ov, np.nan)            # caller-owned output, poisoned
        got = v.unfold_noalloc(y, tgt, ov, out=out)
        assert got is out and got.dtype == np.float64 and got.shape == want.shape
        assert np.array_equal(got, want)
    # and the first shape again: whatever the second call left in the array does not leak into a third
    y = rng.uniform(-1, 1, size=(20, 12100))
    assert np.array_equal(v.unfold_noalloc(y, 11000, 550), v.unfold(y, 11000, 550))
    # edge: no overlap at all, one fold
    y = rng.uniform(-1, 1, size=(1, 50))
    assert np.array_equal(v.unfold_noalloc(y, 50, 0), v.unfold(y, 50, 0))


def test_unfold_noalloc_rejects_sizes_beyond_the_static_array():
    v = _vocoder()
    B, tgt, ov = 23, 11000, 550                               # 278 300 + 266 200 + 1 100 = 545 600 > 524 288 (22 folds still fit)
    assert B * (tgt + 2 * ov) + B * (tgt + ov) + ov + 2 * ov > MAX_DOUBLES
    with pytest.raises(ValueError, match='static'):
        v.unfold_noalloc(np.zeros((B, tgt + 2 * ov)), tgt, ov)
    with pytest.raises(ValueError, match='static'):          # overlap beyond WRNN_NOALLOC_MAX_OVERLAP
        v.unfold_noalloc(np.zeros((2, 100 + 2 * 5000)), 100, 5000)
    # the handle is still usable after a rejected call
    y = np.random.default_rng(1).uniform(-1, 1, size=(3, 140))
    assert np.array_equal(v.unfold_noalloc(y, 100, 20), v.unfold(y, 100, 20))
