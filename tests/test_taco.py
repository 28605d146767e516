"""Tacotron decoder loop (`models/tacotron.py:438-465`).

CPU: the oracle restatement against the fixtures the reference's own `Tacotron.generate` produced
(tests/golden/make_taco_golden.py: bit-equal there).  GPU: `taco_loop_kernel` through the C ABI against the same
fixtures.  Tolerance: the kernel sums in a different order and folds rnn_input into LSTM 1 and prenet.fc1 into the mel
projection (compositions formed in double): mel frames (range about +-1.5) are held to 2e-5 absolute and attention
rows (they sum to 1) to 2e-6 over up to 200 free-running decoder steps (measured on B200: 1.1e-6 and 2.2e-7); the step
count of the stop rule is exact.
"""
import glob
import os

import numpy as np
import pytest

from wavernn_b200 import synth

GOLDEN = os.path.join(os.path.dirname(__file__), 'golden')
CASES = sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, 'taco_*.npz')))
MEL_TOL, ATTN_TOL = 2e-5, 2e-6


def load_case(name):
    z = np.load(os.path.join(GOLDEN, f'taco_{name}.npz'))
    sd = synth.synth_taco_state_dict(seed=int(z['weight_seed']))
    enc = synth.synth_encoder_seq(int(z['L']), seed=int(z['enc_seed']))
    return z, sd, enc


def test_fixtures_present():
    assert {'short_r2', 'long_r2', 'r7', 'r1_one_char', 'stop_rule'} <= set(CASES)


@pytest.mark.parametrize('name', CASES)
def test_oracle_matches_reference_fixture(name):
    from oracle import tacotron as otc
    z, sd, enc = load_case(name)
    mel, attn = otc.decode(sd, enc, steps=int(z['steps']), r=int(z['r']), stop_threshold=float(z['stop_threshold']))
    assert mel.shape == z['mel'].shape and attn.shape == z['attn'].shape
    # same ATen calls as the reference: equal to the last bit on the machine that made the fixture, and to rounding
    # of the BLAS kernels elsewhere
    np.testing.assert_allclose(mel, z['mel'], rtol=0, atol=1e-5)
    np.testing.assert_allclose(attn, z['attn'], rtol=0, atol=1e-6)
    np.testing.assert_allclose(attn.sum(axis=1), 1.0, atol=1e-5)


def test_stop_rule_fixture_ends_at_first_step_after_t10():
    z, _, _ = load_case('stop_rule')
    # r = 2: t = 0, 2, ..., 12 -> the first t > 10 is 12, seven decoder steps
    assert z['attn'].shape[0] == 7 and z['mel'].shape[1] == 14


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_kernel_matches_reference_fixture(name):
    from wavernn_b200.binding import TacotronDecoder
    z, sd, enc = load_case(name)
    dec = TacotronDecoder()
    dec.load_weights(sd)
    mel, attn = dec.decode(enc, steps=int(z['steps']), r=int(z['r']), stop_threshold=float(z['stop_threshold']))
    assert mel.shape == z['mel'].shape and attn.shape == z['attn'].shape
    assert np.abs(mel - z['mel']).max() < MEL_TOL
    assert np.abs(attn - z['attn']).max() < ATTN_TOL
    t = dec.last_timing()
    assert t['launches'] == 2 and t['ms_loop'] > 0
    dec.close()


@pytest.mark.gpu
def test_kernel_repeats_and_switches_r():
    """One handle: the same call twice is bit-identical (state is reset per call), and a change of r repacks."""
    from wavernn_b200.binding import TacotronDecoder
    from oracle import tacotron as otc
    sd = synth.synth_taco_state_dict(seed=3)
    enc = synth.synth_encoder_seq(90, seed=5)
    dec = TacotronDecoder()
    dec.load_weights(sd)
    a = dec.decode(enc, steps=50, r=2)
    b = dec.decode(enc, steps=50, r=2)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    for r, steps in ((5, 45), (20, 60), (3, 10)):
        mel, attn = dec.decode(enc, steps=steps, r=r)
        o_mel, o_attn = otc.decode(sd, enc, steps=steps, r=r)
        assert mel.shape == o_mel.shape
        assert np.abs(mel - o_mel).max() < MEL_TOL and np.abs(attn - o_attn).max() < ATTN_TOL
    dec.close()


@pytest.mark.gpu
def test_kernel_stop_rule_inside_a_longer_budget():
    """A threshold between the quiet and the loud steps: the loop must stop at the oracle's step, not earlier."""
    from wavernn_b200.binding import TacotronDecoder
    from oracle import tacotron as otc
    sd = synth.synth_taco_state_dict(seed=0)
    enc = synth.synth_encoder_seq(64, seed=9)
    ref_mel, _ = otc.decode(sd, enc, steps=120, r=2, stop_threshold=-1e9)
    # per-step maximum over the r frames; pick a threshold that the first step with t > 10 does not meet but a later one does
    per_step = ref_mel.reshape(80, -1, 2).max(axis=(0, 2))
    later = per_step[8:]
    thr = float(np.sort(later)[len(later) // 3]) + 1e-3
    want_mel, want_attn = otc.decode(sd, enc, steps=120, r=2, stop_threshold=thr)
    assert 7 < want_attn.shape[0] < 60
    dec = TacotronDecoder()
    dec.load_weights(sd)
    mel, attn = dec.decode(enc, steps=120, r=2, stop_threshold=thr)
    assert attn.shape == want_attn.shape
    assert np.abs(mel - want_mel).max() < MEL_TOL
    dec.close()


@pytest.mark.gpu
def test_kernel_long_text():
    """400 characters: 20 positions per owner CTA."""
    from wavernn_b200.binding import TacotronDecoder
    from oracle import tacotron as otc
    sd = synth.synth_taco_state_dict(seed=4)
    enc = synth.synth_encoder_seq(400, seed=6)
    dec = TacotronDecoder()
    dec.load_weights(sd)
    mel, attn = dec.decode(enc, steps=40, r=2)
    o_mel, o_attn = otc.decode(sd, enc, steps=40, r=2)
    assert np.abs(mel - o_mel).max() < MEL_TOL and np.abs(attn - o_attn).max() < ATTN_TOL
    dec.close()


@pytest.mark.gpu
def test_kernel_argument_errors():
    from wavernn_b200.binding import TacotronDecoder
    WrnnError = (ValueError, RuntimeError)
    with pytest.raises(WrnnError):
        TacotronDecoder(decoder_dims=128)
    dec = TacotronDecoder()
    with pytest.raises(WrnnError):
        dec.decode(synth.synth_encoder_seq(10), steps=10, r=2)          # no weights
    sd = synth.synth_taco_state_dict(seed=0)
    bad = dict(sd)
    bad.pop('decoder.attn_net.v.weight')
    dec.load_weights(bad)
    with pytest.raises(WrnnError):
        dec.decode(synth.synth_encoder_seq(10), steps=10, r=2)          # missing key
    dec.load_weights(sd)
    with pytest.raises(WrnnError):
        dec.decode(synth.synth_encoder_seq(10), steps=10, r=21)         # r > Decoder.max_r
    dec.close()
