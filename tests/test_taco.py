"""Tacotron decoder loop (`models/tacotron.py:438-465`).

CPU: the oracle restatement against the fixtures the reference's own `Tacotron.generate` produced
(tests/golden/make_taco_golden.py: bit-equal there).  GPU: `taco_loop_kernel` through the C ABI against the same
fixtures.  Tolerance: the kernel sums in a different order and folds rnn_input into LSTM 1 and prenet.fc1 into the mel
projection (compositions formed in double): mel frames (range about +-1.5) are held to 2e-5 absolute and attention
rows (they sum to 1) to 5e-6 over up to 200 free-running decoder steps (measured on B200: 1.1e-6 and 2.2e-7 on the
fixtures, 3.4e-6 and 1.3e-6 at worst over 200 random shapes, tests/manual/taco_stress.py); the step
count of the stop rule is exact.
"""
import glob
import os

import numpy as np
import pytest

from wavernn_b200 import synth

GOLDEN = os.path.join(os.path.dirname(__file__), 'golden')
CASES = sorted(os.path.basename(p)[5:-4] for p in glob.glob(os.path.join(GOLDEN, 'taco_*.npz')))
MEL_TOL, ATTN_TOL = 2e-5, 5e-6


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
def test_kernel_whole_default_budget():
    """`generate`'s default budget, steps = 2000 (1 000 free-running decoder steps at r = 2, no stop): the error does not
    grow along the utterance (measured 2.2e-6 at worst in any block of 100 steps, 4.8e-7 on attention rows)."""
    from wavernn_b200.binding import TacotronDecoder
    from oracle import tacotron as otc
    sd = synth.synth_taco_state_dict(seed=3)
    enc = synth.synth_encoder_seq(120, seed=55)
    dec = TacotronDecoder()
    dec.load_weights(sd)
    mel, attn = dec.decode(enc, steps=2000, r=2, stop_threshold=-1e30)
    o_mel, o_attn = otc.decode(sd, enc, steps=2000, r=2, stop_threshold=-1e30)
    assert mel.shape == o_mel.shape == (80, 2000) and attn.shape == (1000, 120)
    err = np.abs(mel - o_mel).max(axis=0).reshape(10, 200).max(axis=1)
    assert err.max() < MEL_TOL and err[-1] < 4 * max(err[0], 1e-6)
    assert np.abs(attn - o_attn).max() < ATTN_TOL
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
def test_kernel_other_dims():
    """lstm_dims 256 and 40 mel bands (2 LSTM units and fewer rows per CTA), r = 3."""
    from wavernn_b200.binding import TacotronDecoder
    from oracle import tacotron as otc
    sd = synth.synth_taco_state_dict(lstm_dims=256, n_mels=40, seed=5)
    enc = synth.synth_encoder_seq(75, seed=8)
    dec = TacotronDecoder(lstm_dims=256, n_mels=40)
    dec.load_weights(sd)
    mel, attn = dec.decode(enc, steps=90, r=3)
    o_mel, o_attn = otc.decode(sd, enc, steps=90, r=3)
    assert mel.shape == o_mel.shape == (40, 90)
    assert np.abs(mel - o_mel).max() < MEL_TOL and np.abs(attn - o_attn).max() < ATTN_TOL
    dec.close()


@pytest.mark.gpu
def test_kernel_text_too_long_is_an_error_not_a_crash():
    """More characters than the position owners' shared memory holds: a clear error, and the handle stays usable."""
    from wavernn_b200.binding import TacotronDecoder
    sd = synth.synth_taco_state_dict(seed=0)
    dec = TacotronDecoder()
    dec.load_weights(sd)
    with pytest.raises((ValueError, RuntimeError), match='shared memory|512 characters'):
        dec.decode(synth.synth_encoder_seq(3000, seed=1), steps=4, r=2)
    mel, attn = dec.decode(synth.synth_encoder_seq(20, seed=1), steps=8, r=2)
    assert mel.shape == (80, 8) and attn.shape == (4, 20) and np.all(np.isfinite(mel))
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


def test_adapter_plumbing_against_the_reference_model():
    """`wavernn_b200.tacotron.generate` bound to the reference's own Tacotron object (build container only): with the
    decoder loop played by the oracle, the adapter's encoder / postnet plumbing must return exactly what the
    reference's `generate` returns."""
    from oracle import ref_shim, tacotron as otc
    if not ref_shim.available():
        pytest.skip('reference tree not present')
    import wavernn_b200.tacotron as bt

    class OracleDecoder:
        def __init__(self, **kw):
            self.kw = kw

        def load_weights(self, state):
            self.state = {k: v.detach().cpu().numpy() for k, v in state.items()}

        def decode(self, enc, steps, r, stop_threshold):
            return otc.decode(self.state, enc, steps=steps, r=r, stop_threshold=stop_threshold)

    sd = synth.synth_taco_state_dict(seed=0)
    enc = synth.synth_encoder_seq(37, seed=77)
    model = ref_shim.build_tacotron(sd, enc, r=2)
    x = [1 + (i % 60) for i in range(37)]
    want = model.generate(x, steps=60)
    real = bt.TacotronDecoder
    bt.TacotronDecoder = OracleDecoder
    try:
        got = bt.generate(model, x, steps=60)
    finally:
        bt.TacotronDecoder = real
        model.__dict__.pop('_b200_taco_decoder', None)
    for a, b in zip(want, got):
        assert a.shape == b.shape and np.array_equal(a, b)
    assert model.training


@pytest.mark.gpu
def test_adapter_on_gpu_with_a_stand_in_model():
    """The adapter end to end on the GPU: a model object with the reference's attribute names (embedding as the encoder,
    one convolution as the postnet), the decoder loop on the kernel; mel against the oracle, linear against the same
    postnet applied to the oracle's mel."""
    import torch
    from torch import nn
    from oracle import tacotron as otc
    import wavernn_b200.tacotron as bt

    from taco_standin import build_stand_in
    sd = synth.synth_taco_state_dict(seed=2)
    model = build_stand_in(sd).cuda()
    x = [3 + (i * 7) % 60 for i in range(45)]
    mel, linear, attn = bt.generate(model, x, steps=80)
    enc = model.encoder(torch.as_tensor(x, device='cuda')).detach().cpu().numpy()
    o_mel, o_attn = otc.decode(sd, enc, steps=80, r=2)
    assert mel.shape == o_mel.shape == (80, 80) and attn.shape == (40, 45) and linear.shape == (80, 80)
    assert np.abs(mel - o_mel).max() < MEL_TOL and np.abs(attn - o_attn).max() < ATTN_TOL
    with torch.no_grad():
        want_lin = model.post_proj(model.postnet(torch.as_tensor(o_mel, device='cuda').unsqueeze(0))).transpose(1, 2)[0].cpu().numpy()
    assert np.abs(linear - want_lin).max() < 1e-4
    assert model.training
    # the second call reuses the packed decoder; a parameter update re-packs it
    dec1 = model.__dict__['_b200_taco_decoder'][1]
    bt.generate(model, x, steps=20)
    assert model.__dict__['_b200_taco_decoder'][1] is dec1
    with torch.no_grad():
        model.decoder.mel_proj.weight.mul_(0.5)
    mel2, _, _ = bt.generate(model, x, steps=20)
    assert model.__dict__['_b200_taco_decoder'][1] is not dec1
    assert np.abs(mel2[:, :2] - 0.5 * mel[:, :2]).max() < MEL_TOL
