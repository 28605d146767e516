"""GPU: the deepmind coarse/fine variant and the 16x1 block-pruned GRU against the oracle / golden fixtures.

Integer outputs (coarse, fine, combined) are compared bit-exact under identical injected Exp(1) noise.
The pruned model's oracle is the dense reference run on the masked weights.
"""
import numpy as np
import pytest
import torch

from oracle import deepmind as odm
from oracle import fatchord as ofc
from wavernn_b200 import binding, synth
from wavernn_b200.config import VocoderDims
from wavernn_b200.deepmind_version import WaveRNN as DeepmindWaveRNN

from common import CASES, case_inputs, golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('name,hs,S', [('dm_small', 64, 300), ('dm_full', 896, 400)])
def test_deepmind_vs_reference_fixture(name, hs, S):
    sd = synth.synth_dm_state_dict(hs, 256, seed=0)
    qc = synth.synth_noise_exp(S, 1, 256, seed=4321)
    qf = synth.synth_noise_exp(S, 1, 256, seed=4322)
    v = binding.DeepmindVocoder(hs, 256, device=0)
    v.load_weights(sd)
    out, c, f = v.generate(S, 1, qc, qf)
    g = golden('deepmind')
    assert out.dtype == np.int64 and out.shape == (1, S)
    assert np.array_equal(c[0], g[name + '_coarse'])
    assert np.array_equal(f[0], g[name + '_fine'])
    assert np.array_equal(out[0], g[name + '_output'])
    assert np.array_equal(out[0], c[0] * 256 + f[0] - 2 ** 15)          # combine_signal, utils/dsp.py:33-34


def test_deepmind_many_sequences_vs_oracle():
    hs, S, N = 64, 120, 11                      # 11 sequences: two launches (8 + 3)
    sd = synth.synth_dm_state_dict(hs, 256, seed=3)
    qc = synth.synth_noise_exp(S, N, 256, seed=11)
    qf = synth.synth_noise_exp(S, N, 256, seed=12)
    v = binding.DeepmindVocoder(hs, 256, device=0)
    v.load_weights(sd)
    out, c, f = v.generate(S, N, qc, qf)
    o_out, o_c, o_f = odm.generate(ofc.to_torch(sd), S, qc, qf, n_seq=N)
    assert np.array_equal(c, o_c) and np.array_equal(f, o_f) and np.array_equal(out, o_out)
    # every sequence equals a single-sequence run with its own noise column
    out1, c1, f1 = v.generate(S, 1, qc[:, 4:5], qf[:, 4:5])
    assert np.array_equal(c1[0], c[4]) and np.array_equal(f1[0], f[4])


def test_deepmind_module_contract_and_philox():
    model = DeepmindWaveRNN(hidden_size=64, quantisation=256).cuda()
    a = model.generate(200, seed=5)
    b = model.generate(200, seed=5)
    c = model.generate(200, seed=6)
    assert len(a) == 3 and all(x.dtype == np.int64 and x.shape == (200,) for x in a)
    assert all(np.array_equal(x, y) for x, y in zip(a, b)) and not np.array_equal(a[0], c[0])
    assert a[1].min() >= 0 and a[1].max() <= 255 and a[2].min() >= 0 and a[2].max() <= 255
    with pytest.raises(RuntimeError):
        DeepmindWaveRNN(hidden_size=64).generate(10)                   # CPU module: no fallback


def _mask_16x1(w, H, keep):
    """numpy statement of SURVEY.md row S: per gate, 16x1 blocks, keep the top `keep` fraction by L1 norm."""
    w = w.copy()
    K = w.shape[1]
    for g in range(3):
        blk = w[g * H:(g + 1) * H].reshape(H // 16, 16, K)
        norm = np.abs(blk.astype(np.float64)).sum(axis=1).reshape(-1)           # [block_row * K + col]
        n_keep = min(norm.size, max(1, int(np.floor(keep * norm.size + 0.5))))
        order = np.lexsort((np.arange(norm.size), -norm))
        m = np.zeros(norm.size, dtype=bool)
        m[order[:n_keep]] = True
        blk *= m.reshape(H // 16, 1, K)
    return w


@pytest.mark.parametrize('name', ['tiny_mol_batched', 'full_mol_batched'])
def test_pruned_16x1_vs_dense_oracle_on_masked_weights(name):
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    masked = v.prune_16x1(0.1, want_masked=True)
    H = dims.rnn_dims
    sd2 = dict(sd)
    for k, w in masked.items():
        ref = _mask_16x1(sd[k], H, 0.1)
        assert np.array_equal(w, ref), k
        kept = (np.abs(w.reshape(3, H // 16, 16, -1)).sum(axis=2) > 0).mean()
        assert abs(kept - 0.1) < 2e-3
        sd2[k] = w
    wave = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    ref = ofc.generate(ofc.to_torch(sd2), dims, mel, batched, target, overlap, mu_law, noise)
    assert wave.shape == ref.shape and np.abs(wave - ref).max() <= 2e-4
    if name.startswith('full'):
        assert v.loop_kernel_name(B) == 'fma-resident-sparse16x1'       # the dedicated 16x1 block-sparse products
    # loading weights again restores the dense model
    v.load_weights(sd)
    if name.startswith('full'):
        assert v.loop_kernel_name(B) == 'fma-resident'
    dense = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    assert np.abs(dense - golden(name)['wave']).max() <= 2e-4


@pytest.mark.parametrize('mode,keep', [('RAW', 0.1), ('MOL', 0.03), ('MOL', 0.14)])
def test_pruned_checkpoint_is_detected_and_runs_on_the_sparse_tables(mode, keep):
    """A checkpoint that was pruned elsewhere (the masked weights arrive through load_weights) takes the 16x1
    block-sparse tables when every (CTA, gate) keeps <= 80 columns, else the dense products; both against the dense
    oracle on the same masked weights.  RAW mode: sparse products beside the in-row fc3.  21 folds: two groups."""
    dims = VocoderDims(mode=mode)
    H = dims.rnn_dims
    sd = dict(synth.synth_state_dict(dims, seed=2, style='init'))
    for k in ('rnn1.weight_ih_l0', 'rnn1.weight_hh_l0', 'rnn2.weight_ih_l0', 'rnn2.weight_hh_l0'):
        sd[k] = _mask_16x1(sd[k], H, keep)
    mel = synth.synth_mel(26, seed=8, feat_dims=dims.feat_dims)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    assert v.loop_kernel_name(20) == ('fma-resident' if keep > 0.12 else 'fma-resident-sparse16x1')
    W = ofc.to_torch(sd)
    target, overlap = 300, 40
    with torch.no_grad():
        m, aux, wave_len = ofc.conditioning(W, dims, mel, True, target, overlap)
    B, S = int(m.shape[0]), int(m.shape[1])
    noise = synth.synth_noise_mol(S, B, 9) if mode == 'MOL' else synth.synth_noise_exp(S, B, dims.n_classes, 9)
    wave = v.generate([mel[0]], True, target, overlap, mode == 'RAW', noise=noise)[0]
    ref = ofc.generate(W, dims, mel, True, target, overlap, mode == 'RAW', noise)
    assert wave.shape == ref.shape and np.abs(wave - ref).max() <= 2e-4


@pytest.mark.parametrize('kernel', [1, 2])
def test_all_loop_kernels_agree_with_fixture(kernel):
    """The default kernel is covered by test_gpu_parity; force the grid-barrier and the pipelined kernels."""
    name = 'full_mol_batched'
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    v.set_loop_kernel(kernel)
    assert v.loop_kernel_name(B) == {1: 'fma', 2: 'fma-pipelined'}[kernel]
    wave = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    assert np.abs(wave - golden(name)['wave']).max() <= 2e-4
    v.set_loop_kernel(0)
    assert v.loop_kernel_name(B) == 'fma-resident'


def test_raw_default_dims_run_on_the_resident_kernel_and_agree_with_the_barrier_kernel():
    """RAW 9-bit at default dims: fc3 (512 rows) is a fifth layer of the row CTAs in the resident kernel."""
    name = 'full_raw_batched_mulaw'
    dims, sd, mel, noise, (T, batched, target, overlap, mu_law, B, S) = case_inputs(name)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    assert v.loop_kernel_name(B) == 'fma-resident'
    wave = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    labels = v.fetch_samples()
    assert np.abs(wave - golden(name)['wave']).max() <= 2e-4
    v.set_loop_kernel(1)
    assert v.loop_kernel_name(B) == 'fma'
    wave1 = v.generate([mel[0]], batched, target, overlap, mu_law, noise=noise)[0]
    labels1 = v.fetch_samples()
    # identical class draws under identical injected noise (the two kernels differ in summation order only)
    same = np.mean(np.asarray(labels) == np.asarray(labels1))
    assert same >= 0.999, same
    assert np.abs(wave1 - golden(name)['wave']).max() <= 2e-4


@pytest.mark.parametrize('mode,T,target,overlap', [('MOL', 64, 300, 50), ('MOL', 26, 300, 40), ('RAW', 40, 300, 50)])
def test_default_dims_fold_groups_vs_oracle(mode, T, target, overlap):
    """More than 20 folds: the resident kernel takes them in balanced groups of <= 20, each group with its own
    launch; noise and Philox addressing stay per (step, global fold).  21 folds is the smallest split (11 + 10)."""
    dims = VocoderDims(mode=mode)
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    mel = synth.synth_mel(T, seed=77, feat_dims=dims.feat_dims)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    W = ofc.to_torch(sd)
    with torch.no_grad():
        m, aux, wave_len = ofc.conditioning(W, dims, mel, True, target, overlap)
    B, S = int(m.shape[0]), int(m.shape[1])
    assert B > 20 and S == target + 2 * overlap                    # 51, 21 and 32 folds
    noise = synth.synth_noise_mol(S, B, 9) if mode == 'MOL' else synth.synth_noise_exp(S, B, dims.n_classes, 9)
    wave = v.generate([mel[0]], True, target, overlap, mode == 'RAW', noise=noise)[0]
    assert v.loop_kernel_name(20) == 'fma-resident' and v.last_timing()['folds'] == B
    ref = ofc.generate(W, dims, mel, True, target, overlap, mode == 'RAW', noise)
    assert wave.shape == ref.shape == (wave_len,)
    assert np.abs(wave - ref).max() <= 2e-4


def test_edge_inputs_default_dims():
    """Shortest utterance the reference accepts (T = 21: the 5 500-sample fade-out must fit, fatchord_version.py:256-258),
    a batched call that yields a single fold, ragged utterances in one call, and an empty call (ValueError)."""
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    v = binding.Vocoder(dims, device=0)
    v.load_weights(sd)
    W = ofc.to_torch(sd)
    # T = 21, unbatched, first 600 steps against the oracle (the rest of the 5 775 steps only adds time)
    mel = synth.synth_mel(21, seed=5, feat_dims=dims.feat_dims)
    wave = v.generate([mel[0]], False, 11000, 550, False, seed=1)[0]
    assert wave.shape == (20 * dims.hop_length,) and np.all(np.isfinite(wave)) and wave[-1] == 0.0     # linspace(1, 0) ends at 0
    assert v.last_timing()['folds'] == 1 and v.last_timing()['steps'] == 21 * dims.hop_length
    # batched with a target longer than the utterance: one fold, zero padded tail (fatchord_version.py:325-331)
    noise = synth.synth_noise_mol(7000 + 2 * 100, 1, 3)
    wb = v.generate([mel[0]], True, 7000, 100, False, noise=noise)[0]
    ref = ofc.generate(W, dims, mel, True, 7000, 100, False, noise)
    assert v.last_timing()['folds'] == 1 and wb.shape == ref.shape and np.abs(wb - ref).max() <= 2e-4
    # ragged: three utterances of different lengths in one call == three single calls (Philox keyed per utterance fold)
    mels = [synth.synth_mel(T, seed=40 + T, feat_dims=dims.feat_dims)[0] for T in (21, 37, 64)]
    outs = v.generate(mels, True, 2000, 100, False, seed=11)
    assert [len(o) for o in outs] == [(T - 1) * dims.hop_length for T in (21, 37, 64)]
    assert all(np.all(np.isfinite(o)) for o in outs)
    # empty call: an argument error, not a crash
    with pytest.raises(ValueError):
        v.generate([], True, 2000, 100, False, seed=0)


@pytest.mark.parametrize('mode', ['MOL', 'RAW'])
def test_config2_full_size_properties(mode):
    """BASELINE.json config 2 / 3 at full size (T = 801 -> 20 folds x 12 100 steps, 220 000 samples), through properties
    that need no oracle run: the call is deterministic in its seed, the waveform has the reference's length
    (fatchord_version.py:253) and range, an utterance does not depend on what else is in the call (folds are independent,
    :194-196), and the injected-noise path and the Philox path agree on the first fold-steps of teacher-free generation
    only in shape (different noise), so only shape is asserted there."""
    dims = VocoderDims(mode='RAW', bits=9) if mode == 'RAW' else VocoderDims(mode='MOL')
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    mel = synth.synth_mel(801, seed=1234)
    mel2 = synth.synth_mel(260, seed=77)
    mu_law = mode == 'RAW'
    a = v.generate([mel[0]], True, 11000, 550, mu_law, seed=11)[0]
    t = v.last_timing()
    assert (t['folds'], t['steps']) == (20, 12100) and v.loop_kernel_name(20).startswith('fma-resident')
    b = v.generate([mel[0]], True, 11000, 550, mu_law, seed=11)[0]
    c = v.generate([mel[0]], True, 11000, 550, mu_law, seed=12)[0]
    assert a.shape == (800 * 275,) and a.dtype == np.float64
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert np.all(np.isfinite(a)) and np.abs(a).max() <= 1.5   # cross-faded overlaps: sqrt gains add up to <= sqrt(2)
    assert np.all(a[-1:] == 0.0)                      # fade-out ends at exactly 0 (linspace(1, 0, 5500), :256-258)
    # independence: the same utterance as the second of two in one call (its folds then run in another launch group
    # with other fold indices); Philox is keyed by (seed, global fold, step), so compare through injected noise
    if mode == 'MOL':
        B1, B2, S = 20, binding.num_folds(260 * 275, 11000, 550), 12100
        n1, n2 = synth.synth_noise_mol(S, B1, 5), synth.synth_noise_mol(S, B2, 6)
        w2 = v.generate([mel2[0]], True, 11000, 550, False, noise=n2)[0]
        both = (np.concatenate([n1[0], n2[0]], axis=1), np.concatenate([n1[1], n2[1]], axis=1))
        o1, o2 = v.generate([mel[0], mel2[0]], True, 11000, 550, False, noise=both)
        assert np.array_equal(o2, w2) and o1.shape == a.shape


def test_exchange_calibration_is_measured_and_neutral_on_the_tuning_hardware():
    """the gather delays are scaled by a measured L2 round trip.  On a B200 at its default
    clocks the measurement lands inside the dead band and the tuned delays are used as they are."""
    dims = VocoderDims(mode='MOL')
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    assert v.exchange_calibration() == (0.0, 1.0)
    mel = synth.synth_mel(26, seed=3)
    v.set_loop_kernel(3)
    v.generate([mel[0]], True, 300, 40, False, seed=1)
    clk, scale = v.exchange_calibration()
    assert 200.0 < clk < 1000.0, clk
    assert 0.6 <= scale <= 2.0
    if abs(clk / 297.5 - 1.0) < 0.05:
        assert scale == 1.0
    else:
        assert abs(scale - min(2.0, max(0.6, clk / 297.5))) < 1e-3
