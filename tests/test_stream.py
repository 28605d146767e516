"""Chunked delivery of finished folds: wrnn_generate_stream and the chunk-wise WAV writer."""
import os
import time

import numpy as np
import pytest

from wavernn_b200 import wavio


def test_wav_stream_writer_equals_save_wav(tmp_path):
    rng = np.random.default_rng(0)
    x = rng.standard_normal(12345).astype(np.float32) * 0.3
    a, b = tmp_path / 'a.wav', tmp_path / 'b.wav'
    wavio.save_wav(x, a, 22050)
    with wavio.WavStreamWriter(b, 22050) as w:
        for lo in range(0, x.size, 5000):
            w.write(x[lo:lo + 5000])
            data, rate = wavio.load_wav(b)              # readable mid-stream: lengths are patched at close, data is there
            assert rate == 22050
    assert open(a, 'rb').read() == open(b, 'rb').read()


@pytest.mark.gpu
@pytest.mark.parametrize('mode,mu_law,chunk', [('MOL', False, 20), ('MOL', False, 7), ('RAW', True, 20), ('MOL', False, 0)])
def test_streamed_chunks_equal_generate(mode, mu_law, chunk):
    """50 folds in groups: the chunks arrive in time order, end on fold boundaries, and concatenate to generate()'s
    waveform bit for bit (injected noise, so the two runs draw the same samples)."""
    from wavernn_b200 import binding, synth
    from wavernn_b200.config import VocoderDims
    dims = VocoderDims(mode=mode)
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    T, target, overlap = 101, 500, 50
    mel = synth.synth_mel(T, seed=5)
    B, S = binding.num_folds(T * dims.hop_length, target, overlap), target + 2 * overlap
    noise = synth.synth_noise_mol(S, B, 77) if mode == 'MOL' else synth.synth_noise_exp(S, B, dims.n_classes, 77)
    v.set_loop_kernel(3)
    ref = v.generate([mel[0]], True, target, overlap, mu_law, noise=noise)[0]
    got = []
    wave = v.generate_stream(mel[0], target, overlap, mu_law, on_chunk=lambda first, x, last: got.append((first, x, last)),
                             chunk_folds=chunk, noise=noise)
    assert np.array_equal(wave, ref)
    per = chunk if chunk else 20
    assert len(got) == -(-B // per)
    pos = 0
    for i, (first, x, last) in enumerate(got):
        assert first == pos and last == (i == len(got) - 1)
        if not last:
            assert (first + len(x)) % (target + overlap) == 0 and len(x) == per * (target + overlap)
        pos += len(x)
    assert pos == len(ref) and np.array_equal(np.concatenate([x for _, x, _ in got]), ref)
    # a failing callback surfaces as its own exception after the call, the handle stays usable
    with pytest.raises(KeyError):
        v.generate_stream(mel[0], target, overlap, mu_law, on_chunk=lambda *a: (_ for _ in ()).throw(KeyError('x')), noise=noise)
    assert np.array_equal(v.generate([mel[0]], True, target, overlap, mu_law, noise=noise)[0], ref)


@pytest.mark.gpu
def test_first_chunk_arrives_before_the_job_is_done(tmp_path):
    """160 folds x 1 200 steps in 8 groups: the first audio is on the host after about one group, not after eight; the
    chunks go straight into a WAV file."""
    from wavernn_b200 import binding, synth
    from wavernn_b200.config import VocoderDims
    dims = VocoderDims(mode='MOL')
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    target, overlap = 1000, 100
    T = 160 * (target + overlap) // dims.hop_length
    mel = synth.synth_mel(T, seed=6)
    v.set_loop_kernel(3)
    v.generate_stream(mel[0], target, overlap, False)                  # warm-up (workspaces)
    stamps = []
    path = tmp_path / 'stream.wav'
    t0 = time.perf_counter()
    with wavio.WavStreamWriter(path, dims.sample_rate) as w:
        def on_chunk(first, x, last):
            stamps.append(time.perf_counter() - t0)
            w.write(x)
        wave = v.generate_stream(mel[0], target, overlap, False, on_chunk=on_chunk, seed=3)
    total = time.perf_counter() - t0
    # about 1/8 of the call when nothing disturbs the box (measured 13 .. 15 %); the bound leaves room for a slow first group
    assert len(stamps) >= 8 and stamps[0] < 0.6 * total, (stamps, total)
    data, rate = wavio.load_wav(path)
    assert rate == dims.sample_rate and np.array_equal(data, wave.astype(np.float32))
    assert os.path.getsize(path) == 58 + 4 * len(wave)


@pytest.mark.gpu
def test_streamed_chunks_on_the_tensor_core_kernel():
    """Chunks of 64 folds run on the tensor-core kernel (three launches for 160 folds) and still concatenate to the
    waveform of the single 160-fold launch bit for bit: a fold's arithmetic does not depend on where in a tile it sits."""
    from wavernn_b200 import binding, synth
    from wavernn_b200.config import VocoderDims
    dims = VocoderDims(mode='MOL')
    v = binding.Vocoder(dims, device=0)
    v.load_weights(synth.synth_state_dict(dims, seed=0, style='init'))
    target, overlap = 500, 50
    T = 160 * (target + overlap) // dims.hop_length
    mel = synth.synth_mel(T, seed=6)
    B = binding.num_folds(T * dims.hop_length, target, overlap)
    noise = synth.synth_noise_mol(target + 2 * overlap, B, 5)
    ref = v.generate([mel[0]], True, target, overlap, False, noise=noise)[0]
    assert v.loop_kernel_name(B) == 'tcgen05-bf16x3' and v.loop_kernel_name(64) == 'tcgen05-bf16x3'
    got = []
    wave = v.generate_stream(mel[0], target, overlap, False, on_chunk=lambda first, x, last: got.append(len(x)), chunk_folds=64, noise=noise)
    assert np.array_equal(wave, ref) and len(got) == 3 and got[0] == 64 * (target + overlap)
