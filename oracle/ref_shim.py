"""Import the real reference (build container only) to pin the oracle.  TEST INFRASTRUCTURE.

`/root/reference` is read-only and exists only in the build container; on the GPU box the
un-modified copies staged by `oracle/stage_ref.py` into git-ignored `oracle/_ref/reference/` are
used instead (timed CPU baseline and call-compatibility test only).  Nothing on the product
path may call this.  The shims are the ones listed in SURVEY.md F4/F6 and touch no arithmetic:
  * stub `matplotlib` / `librosa` modules (`utils/display.py:1-3`, `utils/dsp.py:3`);
  * `np.cumproduct = np.cumprod` (`models/fatchord_version.py:68`; removed in NumPy 2);
  * `hp.configure(<reference>/hparams.py)`;
  * `fatchord_version.save_wav` -> no-op (librosa.output.write_wav no longer exists);
  * `deepmind_version.stream` -> accepts the 2-argument call at `deepmind_version.py:159`.
"""
import os
import sys
import types

_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref', 'reference')


def _find_root():
    # the build container has the tree itself; the GPU box has the copies `oracle/stage_ref.py` made at build time
    for cand in (os.environ.get('WAVERNN_REFERENCE'), '/root/reference', _STAGED):
        if cand and os.path.isfile(os.path.join(cand, 'models', 'fatchord_version.py')):
            return cand
    return '/root/reference'


REF_ROOT = _find_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, 'models', 'fatchord_version.py'))


_loaded = None


def load():
    """-> (fatchord_version module, deepmind_version module, distribution module, dsp module)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError('reference tree not present (expected in the build container only)')
    import numpy as np
    if not hasattr(np, 'cumproduct'):
        np.cumproduct = np.cumprod
    for name in ('matplotlib', 'matplotlib.pyplot', 'librosa'):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.use = lambda *a, **k: None
            sys.modules[name] = m
    sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    clash = [k for k in sys.modules if k == 'utils' or k.startswith('utils.') or k == 'models'
             or k.startswith('models.')]
    if clash:
        raise RuntimeError(f'module name clash with the reference tree: {clash}')
    sys.path.insert(0, REF_ROOT)
    try:
        from utils import hparams as hp
        if not hp.is_configured():
            hp.configure(os.path.join(REF_ROOT, 'hparams.py'))
        import models.fatchord_version as fv
        import models.deepmind_version as dv
        import utils.distribution as dist
        import utils.dsp as dsp
    finally:
        sys.path.remove(REF_ROOT)
    fv.save_wav = lambda x, path: None
    dv.stream = lambda *a, **k: None
    fv.stream = lambda *a, **k: None
    _loaded = (fv, dv, dist, dsp)
    return _loaded


def load_gen_wavernn():
    """The reference's CLI module `gen_wavernn.py` (its `gen_from_file` / `gen_testset` are the callers of the
    drop-in boundary), imported un-modified; its `save_wav` for the target file is a no-op."""
    load()
    # gen_wavernn imports utils.dataset -> utils.text (Tacotron's text front end, not on the path), which wants
    # `unidecode` and `inflect`: absent here, stubbed like matplotlib / librosa
    if 'unidecode' not in sys.modules:
        m = types.ModuleType('unidecode')
        m.unidecode = lambda s: s
        sys.modules['unidecode'] = m
    if 'inflect' not in sys.modules:
        m = types.ModuleType('inflect')
        m.engine = lambda: types.SimpleNamespace(number_to_words=lambda *a, **k: '')
        sys.modules['inflect'] = m
    sys.path.insert(0, REF_ROOT)
    try:
        import gen_wavernn as gw
    finally:
        sys.path.remove(REF_ROOT)
    gw.save_wav = lambda x, path: None
    return gw


def build_fatchord(dims, state_dict_np):
    """Reference `WaveRNN(**dims)` on CPU with the given checkpoint loaded (strict)."""
    import contextlib
    import io
    import torch
    fv, _, _, _ = load()
    with contextlib.redirect_stdout(io.StringIO()):
        model = fv.WaveRNN(**dims.ctor_kwargs())
    sd = {k: torch.from_numpy(v.copy()) for k, v in state_dict_np.items()}
    model.load_state_dict(sd, strict=True)
    return model


def build_deepmind(state_dict_np, hidden_size=896, quantisation=256):
    import contextlib
    import io
    import torch
    _, dv, _, _ = load()
    with contextlib.redirect_stdout(io.StringIO()):
        model = dv.WaveRNN(hidden_size=hidden_size, quantisation=quantisation)
    model.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in state_dict_np.items()}, strict=True)
    return model


class injected_noise:
    """Context manager: make the reference consume pre-drawn noise.

    MoL: `Tensor.uniform_` is replaced so the k-th call returns the k-th pre-drawn tensor
    (shapes checked) — draw order per step is u1 `(1,B,10)` then u2 `(1,B)`
    (`utils/distribution.py:106,118`).  RAW / deepmind: `Categorical.sample` is replaced by
    `argmax(probs / q)` with pre-drawn `q ~ Exp(1)`, which is what ATen multinomial computes
    (SURVEY.md F8).
    """

    def __init__(self, uniforms=None, exps=None):
        self.uniforms = list(uniforms) if uniforms is not None else None
        self.exps = list(exps) if exps is not None else None

    def __enter__(self):
        import torch
        self._u = torch.Tensor.uniform_
        self._s = torch.distributions.Categorical.sample
        ctx = self
        if self.uniforms is not None:
            it = iter(self.uniforms)

            def uniform_(t, *a, **k):
                # only the sampler's draws; `get_gru_cell` builds nn.GRUCell whose
                # reset_parameters also calls uniform_ (`fatchord_version.py:274`)
                if len(a) != 2 or a[0] != 1e-5:
                    return ctx._u(t, *a, **k)
                v = next(it)
                assert tuple(v.shape) == tuple(t.shape), (v.shape, t.shape)
                return t.copy_(v)
            torch.Tensor.uniform_ = uniform_
        if self.exps is not None:
            ie = iter(self.exps)

            def sample(dist, sample_shape=torch.Size()):
                q = next(ie)
                assert tuple(q.shape) == tuple(dist.probs.shape), (q.shape, dist.probs.shape)
                return torch.argmax(dist.probs / q, dim=-1)
            torch.distributions.Categorical.sample = sample
        return ctx

    def __exit__(self, *exc):
        import torch
        torch.Tensor.uniform_ = self._u
        torch.distributions.Categorical.sample = self._s
        return False


def build_tacotron(state_dict_np, enc_seq, r=2, stop_threshold=-3.4, decoder_dims=256, lstm_dims=512, n_mels=80):
    """Reference `Tacotron` (hparams.py:67-77 dims) on CPU with the decoder-loop parameters of `state_dict_np` loaded and
    its CBHG encoder's output replaced by `enc_seq` (a forward hook on `model.encoder`): the reference's own
    `generate` then runs its decoder loop on a known input.  -> model; call `model.generate(x, steps)` with any
    `x` of length L."""
    import contextlib
    import io
    import torch
    load()
    sys.path.insert(0, REF_ROOT)
    try:
        import models.tacotron as tc
    finally:
        sys.path.remove(REF_ROOT)
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        model = tc.Tacotron(embed_dims=256, num_chars=70, encoder_dims=decoder_dims // 2, decoder_dims=decoder_dims,
                            n_mels=n_mels, fft_bins=n_mels, postnet_dims=128, encoder_K=16, lstm_dims=lstm_dims,
                            postnet_K=8, num_highways=4, dropout=0.5, stop_threshold=stop_threshold)
    missing, unexpected = model.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in state_dict_np.items()},
                                                strict=False)
    if unexpected or any(k.startswith('decoder.') and k != 'decoder.r' for k in missing):
        raise RuntimeError(f'decoder keys do not match the reference: missing {missing}, unexpected {unexpected}')
    model.r = r
    enc = torch.as_tensor(enc_seq).unsqueeze(0)
    model.encoder.register_forward_hook(lambda mod, inp, out: enc)
    return model
