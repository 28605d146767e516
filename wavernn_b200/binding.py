"""ctypes binding of libwavernn_b200.so (include/wavernn_b200.h).

This is the stub a maintainer of the reference would add behind `WaveRNN.generate`
(INTEGRATION.md).  The library is loaded from this package directory (built in-tree by
`__graft_entry__.build()` / `make -C wavernn_b200/csrc`).  There is no fallback: if the
library is missing or no CUDA device is present the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libwavernn_b200.so')

WRNN_MODE_RAW, WRNN_MODE_MOL = 0, 1


class wrnn_config(C.Structure):
    _fields_ = [('rnn_dims', C.c_int32), ('fc_dims', C.c_int32), ('bits', C.c_int32), ('pad', C.c_int32),
                ('n_upsample', C.c_int32), ('upsample_factors', C.c_int32 * 8),
                ('feat_dims', C.c_int32), ('compute_dims', C.c_int32), ('res_out_dims', C.c_int32),
                ('res_blocks', C.c_int32), ('hop_length', C.c_int32), ('sample_rate', C.c_int32),
                ('mode', C.c_int32)]


class wrnn_tensor(C.Structure):
    _fields_ = [('name', C.c_char_p), ('data', C.c_void_p), ('dtype', C.c_int32), ('ndim', C.c_int32),
                ('shape', C.c_int64 * 4)]


class wrnn_noise(C.Structure):
    _fields_ = [('u1', C.c_void_p), ('u2', C.c_void_p), ('q', C.c_void_p), ('steps', C.c_int64),
                ('folds', C.c_int64)]


class MelParams(C.Structure):          # wrnn_mel_params
    _fields_ = [('sample_rate', C.c_int32), ('n_fft', C.c_int32), ('hop_length', C.c_int32), ('win_length', C.c_int32),
                ('num_mels', C.c_int32), ('fmin', C.c_float), ('min_level_db', C.c_float)]


_P = C.c_void_p
CHUNK_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32)      # wrnn_chunk_fn
_SIGS = {
    'wrnn_create': (C.c_int, [C.POINTER(wrnn_config), C.c_int, C.POINTER(_P)]),
    'wrnn_destroy': (None, [_P]),
    'wrnn_last_error': (C.c_char_p, [_P]),
    'wrnn_load_weights': (C.c_int, [_P, C.POINTER(wrnn_tensor), C.c_int32]),
    'wrnn_prune_16x1': (C.c_int, [_P, C.c_float, _P]),
    'wrnn_generate_stream': (C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, C.c_uint64, C.c_int32, _P, _P, _P, _P, _P]),
    'wrnn_melspectrogram': (C.c_int, [_P, _P, C.c_int64, _P, _P, C.c_int64, _P, _P]),
    'wrnn_generate': (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32,
                                C.c_int32, C.c_int32, C.POINTER(wrnn_noise), C.c_uint64, C.POINTER(_P),
                                C.POINTER(C.c_int64), _P]),
    'wrnn_stage': (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32,
                             C.c_int32, C.c_int32, C.POINTER(wrnn_noise), _P]),
    'wrnn_run_staged': (C.c_int, [_P, C.c_uint64, _P]),
    'wrnn_fetch': (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_int64), _P]),
    'wrnn_last_timing': (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float),
                                   C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    'wrnn_fetch_samples': (C.c_int, [_P, _P, _P]),
    'wrnn_set_fold_range': (C.c_int, [_P, C.c_int64, C.c_int64]),
    'wrnn_upsample': (C.c_int, [_P, _P, C.c_int32, _P, _P]),
    'wrnn_num_folds': (C.c_int64, [C.c_int64, C.c_int64, C.c_int64]),
    'wrnn_fold': (C.c_int, [_P, _P, C.c_int64, C.c_int32, C.c_int64, C.c_int64, _P]),
    'wrnn_unfold': (C.c_int, [_P, _P, C.c_int64, C.c_int64, C.c_int64, _P]),
    'wrnn_unfold_noalloc': (C.c_int, [_P, _P, C.c_int64, C.c_int64, C.c_int64, _P, _P]),
    'wrnn_decode_mu_law': (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P]),
    'wrnn_mulaw_lut': (C.c_int, [C.c_int32, _P]),
    'wrnn_set_mulaw_lut': (C.c_int, [_P, C.c_int32, _P]),
    'wrnn_xfade_tables': (C.c_int, [C.c_int64, _P, _P]),
    'wrnn_forward_logits': (C.c_int, [_P, _P, _P, C.c_int32, C.c_int64, _P]),
    'wrnn_loop_kernel_name': (C.c_char_p, [_P, C.c_int64]),
    'wrnn_set_loop_kernel': (C.c_int, [_P, C.c_int32]),
    'wrnn_set_mma_fold_tile': (C.c_int, [_P, C.c_int32]),
    'wrnn_exchange_calibration': (C.c_int, [_P, _P, _P]),
    'wrnn_debug_mma_state': (C.c_int, [_P, C.c_int32, C.c_int64, _P]),
    'wrnn_dm_create': (C.c_int, [C.c_int32, C.c_int32, C.c_int, C.POINTER(_P)]),
    'wrnn_dm_destroy': (None, [_P]),
    'wrnn_dm_last_error': (C.c_char_p, [_P]),
    'wrnn_dm_load_weights': (C.c_int, [_P, C.POINTER(wrnn_tensor), C.c_int32]),
    'wrnn_dm_generate': (C.c_int, [_P, C.c_int64, C.c_int32, _P, _P, C.c_uint64, _P, _P, _P, _P]),
    'wrnn_dm_last_timing': (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    'wrnn_taco_create': (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int, C.POINTER(_P)]),
    'wrnn_taco_destroy': (None, [_P]),
    'wrnn_taco_last_error': (C.c_char_p, [_P]),
    'wrnn_taco_load_weights': (C.c_int, [_P, C.POINTER(wrnn_tensor), C.c_int32]),
    'wrnn_taco_decode': (C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P, _P, C.POINTER(C.c_int32), _P]),
    'wrnn_taco_last_timing': (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
}

_lib = None


def lib():
    """Load the C-ABI library (once).  Raises if it has not been built — no fallback."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(f'{LIB_PATH} not built: run `python -c "import __graft_entry__ as g; g.build()"` '
                               'or `make -C wavernn_b200/csrc`; there is no CPU fallback')
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def exported_symbols():
    return sorted(_SIGS)


def _ptr(a):
    return a.ctypes.data_as(_P) if a is not None else None


def _raise(code, msg):
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    if code == -1:
        raise ValueError(msg)
    if code == -4:
        raise MemoryError(msg)
    raise RuntimeError(msg)


def _tensors(state):
    """dict name -> np.ndarray  ->  (wrnn_tensor array, keep-alive list)."""
    arr = (wrnn_tensor * len(state))()
    keep = []
    for i, (k, v) in enumerate(state.items()):
        v = np.asarray(v)
        if v.dtype == np.int64:
            dt = 1
        else:
            v = np.ascontiguousarray(v, dtype=np.float32)
            dt = 0
        if v.ndim > 4:
            raise ValueError(f'{k}: more than 4 dims')
        name = k.encode()
        keep += [v, name]
        arr[i].name, arr[i].data, arr[i].dtype, arr[i].ndim = name, v.ctypes.data, dt, v.ndim
        for j, s in enumerate(v.shape):
            arr[i].shape[j] = s
    return arr, keep


class Vocoder:
    """Owns one `wrnn_handle` (one per GPU).  Thin, typed wrappers over the C ABI."""

    def __init__(self, dims, device=0):
        self.dims = dims
        cfg = wrnn_config()
        cfg.rnn_dims, cfg.fc_dims, cfg.bits, cfg.pad = dims.rnn_dims, dims.fc_dims, dims.bits, dims.pad
        cfg.n_upsample = len(dims.upsample_factors)
        for i, f in enumerate(dims.upsample_factors):
            cfg.upsample_factors[i] = f
        cfg.feat_dims, cfg.compute_dims, cfg.res_out_dims = dims.feat_dims, dims.compute_dims, dims.res_out_dims
        cfg.res_blocks, cfg.hop_length, cfg.sample_rate = dims.res_blocks, dims.hop_length, dims.sample_rate
        if dims.mode not in ('RAW', 'MOL'):
            raise RuntimeError("Unknown model mode value - ", dims.mode)
        cfg.mode = WRNN_MODE_MOL if dims.mode == 'MOL' else WRNN_MODE_RAW
        self._h = _P()
        rc = lib().wrnn_create(C.byref(cfg), int(device), C.byref(self._h))
        if rc != 0:
            _raise(rc, lib().wrnn_last_error(None))
        if dims.mode == 'RAW':
            self.set_mulaw_lut(dims.n_classes)

    def close(self):
        if getattr(self, '_h', None):
            lib().wrnn_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            _raise(rc, lib().wrnn_last_error(self._h))

    @staticmethod
    def mulaw_table(mu):
        """decode_mu_law(y, mu, from_labels=False) (`utils/dsp.py:98-103`) at the mu sample values y = float32(2k/(mu-1)-1),
        evaluated with the reference's own numpy expression."""
        k = np.arange(mu, dtype=np.float32)
        y = (np.float32(2) * k / np.float32(mu - 1.) - np.float32(1)).astype(np.float64)
        m = mu - 1
        return np.ascontiguousarray(np.sign(y) / m * ((1 + m) ** np.abs(y) - 1), dtype=np.float64)

    @staticmethod
    def num_folds(total_len, target, overlap):
        return num_folds(total_len, target, overlap)

    def set_mulaw_lut(self, mu):
        """Install that table, so decode is bit-equal on this host."""
        self._ck(lib().wrnn_set_mulaw_lut(self._h, int(mu), _ptr(self.mulaw_table(mu))))

    def load_weights(self, state):
        arr, keep = _tensors(state)
        self._ck(lib().wrnn_load_weights(self._h, arr, len(state)))

    # ---- the path -------------------------------------------------------------------------
    def _noise(self, noise):
        if noise is None:
            return None, []
        n = wrnn_noise()
        keep = []
        if self.dims.mode == 'MOL':
            u1 = np.ascontiguousarray(noise[0], dtype=np.float32)
            u2 = np.ascontiguousarray(noise[1], dtype=np.float32)
            n.u1, n.u2, n.q = u1.ctypes.data, u2.ctypes.data, None
            n.steps, n.folds = u2.shape
            keep = [u1, u2]
        else:
            q = np.ascontiguousarray(noise, dtype=np.float32)
            n.u1, n.u2, n.q = None, None, q.ctypes.data
            n.steps, n.folds = q.shape[0], q.shape[1]
            keep = [q]
        return C.pointer(n), keep + [n]

    def _mel_args(self, mels):
        mels = [np.ascontiguousarray(m, dtype=np.float32) for m in mels]
        for m in mels:
            if m.ndim != 2 or m.shape[0] != self.dims.feat_dims:
                raise ValueError(f'mel must be ({self.dims.feat_dims}, T), got {m.shape}')
        ptrs = (_P * len(mels))(*[m.ctypes.data for m in mels])
        Ts = (C.c_int32 * len(mels))(*[m.shape[1] for m in mels])
        return mels, ptrs, Ts

    def generate(self, mels, batched, target, overlap, mu_law, noise=None, seed=0, stream=None):
        """mels: list of (feat, T) float32 arrays -> list of float64 waves of (T-1)*hop samples."""
        mels, ptrs, Ts = self._mel_args(mels)
        outs = [np.empty(max((m.shape[1] - 1) * self.dims.hop_length, 0), dtype=np.float64) for m in mels]
        optrs = (_P * len(outs))(*[o.ctypes.data for o in outs])
        lens = (C.c_int64 * len(outs))()
        nz, keep = self._noise(noise)
        self._ck(lib().wrnn_generate(self._h, ptrs, Ts, len(mels), int(bool(batched)), int(target), int(overlap),
                                     int(bool(mu_law)), nz, int(seed), optrs, lens, stream))
        return outs

    def generate_stream(self, mel, target, overlap, mu_law, on_chunk=None, chunk_folds=0, noise=None, seed=0, stream=None):
        """One batched utterance, delivered in time order as fold groups finish (wrnn_generate_stream).

        on_chunk(first_sample, samples (float64 copy), last) is called from this thread while later groups are still
        running on the GPU.  Returns the whole float64 waveform (equal to `generate`)."""
        mels, ptrs, Ts = self._mel_args([mel])
        out = np.empty(max((mels[0].shape[1] - 1) * self.dims.hop_length, 0), dtype=np.float64)
        n_out = C.c_int64(0)
        nz, keep = self._noise(noise)
        errors = []

        def trampoline(user, first, ptr, n, last):
            if on_chunk is None or errors:
                return
            try:
                on_chunk(int(first), np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(int(n),)).copy(), bool(last))
            except BaseException as e:      # noqa: BLE001  (must not unwind through the C frame)
                errors.append(e)

        cb = CHUNK_FN(trampoline)
        self._ck(lib().wrnn_generate_stream(self._h, _ptr(mels[0]), int(Ts[0]), int(target), int(overlap), int(bool(mu_law)), nz,
                                            int(seed), int(chunk_folds), cb, None, _ptr(out), C.byref(n_out), stream))
        if errors:
            raise errors[0]
        return out

    def stage(self, mels, batched, target, overlap, mu_law, noise=None, stream=None):
        mels, ptrs, Ts = self._mel_args(mels)
        nz, keep = self._noise(noise)
        self._ck(lib().wrnn_stage(self._h, ptrs, Ts, len(mels), int(bool(batched)), int(target), int(overlap),
                                  int(bool(mu_law)), nz, stream))
        self._staged_T = [m.shape[1] for m in mels]

    def run_staged(self, seed=0, stream=None):
        self._ck(lib().wrnn_run_staged(self._h, int(seed), stream))

    def fetch(self, stream=None):
        outs = [np.empty((T - 1) * self.dims.hop_length, dtype=np.float64) for T in self._staged_T]
        optrs = (_P * len(outs))(*[o.ctypes.data for o in outs])
        lens = (C.c_int64 * len(outs))()
        self._ck(lib().wrnn_fetch(self._h, optrs, lens, stream))
        return outs

    def last_timing(self):
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        s, f, k = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(lib().wrnn_last_timing(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(s), C.byref(f), C.byref(k)))
        return dict(ms_cond=a.value, ms_loop=b.value, ms_tail=c.value, steps=s.value, folds=f.value,
                    launches=k.value)

    def set_fold_range(self, lo, hi):
        """Run only folds [lo, hi) of the next single-utterance batched stage (0, 0 = all): see shard.py."""
        self._ck(lib().wrnn_set_fold_range(self._h, int(lo), int(hi)))

    def fetch_samples(self, stream=None):
        t = self.last_timing()
        out = np.empty((t['folds'], t['steps']), dtype=np.float32)
        self._ck(lib().wrnn_fetch_samples(self._h, _ptr(out), stream))
        return out

    def loop_kernel_name(self, folds):
        return lib().wrnn_loop_kernel_name(self._h, int(folds)).decode()

    def set_loop_kernel(self, which):
        """0 auto, 1 fma, 2 fma-pipelined, 4 tcgen05-bf16x3 (tensor cores; auto from 41 folds up)."""
        self._ck(lib().wrnn_set_loop_kernel(self._h, int(which)))

    def debug_mma_state(self, which, folds):
        out = np.empty((folds, 512), dtype=np.float32)
        self._ck(lib().wrnn_debug_mma_state(self._h, int(which), int(folds), _ptr(out)))
        return out

    def exchange_calibration(self):
        """(L2 round trip in SM clocks, gather-delay scale) of the resident kernel's calibration pass; (0, 1) before its first run."""
        a, b = C.c_float(), C.c_float()
        self._ck(lib().wrnn_exchange_calibration(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def set_mma_fold_tile(self, nt):
        self._ck(lib().wrnn_set_mma_fold_tile(self._h, int(nt)))

    def melspectrogram(self, wav, sample_rate=22050, n_fft=2048, hop_length=275, win_length=1100, num_mels=80, fmin=40.0,
                       min_level_db=-100.0, stream=None):
        """`melspectrogram(y)` of utils/dsp.py:72-75 on the GPU: float32 [num_mels][1 + len(wav) // hop_length]."""
        wav = np.ascontiguousarray(wav, dtype=np.float32).reshape(-1)
        mp = MelParams(int(sample_rate), int(n_fft), int(hop_length), int(win_length), int(num_mels), float(fmin), float(min_level_db))
        frames = C.c_int64(0)
        self._ck(lib().wrnn_melspectrogram(self._h, _ptr(wav), wav.size, C.byref(mp), None, 0, C.byref(frames), stream))
        out = np.empty((int(num_mels), frames.value), dtype=np.float32)
        self._ck(lib().wrnn_melspectrogram(self._h, _ptr(wav), wav.size, C.byref(mp), _ptr(out), frames.value, C.byref(frames), stream))
        return out

    def prune_at_step(self, t, start_prune, prune_steps, target_sparsity, want_masked=False):
        """16x1 block-prune the loaded GRU matrices to the sparsity the reference's schedule asks for at training step t
        (`notebooks/Pruning - Scratchpad.ipynb:127-131`; wavernn_b200/prune.py drives it during training)."""
        from .prune import sparsity_at
        return self.prune_16x1(1.0 - sparsity_at(t, start_prune, prune_steps, target_sparsity), want_masked=want_masked)

    def prune_16x1(self, keep_fraction, want_masked=False):
        d = self.dims
        H, A = d.rnn_dims, d.aux_dims
        n = 3 * H * (H + H + (H + A) + H)
        out = np.empty(n, dtype=np.float32) if want_masked else None
        self._ck(lib().wrnn_prune_16x1(self._h, float(keep_fraction), _ptr(out)))
        if not want_masked:
            return None
        sizes = [('rnn1.weight_ih_l0', H), ('rnn1.weight_hh_l0', H), ('rnn2.weight_ih_l0', H + A), ('rnn2.weight_hh_l0', H)]
        res, o = {}, 0
        for k, K in sizes:
            res[k] = out[o:o + 3 * H * K].reshape(3 * H, K).copy()
            o += 3 * H * K
        return res

    # ---- components -----------------------------------------------------------------------
    def upsample(self, mel):
        mel = np.ascontiguousarray(mel, dtype=np.float32)
        T = mel.shape[1]
        L = T * self.dims.hop_length
        m = np.empty((L, self.dims.feat_dims), dtype=np.float32)
        a = np.empty((L, self.dims.res_out_dims), dtype=np.float32)
        self._ck(lib().wrnn_upsample(self._h, _ptr(mel), T, _ptr(m), _ptr(a)))
        return m, a

    def fold(self, x, target, overlap):
        x = np.ascontiguousarray(x, dtype=np.float32)
        L, F = x.shape
        if L <= overlap or target < 1:
            raise ValueError('sequence shorter than the overlap')
        n = lib().wrnn_num_folds(L, target, overlap)
        out = np.empty((n, target + 2 * overlap, F), dtype=np.float32)
        self._ck(lib().wrnn_fold(self._h, _ptr(x), L, F, target, overlap, _ptr(out)))
        return out

    def unfold(self, y, target, overlap):
        y = np.ascontiguousarray(y, dtype=np.float64)
        folds, length = y.shape
        target = length - 2 * overlap          # recomputed like fatchord_version.py:375
        out = np.empty(folds * (target + overlap) + overlap, dtype=np.float64)
        self._ck(lib().wrnn_unfold(self._h, _ptr(y), folds, target, overlap, _ptr(out)))
        return out

    def unfold_noalloc(self, y, target, overlap, out=None, stream=None):
        """wrnn_unfold_noalloc: same result as `unfold`, device scratch in a static array (no cudaMalloc); `out` may be
        a caller-owned float64 array so that the call allocates nothing on the host side either."""
        y = np.ascontiguousarray(y, dtype=np.float64)
        folds, length = y.shape
        target = length - 2 * overlap
        if out is None:
            out = np.empty(folds * (target + overlap) + overlap, dtype=np.float64)
        self._ck(lib().wrnn_unfold_noalloc(self._h, _ptr(y), folds, target, overlap, _ptr(out), stream))
        return out

    def decode_mu_law(self, y, mu):
        y = np.ascontiguousarray(y, dtype=np.float32)
        out = np.empty(y.shape, dtype=np.float64)
        self.set_mulaw_lut(mu)
        self._ck(lib().wrnn_decode_mu_law(self._h, _ptr(y), y.size, int(mu), _ptr(out)))
        return out

    def forward_logits(self, x, mels):
        x = np.ascontiguousarray(x, dtype=np.float32)
        mels = np.ascontiguousarray(mels, dtype=np.float32)
        n, S = x.shape
        out = np.empty((n, S, self.dims.n_classes), dtype=np.float32)
        self._ck(lib().wrnn_forward_logits(self._h, _ptr(x), _ptr(mels), n, S, _ptr(out)))
        return out


def mulaw_lut(mu):
    out = np.empty(mu, dtype=np.float64)
    if lib().wrnn_mulaw_lut(int(mu), _ptr(out)) != 0:
        raise ValueError('bad mu')
    return out


def xfade_tables(overlap):
    fi = np.empty(max(overlap, 1), dtype=np.float64)
    fo = np.empty(max(overlap, 1), dtype=np.float64)
    if lib().wrnn_xfade_tables(int(overlap), _ptr(fi), _ptr(fo)) != 0:
        raise ValueError('bad overlap')
    return fi[:overlap], fo[:overlap]


def num_folds(total_len, target, overlap):
    return int(lib().wrnn_num_folds(int(total_len), int(target), int(overlap)))


class DeepmindVocoder:
    """Owns one `wrnn_dm_handle`: the coarse/fine dual-softmax variant (deepmind_version.py)."""

    def __init__(self, hidden_size=896, quantisation=256, device=0):
        self.hidden_size, self.quantisation = hidden_size, quantisation
        self._h = _P()
        rc = lib().wrnn_dm_create(int(hidden_size), int(quantisation), int(device), C.byref(self._h))
        if rc != 0:
            _raise(rc, lib().wrnn_dm_last_error(None))

    def close(self):
        if getattr(self, '_h', None):
            lib().wrnn_dm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            _raise(rc, lib().wrnn_dm_last_error(self._h))

    def load_weights(self, state):
        arr, keep = _tensors(state)
        self._ck(lib().wrnn_dm_load_weights(self._h, arr, len(state)))

    def generate(self, seq_len, n_seq=1, q_coarse=None, q_fine=None, seed=0, stream=None):
        """-> output, coarse, fine : int64 (n_seq, seq_len)."""
        qc = np.ascontiguousarray(q_coarse, dtype=np.float32) if q_coarse is not None else None
        qf = np.ascontiguousarray(q_fine, dtype=np.float32) if q_fine is not None else None
        for q in (qc, qf):
            if q is not None and q.shape != (seq_len, n_seq, self.quantisation):
                raise ValueError(f'noise must be ({seq_len}, {n_seq}, {self.quantisation}), got {q.shape}')
        out = np.empty((n_seq, seq_len), dtype=np.int64)
        c = np.empty((n_seq, seq_len), dtype=np.int64)
        f = np.empty((n_seq, seq_len), dtype=np.int64)
        self._ck(lib().wrnn_dm_generate(self._h, int(seq_len), int(n_seq), _ptr(qc), _ptr(qf), int(seed),
                                        _ptr(out), _ptr(c), _ptr(f), stream))
        return out, c, f

    def last_timing(self):
        ms, k = C.c_float(), C.c_int64()
        self._ck(lib().wrnn_dm_last_timing(self._h, C.byref(ms), C.byref(k)))
        return dict(ms_loop=ms.value, launches=k.value)


class TacotronDecoder:
    """Owns one `wrnn_taco_handle`: the decoder loop of `Tacotron.generate` (models/tacotron.py:438-465)."""

    def __init__(self, decoder_dims=256, lstm_dims=512, n_mels=80, device=0):
        self.decoder_dims, self.lstm_dims, self.n_mels = decoder_dims, lstm_dims, n_mels
        self._h = _P()
        rc = lib().wrnn_taco_create(int(decoder_dims), int(lstm_dims), int(n_mels), int(device), C.byref(self._h))
        if rc != 0:
            _raise(rc, lib().wrnn_taco_last_error(None))

    def close(self):
        if getattr(self, '_h', None):
            lib().wrnn_taco_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            _raise(rc, lib().wrnn_taco_last_error(self._h))

    def load_weights(self, state):
        """state: `Tacotron.state_dict()` (numpy arrays or torch tensors); the `decoder.*` and `encoder_proj.weight`
        entries are used."""
        state = {k: v for k, v in state.items() if k.startswith('decoder.') or k == 'encoder_proj.weight'}
        state = {k: (v.detach().cpu().numpy() if hasattr(v, 'detach') else np.asarray(v)) for k, v in state.items()}
        state = {k: np.ascontiguousarray(v, dtype=np.float32) for k, v in state.items() if v.dtype.kind == 'f'}
        arr, keep = _tensors(state)
        self._ck(lib().wrnn_taco_load_weights(self._h, arr, len(state)))

    def decode(self, enc_seq, steps=2000, r=2, stop_threshold=-3.4, stream=None):
        """enc_seq (L, decoder_dims) -> mel (n_mels, n_iter * r), attn (n_iter, L)."""
        enc = np.ascontiguousarray(enc_seq, dtype=np.float32)
        if enc.ndim != 2 or enc.shape[1] != self.decoder_dims:
            raise ValueError(f'enc_seq must be (L, {self.decoder_dims}), got {enc.shape}')
        L = enc.shape[0]
        max_iter = (int(steps) + int(r) - 1) // int(r)
        mel = np.empty(self.n_mels * max_iter * r, dtype=np.float32)
        attn = np.empty(max_iter * L, dtype=np.float32)
        n = C.c_int32(0)
        self._ck(lib().wrnn_taco_decode(self._h, _ptr(enc), L, int(r), int(steps), float(stop_threshold), _ptr(mel),
                                        _ptr(attn), C.byref(n), stream))
        n = n.value
        return mel[:self.n_mels * n * r].reshape(self.n_mels, n * r).copy(), attn[:n * L].reshape(n, L).copy()

    def last_timing(self):
        ms, k = C.c_float(), C.c_int64()
        self._ck(lib().wrnn_taco_last_timing(self._h, C.byref(ms), C.byref(k)))
        return dict(ms_loop=ms.value, launches=k.value)
