"""CPU oracle for the WaveRNN generation path.  TEST INFRASTRUCTURE ONLY.

A restatement of the reference's algorithm (`/root/reference/models/fatchord_version.py`,
`models/deepmind_version.py`, `utils/distribution.py`, `utils/dsp.py`) on torch-CPU fp32 and
numpy float64, with the two changes that make it usable as a checker:
  * the random draws are INJECTED (pre-drawn tensors, in the reference's draw order) instead of
    being drawn inside, so a CUDA run and an oracle run can consume identical noise;
  * weights are a plain dict of arrays in checkpoint (`state_dict`) layout.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference`
legs may import this package, and only as the checker or the timed CPU baseline.  The product
(`wavernn_b200/`) never imports it and has no CPU fallback.

Pinning: the reference ships no tests and no golden vectors.  The oracle is
pinned by running the reference's own Python in the build container
(`oracle/ref_shim.py` + `tests/golden/make_golden.py`), which (a) checks oracle == reference
bit-for-bit on CPU and (b) writes the reference's outputs to `tests/golden/*.npz`; the CPU test
suite re-checks the oracle against those fixtures everywhere (`/root/reference` does not exist
on the GPU box).  Third-party arithmetic (ATen `linear`, `gru_cell`, `conv1d`, `softmax`) is
called at the same call sites as the reference calls it.
"""
