#!/usr/bin/env python
"""Benchmark of the WaveRNN batched generation path (BASELINE.json: generated audio samples/s per B200 and whole box).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one full `generate` of the workload.  The headline workload (`value`, `e2e`) is BASELINE.json configs[1]:
fatchord WaveRNN MoL (rnn 512, fc 512, hop 275), batched target=11000 overlap=550 on a synthetic 10 s mel
(T=801 -> 20 folds x 12100 loop steps -> 220 000 samples), one utterance per GPU (weak scaling, no collective on the path).
  value : samples/s, device time (CUDA events) of stage-resident runs: conditioning + fold + per-sample loop + float64
          tail, inputs already in HBM.
  e2e   : samples/s through the public host call (wrnn_generate via ctypes): host mel in, host float64 waveform out,
          copies inside the timed region.
  roofline : the per-sample loop kernel of that workload: algorithmic weight bytes per loop step (15 331 448 B, fp32 MoL
          model) x steps / kernel time against measured HBM bandwidth, PLUS what actually binds
          the register-resident kernel: the exchange-latency floor and the fraction of it reached.
  config5 : the same line carries a second record, BASELINE.json configs[4] per GPU: 32 utterances of T=801 in ONE
          call (640 folds x 12100 steps; 256 utterances on 8 GPUs), through the tensor-core loop kernel (tcgen05, bf16
          hi/lo operands), with its own value / e2e / roofline (bound: tensor, against the measured sustained bf16 rate).
  tacotron : a third record, SURVEY.md section 8(f)-3: the decoder loop of Tacotron.generate on `taco_loop_kernel`
          (us per decoder step, mel frames/s, end to end with host buffers, the CPU restatement beside it).
  cpu_baseline : the reference's own un-patched generate() (staged copy, oracle/_ref) on the host cores, whole utterance.
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

T_FRAMES, TARGET, OVERLAP = 801, 11000, 550
WEIGHT_BYTES_PER_STEP = 15_331_448        # 3 825 152 matrix + 7 710 bias fp32 elements
FLOP_PER_FOLD_STEP = 7_650_304            # 2 x matrix elements (MoL)
CONFIG5_UTTERANCES = 32                   # per GPU: 256 utterances on the 8-GPU box
SAMPLE_RATE = 22050
METRIC = 'generated audio samples/sec'
# Exchange-latency floor of the resident kernel (measured with scripts/microbench*.cu on B200,
# profiles/r01_microbench_b200.txt): a step is five dependent L2 exchanges of ~2 400 clocks each (including the delayed
# first look) plus the fp32 FMA floor of the four products on the chain (153 MFLOP at 20 folds on 128 SMs = 3 900 clocks).
EXCHANGE_CLOCKS, EXCHANGES_PER_STEP, FMA_FLOOR_CLOCKS = 2400, 5, 3900


def loop_traffic(kernel, folds=None, steps=None):
    """DRAM bytes per launch of the loop kernel that ran, from the newest ncu capture summary under profiles/
    (`profiles/*_traffic.json`, written from `ncu --set full` captures by hand at the HEAD they name).  The captures
    run fewer steps than the benchmark (ncu replays a kernel ~40 times); DRAM traffic of these kernels is the
    conditioning rows, i.e. proportional to fold-steps, so the file holds bytes per fold-step and the per-launch figure
    is that times the folds x steps of the launch that was timed.  None when the kernel has no capture."""
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, 'profiles', '*_traffic.json'))):
        try:
            d = json.load(open(path))
        except Exception:
            continue
        if kernel in d:
            best = dict(d[kernel], file=os.path.relpath(path, ROOT))
    if best and folds and steps:
        best['dram_bytes_per_launch'] = int(best['dram_bytes_per_fold_step'] * folds * steps)
        best['source'] = (f"{best['dram_bytes_per_fold_step']} B per fold-step x {folds} folds x {steps} steps; capture: "
                          f"{best['capture']}, {best['source']}")
    return best


def measured_peaks():
    """-> (hbm GB/s, bf16 sustained TFLOP/s, source)."""
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.isfile(p):
        try:
            d = json.load(open(p))
            return float(d['hbm_gbs']), float(d.get('bf16_tflops_sustained', d['bf16_tflops'])), 'measured (MEASURED_PEAKS.json)'
        except Exception:
            pass
    return 6650.0, 1400.0, 'fallback (B200_PROFILING.md)'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                          '-i', str(self.index), '-lms', '100'], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith('active'):
                        reasons.add(n)
            except Exception:
                pass
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        return {'sm_mhz': float(np.median(sm)), 'sm_max_mhz': max(mx), 'reasons': sorted(reasons), 'samples': len(sm)}


def workload(rank=0):
    from wavernn_b200 import synth
    from wavernn_b200.config import VocoderDims
    dims = VocoderDims(mode='MOL')
    sd = synth.synth_state_dict(dims, seed=0, style='init')
    mel = synth.synth_mel(T_FRAMES, seed=1234 + rank)
    return dims, sd, mel


def config_dict(n_gpus):
    return {'workload': 'fatchord WaveRNN MoL rnn512 fc512 hop275 22.05kHz, batched generate target=11000 '
                        'overlap=550, synthetic 10 s mel T=801 (20 folds x 12100 steps), one utterance per GPU',
            'utterances_per_gpu': 1, 'folds_per_gpu': 20, 'loop_steps': 12100, 'samples_per_utterance': 220000,
            'weights': 'random-init seed 0 (pretrained ljspeech.wavernn.mol.800k unavailable on this box)',
            'noise': 'in-kernel Philox', 'parallelism': f'independent utterances x{n_gpus}, no collective',
            'l2': 'flushed between timed iterations (256 MiB write)'}


def cpu_generate(dims, sd, mel, threads, target=TARGET, overlap=OVERLAP):
    """One whole `generate()` of the workload on the host cores -> (samples/s, seconds, kind).

    kind "reference": the reference's own un-patched `WaveRNN.generate` (models/fatchord_version.py:169-264) from the
    copies `oracle/stage_ref.py` staged at build time (shims only: stub matplotlib / librosa, np.cumproduct, no-op
    save_wav and progress print), drawing its own noise.  kind "port": the oracle restatement, when the staged tree is
    absent.  Conditioning, all loop steps, fold / unfold are inside the timed region: nothing is extrapolated."""
    import torch
    from oracle import ref_shim
    torch.set_num_threads(threads)
    T = mel.shape[-1]
    wave_len = (T - 1) * dims.hop_length
    if ref_shim.available():
        model = _REF_MODEL.get('m')
        if model is None:
            model = _REF_MODEL['m'] = ref_shim.build_fatchord(dims, sd)
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            t0 = time.time()
            wave = model.generate(torch.as_tensor(mel), '/dev/null', True, target, overlap, False)
            dt = time.time() - t0
        kind = 'reference'
    else:
        from oracle import fatchord as ofc
        from wavernn_b200 import binding as _b, synth
        W = ofc.to_torch(sd)
        B = ofc.num_folds_for(T * dims.hop_length, target, overlap)
        noise = synth.synth_noise_mol(target + 2 * overlap, B, 4321)
        t0 = time.time()
        wave = ofc.generate(W, dims, mel, True, target, overlap, False, noise)
        dt = time.time() - t0
        kind = 'port'
    assert wave.shape == (wave_len,)
    return wave_len / dt, dt, kind


_REF_MODEL = {}


def run_reference(args):
    """--impl reference: the reference's own CPU generate() of the SAME utterance, whole, once per step."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    dims, sd, mel = workload(0)
    threads = os.cpu_count() or 1
    from wavernn_b200 import synth
    short = synth.synth_mel(41, seed=7)                 # warm-up steps: a 0.5 s utterance (thread pool, allocator, caches)
    for _ in range(max(args.warmup, 1)):
        cpu_generate(dims, sd, short, threads, target=1100, overlap=550)
    rates, total, kind = [], 0.0, 'port'
    for _ in range(args.steps):
        r, dt, kind = cpu_generate(dims, sd, mel, threads)
        rates.append(r)
        total += dt
    v = float(np.median(rates))
    sample = (f'whole utterance per step: conditioning + 12100 of 12100 loop steps at 20 folds + unfold, {args.steps} steps '
              f'({total:.1f} s of CPU work); warm-up steps run a T=41 utterance')
    cfg = config_dict(args.gpus)
    cfg['cpu_arm'] = 'ONE utterance on the host cores of rank 0, whatever --gpus is (the GPU arm runs one utterance per GPU)'
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': 'samples/s', 'rtf': v / SAMPLE_RATE,
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1000 * total / max(args.steps, 1),
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': cfg,
        'cpu_baseline': {'value': v, 'unit': 'samples/s', 'cores': threads, 'kind': kind, 'sample': sample},
        'e2e': {'value': v, 'unit': 'samples/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0}))


def run_config5(args, voc, dims, rank, world, stream, flush, barrier):
    """BASELINE.json configs[4], one GPU's share: 32 utterances of T=801 (seed 1234 + global utterance index) in ONE
    generate call = 640 folds x 12100 steps through the tensor-core loop kernel.  Same protocol as the headline record:
    device-timed runs with inputs resident, then end-to-end calls with host buffers."""
    import torch
    from wavernn_b200 import synth
    from wavernn_b200.shard import whole_job_rate
    n_utt = CONFIG5_UTTERANCES
    mels = [synth.synth_mel(T_FRAMES, seed=1234 + rank * n_utt + u)[0] for u in range(n_utt)]
    wave_len = (T_FRAMES - 1) * dims.hop_length
    k = max(1, min(args.steps, 3))
    voc.stage(mels, True, TARGET, OVERLAP, False, stream=stream)
    voc.run_staged(seed=50, stream=stream)                       # warm-up (weights / workspaces already packed)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
    loop_ms, launches = [], 0
    barrier()
    for i in range(k):
        flush.fill_(i & 0xff)
        ev[i][0].record()
        voc.run_staged(seed=500 + i, stream=stream)
        ev[i][1].record()
        ev[i][1].synchronize()
        t = voc.last_timing()
        loop_ms.append(t['ms_loop'])
        launches += t['launches']
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    waves = voc.fetch(stream=stream)
    assert len(waves) == n_utt and all(w.shape == (wave_len,) and np.all(np.isfinite(w)) for w in waves)
    voc.generate(mels, True, TARGET, OVERLAP, False, seed=3, stream=stream)
    barrier()
    t0 = time.time()
    for i in range(k):
        voc.generate(mels, True, TARGET, OVERLAP, False, seed=600 + i, stream=stream)
    barrier()
    e2e_s = time.time() - t0
    value, (dev_ms_max, e2e_ms_max) = whole_job_rate(n_utt * wave_len * k, [dev_ms, e2e_s * 1000.0], device='cuda')
    import torch
    dm = run_deepmind(torch.cuda.current_device(), world, barrier)
    if rank != 0:
        return None
    k_ms = float(np.mean(loop_ms))
    fold_steps = t['folds'] * t['steps'] / (k_ms * 1e-3)
    kernel = voc.loop_kernel_name(t['folds'])
    traffic = loop_traffic(kernel, t['folds'], t['steps'])
    return {
        'workload': f'{n_utt} utterances per GPU (= {n_utt * world} in the job), each a synthetic 10 s mel T=801, one batched '
                    f'generate call: {t["folds"]} folds x {t["steps"]} steps per GPU',
        'value': value, 'unit': 'samples/s', 'rtf': value / SAMPLE_RATE, 'steps': k, 'ms_per_step': dev_ms_max / k,
        'e2e': {'value': n_utt * wave_len * k * world / (e2e_ms_max / 1000.0), 'unit': 'samples/s',
                'h2d_bytes_per_step': int(sum(m.nbytes for m in mels)), 'd2h_bytes_per_step': int(n_utt * wave_len * 8),
                'ms_per_step': e2e_ms_max / k},
        'loop': {'kernel': kernel, 'ms': k_ms, 'us_per_step': 1000 * k_ms / t['steps'], 'folds': t['folds'], 'steps': t['steps'],
                 'ms_cond': t['ms_cond'], 'ms_tail': t['ms_tail'], 'fold_steps_per_s': fold_steps,
                 'fold_steps_per_s_resident_groups_r01': 1.75e6},
        'dtype': 'bf16 hi+lo split operands (3 MMAs per product), fp32 accumulate in TMEM, all five layers; gates and the MoL draw fp32',
        'roofline': {'bound': 'tensor', 'achieved': FLOP_PER_FOLD_STEP * fold_steps / 1e12, 'unit': 'TFLOP/s',
                     'algorithmic_flop_per_fold_step': FLOP_PER_FOLD_STEP,
                     'executed_flop_factor': 'x3 for the hi/lo split (the tensor pipe executes three bf16 MMAs per product); no padded rows',
                     'traffic': traffic['dram_bytes_per_launch'] if traffic else None,
                     'traffic_source': (traffic['source'] + ' via ' + traffic['file']) if traffic else
                     'no ncu capture of this kernel under profiles/ at this HEAD'},
        # north_star: "whole-box batched samples/sec as a fraction of the weight-bandwidth roofline": a step of the loop
        # re-reads every weight once whatever the fold count, so the bound is hbm_gbs / weight bytes steps per second
        'weight_bandwidth_roofline': {'achieved': WEIGHT_BYTES_PER_STEP * t['steps'] / (k_ms * 1e-3) / 1e9, 'unit': 'GB/s',
                                      'bound_steps_per_s': measured_peaks()[0] * 1e9 / WEIGHT_BYTES_PER_STEP,
                                      'steps_per_s': t['steps'] / (k_ms * 1e-3),
                                      'frac': WEIGHT_BYTES_PER_STEP * t['steps'] / (k_ms * 1e-3) / 1e9 / measured_peaks()[0],
                                      'note': 'the weights are streamed from L2 (15.3 MB of bf16 hi/lo images, resident there), not HBM; '
                                              'what binds is in roofline (tensor) and DESIGN.md 4.5'},
        # weight / operand read rate against the measured L2 -> SM streaming rate (profiles/r02_tcgen05_microbench.txt:
        # 69 B/clk/SM with all 148 SMs streaming 32 KB bulk copies).  Bytes per step from the kernel's job table: a CTA with a
        # GRU job streams 29 stages of 24 KB weights + 32 KB activations, an fc CTA 20 stages of 16 + 32 KB; 16 + 8 CTAs per tile
        'l2_stream': (lambda tiles, us: {
            'bytes_per_step': tiles * (16 * 29 * 57344 + 8 * 20 * 49152),
            'weight_bytes_per_step': tiles * (16 * 29 * 24576 + 8 * 20 * 16384),
            'achieved': tiles * (16 * 29 * 57344 + 8 * 20 * 49152) / (us * 1e-6) / 1e9, 'unit': 'GB/s',
            'peak': 69.0 * 148 * 1.965, 'peak_source': '69 B/clk/SM x 148 SMs x 1965 MHz, measured (profiles/r02_tcgen05_microbench.txt)',
            'frac': tiles * (16 * 29 * 57344 + 8 * 20 * 49152) / (us * 1e-6) / 1e9 / (69.0 * 148 * 1.965)})(
                -(-t['folds'] // 128), 1000 * k_ms / t['steps']),
        'deepmind': dm,
        'launches': launches + (dm['launches'] if dm else 0),
    }


def run_deepmind(device, world, barrier, n_seq=8, seconds=2.0):      # seq_len = 44 100
    """configs[4], second half: deepmind_version coarse/fine dual-softmax generation (models/deepmind_version.py:75-165),
    hidden 896, `n_seq` independent sequences per GPU (the reference generates one), Philox noise, through the C ABI with
    the int64 outputs copied to host arrays inside the timed call."""
    from wavernn_b200 import binding, synth
    from wavernn_b200.shard import whole_job_rate
    import torch
    steps = int(SAMPLE_RATE * seconds)
    v = binding.DeepmindVocoder(896, 256, device=device)
    v.load_weights(synth.synth_dm_state_dict(896, 256, seed=0))
    v.generate(steps // 8, n_seq, seed=1)                      # warm-up: weight images, workspaces
    barrier()
    t0 = time.time()
    out, c, f = v.generate(steps, n_seq, seed=2)
    torch.cuda.synchronize()
    dt = time.time() - t0
    t = v.last_timing()
    assert out.shape == (n_seq, steps) and out.min() >= -32768 and out.max() <= 32767
    rate, (ms_max,) = whole_job_rate(n_seq * steps, [dt * 1000.0], device='cuda')
    v.close()
    return {'workload': f'{n_seq} sequences per GPU x {steps} samples, hidden 896, 16-bit coarse/fine', 'value': rate, 'unit': 'samples/s',
            'rtf': rate / SAMPLE_RATE, 'ms': ms_max, 'loop_us_per_step': 1000 * t['ms_loop'] / steps, 'launches': int(t['launches']),
            'timed': 'wall clock around the C-ABI call, host int64 outputs included'}


def run_tacotron(device, cpu_leg, L=150, r=2, frames=1600):
    """the decoder loop of `Tacotron.generate` (models/tacotron.py:438-465) on `taco_loop_kernel`
    through the C ABI: encoder output (host) in, mel frames and attention rows (host) out, no stop (threshold -inf) so
    that the step count is fixed.  `cpu` = the oracle restatement (the reference's own ATen calls, bit-equal to the
    reference's generate on CPU) on the host cores for a bounded number of the same steps."""
    from wavernn_b200 import binding, synth
    sd = synth.synth_taco_state_dict(seed=0)
    enc = synth.synth_encoder_seq(L, seed=78)
    dec = binding.TacotronDecoder(device=device)
    dec.load_weights(sd)
    dec.decode(enc, steps=64, r=r, stop_threshold=-1e30)                       # warm-up: packing, workspace
    ms, wall = [], []
    for _ in range(3):
        t0 = time.time()
        mel, attn = dec.decode(enc, steps=frames, r=r, stop_threshold=-1e30)
        wall.append(time.time() - t0)
        ms.append(dec.last_timing()['ms_loop'])
    n = attn.shape[0]
    assert mel.shape == (80, n * r) and n == frames // r and np.all(np.isfinite(mel))
    rec = {'workload': f'one utterance, {L} characters, r = {r}, {n} decoder steps = {n * r} mel frames (no stop), decoder_dims 256, '
                       f'lstm_dims 512 (hparams.py:67-73)',
           'kernel': 'taco_loop_kernel', 'us_per_step': 1000 * min(ms) / n, 'frames_per_s': n * r / (min(ms) * 1e-3),
           'rtf': n * r * 275 / SAMPLE_RATE / (min(ms) * 1e-3),
           'e2e': {'frames_per_s': n * r / min(wall), 'ms': 1000 * min(wall), 'h2d_bytes': int(enc.nbytes),
                   'd2h_bytes': int(mel.nbytes + attn.nbytes)},
           'launches': 2 * 4, 'dtype': 'f32'}
    # what bounds it: seven dependent hand-overs per decoder step (tagged words through L2), each at best one
    # store propagation + one poll round trip (1 300 clk measured with the 450-clk delayed look) + a one-warp product and
    # its publish (350 clk); the weights never leave shared memory, so HBM sees the 26 MB image once per launch
    sm_mhz = 1965.0
    floor_us = 7 * (1300 + 350) / sm_mhz
    tr = None
    for path in sorted(glob.glob(os.path.join(ROOT, 'profiles', '*_traffic.json'))):
        try:
            d = json.load(open(path))
        except Exception:
            continue
        if 'taco_loop_kernel' in d:
            tr = dict(d['taco_loop_kernel'], file=os.path.relpath(path, ROOT))
    weight_bytes = 20_680_960        # the 5 170 240 fp32 decoder parameters a step touches at r = 2
    rec['roofline'] = {'bound': 'latency', 'hand_overs_per_step': 7, 'latency_floor_us': floor_us,
                       'frac_of_floor': floor_us / rec['us_per_step'],
                       'algorithmic_weight_bytes_per_step': weight_bytes, 'weights': 'resident in shared memory for the whole call',
                       'shared_memory_weight_rate_GBps': weight_bytes / (rec['us_per_step'] * 1e-6) / 1e9,
                       'traffic': tr['dram_bytes_captured'] if tr else None,
                       'traffic_source': (tr['capture'] + ', ' + tr['source'] + ' via ' + tr['file']) if tr else None}
    dec.close()
    if cpu_leg:
        from oracle import tacotron as otc
        import torch
        torch.set_num_threads(os.cpu_count() or 1)
        otc.decode(sd, enc, steps=8, r=r, stop_threshold=-1e30)
        t0 = time.time()
        cpu_steps = 200
        otc.decode(sd, enc, steps=cpu_steps * r, r=r, stop_threshold=-1e30)
        dt = time.time() - t0
        rec['cpu'] = {'us_per_step': 1e6 * dt / cpu_steps, 'frames_per_s': cpu_steps * r / dt, 'cores': os.cpu_count() or 1,
                      'kind': 'port', 'sample': f'{cpu_steps} of the {n} decoder steps ({dt:.2f} s)'}
    return rec


def run_ours(args):
    import torch
    import torch.distributed as dist
    from wavernn_b200 import binding
    from wavernn_b200.shard import whole_job_rate

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device and no CPU fallback for the product path')
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    dims, sd, mel = workload(rank)
    voc = binding.Vocoder(dims, device=local)
    voc.load_weights(sd)
    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device='cuda')
    wave_len = (T_FRAMES - 1) * dims.hop_length

    # ---------------- value: inputs resident in HBM, device time by CUDA events ----------------
    voc.stage([mel[0]], True, TARGET, OVERLAP, False, stream=stream)
    for i in range(args.warmup):
        voc.run_staged(seed=100 + i, stream=stream)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    loop_ms, launches = [], 0
    barrier()
    t_wall = time.time()
    for i in range(args.steps):
        flush.fill_(i & 0xff)                      # L2 flush between timed iterations (not timed)
        ev[i][0].record()
        voc.run_staged(seed=1000 + i, stream=stream)
        ev[i][1].record()
        ev[i][1].synchronize()
        t = voc.last_timing()
        loop_ms.append(t['ms_loop'])
        launches += t['launches']
    barrier()
    t_wall = time.time() - t_wall
    folds_loop, steps_loop = t['folds'], t['steps']
    loop_kernel = voc.loop_kernel_name(folds_loop)
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    clocks = sampler.stop() if rank == 0 else None
    wave = voc.fetch(stream=stream)[0]
    assert wave.shape == (wave_len,) and np.all(np.isfinite(wave))

    # ---------------- e2e: host buffers in, host waveform out, through the public call ----------
    for i in range(min(args.warmup, 2)):
        voc.generate([mel[0]], True, TARGET, OVERLAP, False, seed=i, stream=stream)
    barrier()
    t0 = time.time()
    for i in range(args.steps):
        out = voc.generate([mel[0]], True, TARGET, OVERLAP, False, seed=2000 + i, stream=stream)[0]
    barrier()
    e2e_s = time.time() - t0

    # ---------------- config 5 per GPU: 32 utterances in one call, tensor-core loop kernel ----------------
    c5 = None
    if not args.no_config5:
        c5 = run_config5(args, voc, dims, rank, world, stream, flush, barrier)

    # whole-job rate: samples of all ranks over the slowest rank's time (wavernn_b200/shard.py, gloo-tested on CPU)
    value, (dev_ms_max, e2e_ms_max) = whole_job_rate(wave_len * args.steps, [dev_ms, e2e_s * 1000.0], device='cuda')
    total_samples = wave_len * args.steps * world
    e2e = total_samples / (e2e_ms_max / 1000.0)
    if rank == 0:
        hbm, bf16, peak_src = measured_peaks()
        t_last = t
        k_ms = float(np.mean(loop_ms))
        kernel = loop_kernel
        us_step = 1000 * k_ms / steps_loop
        achieved = WEIGHT_BYTES_PER_STEP * steps_loop / (k_ms * 1e-3) / 1e9
        sm_mhz = (clocks or {}).get('sm_mhz') or 1965.0
        floor_us = (EXCHANGES_PER_STEP * EXCHANGE_CLOCKS + FMA_FLOOR_CLOCKS) / sm_mhz
        traffic = loop_traffic(kernel, folds_loop, steps_loop)
        line = {
            'metric': METRIC, 'value': value, 'unit': 'samples/s', 'rtf': value / SAMPLE_RATE / world,
            'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dev_ms_max / args.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': config_dict(world),
            'e2e': {'value': e2e, 'unit': 'samples/s', 'rtf_per_gpu': e2e / SAMPLE_RATE / world,
                    'h2d_bytes_per_step': int(mel[0].nbytes), 'd2h_bytes_per_step': int(wave_len * 8),
                    'ms_per_step': e2e_ms_max / args.steps},
            'gpu_launches': int(launches) + (c5['launches'] if c5 else 0),
            'loop': {'kernel': kernel, 'ms': k_ms, 'us_per_step': us_step,
                     'folds': folds_loop, 'steps': steps_loop, 'ms_cond': t_last['ms_cond'], 'ms_tail': t_last['ms_tail'],
                     'fold_steps_per_s': folds_loop * steps_loop / (k_ms * 1e-3),
                     'exchange_calibration': dict(zip(('l2_round_trip_clk', 'gather_delay_scale'), voc.exchange_calibration()))},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': hbm, 'unit': 'GB/s', 'frac': achieved / hbm,
                         'traffic': traffic['dram_bytes_per_launch'] if traffic else None,
                         'traffic_source': (traffic['source'] + ' via ' + traffic['file']) if traffic else
                         'no ncu capture of this kernel under profiles/ at this HEAD',
                         'algorithmic_bytes_per_launch': WEIGHT_BYTES_PER_STEP * steps_loop, 'peak_source': peak_src,
                         'binding_limit': 'exchange latency (five dependent L2 exchanges per step), then the LSU pipe',
                         'latency_floor_us': floor_us, 'frac_of_floor': floor_us / us_step,
                         'latency_floor': f'{EXCHANGES_PER_STEP} exchanges x {EXCHANGE_CLOCKS} clk + {FMA_FLOOR_CLOCKS} clk fp32 FMA '
                                          f'floor at {sm_mhz:.0f} MHz (DESIGN.md 4.1, profiles/r01_microbench_b200.txt)',
                         'fp32_fma_frac': folds_loop * FLOP_PER_FOLD_STEP / (us_step * 1e-6) / (148 * 128 * 2 * sm_mhz * 1e6),
                         'note': 'algorithmic = fp32 weight bytes re-read once per loop step; the weights are register-resident, '
                                 'so DRAM traffic is the conditioning rows only and HBM does not bind: frac_of_floor is the '
                                 'figure that says how close the kernel is to what bounds it'},
            'clocks': clocks, 'wall_s_value_leg': t_wall,
        }
        if c5:
            c5.pop('launches', None)
            c5['roofline']['peak'] = bf16
            c5['roofline']['frac'] = c5['roofline']['achieved'] / bf16
            c5['roofline']['peak_source'] = peak_src + ' bf16_tflops_sustained'
            line['config5'] = c5
        if not args.no_tacotron:
            taco = run_tacotron(torch.cuda.current_device(), world == 1 and not args.no_cpu)
            line['gpu_launches'] += taco.pop('launches')
            line['tacotron'] = taco
        if world == 1 and not args.no_cpu:
            threads = os.cpu_count() or 1
            from wavernn_b200 import synth
            cpu_generate(dims, sd, synth.synth_mel(41, seed=7), threads, target=1100, overlap=550)     # warm-up
            runs = [cpu_generate(dims, sd, mel, threads) for _ in range(2)]
            v = float(np.median([r[0] for r in runs]))
            line['cpu_baseline'] = {'value': v, 'unit': 'samples/s', 'cores': threads, 'kind': runs[0][2],
                                    'sample': 'whole utterance, twice (median): conditioning + 12100 of 12100 loop steps at 20 '
                                              f'folds + unfold ({sum(r[1] for r in runs):.1f} s of CPU work)'}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--no-config5', action='store_true', help='skip the 32-utterance record')
    ap.add_argument('--no-tacotron', action='store_true', help='skip the Tacotron decoder-loop record')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
