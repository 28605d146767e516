// C ABI, part 2: stage -> run -> fetch (and wrnn_generate = the three in sequence).
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include <thread>

#include "internal.h"

namespace wrnn {

struct StageOpts {
    bool prepadded = false;        // mels already carry `pad` frames on both sides (forward_logits)
    const float *teacher = nullptr;   // host [B][S]
    bool want_logits = false;
    // chunked delivery (wrnn_generate_stream): fold groups of at most `chunk_folds`; after each group the part of the
    // waveform that is final goes to the host and `chunk_events` / `chunk_end` say when and how far
    int64_t chunk_folds = 0;
    std::vector<cudaEvent_t> *chunk_events = nullptr;
    std::vector<int64_t> *chunk_end = nullptr;
};

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Layout of ws_meta (device): FoldMeta[B] | TailMeta[n_utt] | int32 fold_len[B] | UttMeta[n_utt] | ResTile[n_tiles]
struct MetaLayout { size_t off_fold, off_tail, off_len, off_utt, off_tile, bytes; };
static MetaLayout meta_layout(int64_t B, int n_utt, int64_t n_tiles) {
    MetaLayout m;
    m.off_fold = 0;
    m.off_tail = align_up(sizeof(FoldMeta) * B, 64);
    m.off_len = m.off_tail + align_up(sizeof(TailMeta) * n_utt, 64);
    m.off_utt = m.off_len + align_up(sizeof(int32_t) * B, 64);
    m.off_tile = m.off_utt + align_up(sizeof(UttMeta) * n_utt, 64);
    m.bytes = m.off_tile + align_up(sizeof(ResTile) * (size_t)n_tiles, 64);
    return m;
}
// MelResNet tiles an utterance needs: all frames, or, under a fold range, only the frames its folds cover
static void res_tile_span(const wrnn_handle *h, const wrnn_handle::Utt &ut, int *t_first, int *t_end) {
    const int TT = resnet_tile_frames();
    *t_first = 0;
    *t_end = ut.T;
    if (h->Btot != h->B && h->batched) {
        const int64_t stride = h->target + h->overlap, s0 = h->fold_lo * stride, s1 = (h->fold_lo + h->B - 1) * stride + h->S;
        *t_first = (int)(s0 / h->d.hop / TT * TT);
        *t_end = (int)std::min<int64_t>(ut.T, s1 / h->d.hop + 1);
    }
}
static int64_t count_res_tiles(const wrnn_handle *h) {
    int64_t n = 0;
    for (auto &ut : h->utts) {
        int a, b;
        res_tile_span(h, ut, &a, &b);
        n += (b - a + resnet_tile_frames() - 1) / resnet_tile_frames();
    }
    return n;
}
// Layout of ws_tables (device doubles): fade_in[ov] | fade_out[ov] | fo[20*hop] | lut[NC]
struct TableLayout { size_t fin, fout, fo, lut, n; };
static TableLayout table_layout(const wrnn_handle *h) {
    TableLayout t;
    t.fin = 0;
    t.fout = t.fin + (size_t)std::max<int64_t>(h->overlap, 1);
    t.fo = t.fout + (size_t)std::max<int64_t>(h->overlap, 1);
    t.lut = t.fo + (size_t)20 * h->d.hop;
    t.n = t.lut + (size_t)h->d.NC;
    return t;
}

int do_stage(wrnn_handle *h, const float *const *mels, const int32_t *T, int32_t n_utt, int32_t batched,
             int32_t target, int32_t overlap, int32_t mu_law, const wrnn_noise *noise, cudaStream_t st,
             const StageOpts &o) {
    const Dims &d = h->d;
    h->staged = false;
    if (!mels || !T || n_utt < 1) return fail(h, WRNN_ERR_ARG, "need at least one utterance");
    WRNN_DEVICE(h);
    WRNN_TRY(ensure_packed_public(h));
    if (batched && (target < 1 || overlap < 0)) return fail(h, WRNN_ERR_ARG, "target must be >= 1 and overlap >= 0");
    h->batched = batched ? 1 : 0;
    h->target = target;
    h->overlap = batched ? overlap : 0;
    h->mu_law = (d.mode == WRNN_MODE_RAW && mu_law) ? 1 : 0;     // fatchord_version.py:174
    h->utts.clear();
    int64_t mel_off = 0, up_off = 0, res_off = 0, out_off = 0, fold0 = 0, Smax = 0, pm_off = 0, range_total = -1, range_first = 0;
    for (int u = 0; u < n_utt; u++) {
        wrnn_handle::Utt ut;
        const int Tin = T[u];
        ut.T = o.prepadded ? Tin - 2 * d.pad : Tin;
        if (!mels[u] || ut.T < 1) return fail(h, WRNN_ERR_ARG, "utterance %d: empty mel", u);
        ut.L = (int64_t)ut.T * d.hop;
        ut.wave_len = (int64_t)(ut.T - 1) * d.hop;
        // the reference's fade-out multiplies the last 20*hop samples (:256-258) and raises on
        // shorter output; same condition, raised early
        if (!o.want_logits && ut.wave_len < 20 * (int64_t)d.hop)
            return fail(h, WRNN_ERR_ARG, "utterance %d: mel too short (T=%d, need T >= 21 for the 20-hop fade-out)", u, ut.T);
        if (batched) {
            if (ut.L <= overlap) return fail(h, WRNN_ERR_ARG, "utterance %d: shorter than the overlap", u);
            ut.nfolds = wrnn_num_folds(ut.L, target, overlap);
        } else {
            ut.nfolds = 1;
            Smax = std::max(Smax, ut.L);
        }
        if (h->range_hi > h->range_lo && !o.want_logits) {
            // fold-range sharding: this handle runs folds [lo, hi) of the one utterance
            if (n_utt != 1 || !batched) return fail(h, WRNN_ERR_ARG, "a fold range needs one utterance and batched generation");
            range_total = ut.nfolds;
            range_first = std::min(h->range_lo, ut.nfolds);
            ut.nfolds = std::min(h->range_hi, range_total) - range_first;
            if (ut.nfolds < 1) return fail(h, WRNN_ERR_ARG, "fold range [%lld, %lld) is empty: the utterance has %lld folds",
                                           (long long)h->range_lo, (long long)h->range_hi, (long long)range_total);
        }
        ut.fold0 = fold0; fold0 += ut.nfolds;
        ut.mel_off = mel_off; mel_off += (int64_t)d.F * Tin;
        ut.Tin = Tin; ut.Tp = ut.T + 2 * d.pad; ut.pm_off = pm_off; pm_off += ut.Tp;
        ut.up_off = up_off; up_off += ut.L;
        ut.res_off = res_off; res_off += ut.T;
        ut.out_off = out_off; out_off += ut.wave_len;
        h->utts.push_back(ut);
    }
    h->B = fold0;
    h->fold_lo = range_total >= 0 ? range_first : 0;
    h->Btot = range_total >= 0 ? range_total : fold0;
    h->S = batched ? (int64_t)target + 2 * (int64_t)overlap : Smax;
    h->total_out = out_off;
    if (h->B > 0x7fffffff / 4 || h->S > 0x7fffffff) return fail(h, WRNN_ERR_ARG, "too many folds / steps");

    // ---- mels: host -> pinned -> device (raw region, then a padded region filled by run) ----
    const size_t mel_bytes = (size_t)mel_off * sizeof(float);
    if (h->h_pinned_in_bytes < mel_bytes) {
        if (h->h_pinned_in) cudaFreeHost(h->h_pinned_in);
        h->h_pinned_in = nullptr;
        WRNN_CUDA(h, cudaMallocHost((void **)&h->h_pinned_in, mel_bytes + mel_bytes / 4));
        h->h_pinned_in_bytes = mel_bytes + mel_bytes / 4;
    }
    for (int u = 0; u < n_utt; u++)
        memcpy(h->h_pinned_in + h->utts[u].mel_off, mels[u], (size_t)d.F * T[u] * sizeof(float));
    h->total_frames = pm_off;
    const size_t padded_floats = (size_t)pm_off * d.F;       // frame-major padded mels, filled by run
    WRNN_TRY(ws_reserve(h, h->ws_mel, mel_bytes + padded_floats * sizeof(float) + 512));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_mel.p, h->h_pinned_in, mel_bytes, cudaMemcpyHostToDevice, st));

    // ---- meta ----
    const int64_t n_tiles = count_res_tiles(h);
    const MetaLayout ml = meta_layout(h->B, n_utt, n_tiles);
    std::vector<char> hm(ml.bytes, 0);
    FoldMeta *fm = (FoldMeta *)(hm.data() + ml.off_fold);
    TailMeta *tm = (TailMeta *)(hm.data() + ml.off_tail);
    int32_t *fl = (int32_t *)(hm.data() + ml.off_len);
    UttMeta *um = (UttMeta *)(hm.data() + ml.off_utt);
    ResTile *rt = (ResTile *)(hm.data() + ml.off_tile);
    int64_t ti = 0;
    for (int u = 0; u < n_utt; u++) {
        const auto &ut = h->utts[u];
        for (int64_t i = 0; i < ut.nfolds; i++) {
            fm[ut.fold0 + i] = FoldMeta{ut.pm_off, ut.res_off, batched ? (h->fold_lo + i) * ((int64_t)target + overlap) : 0, ut.L, ut.Tp, d.pad};
            fl[ut.fold0 + i] = (int32_t)(batched ? h->S : ut.L);
        }
        tm[u] = TailMeta{ut.fold0, ut.nfolds, ut.wave_len, ut.out_off};
        // already-padded input (forward_logits): all Tp frames come from the caller, no zero frames added
        um[u] = UttMeta{ut.mel_off, ut.pm_off, ut.res_off, ut.T, ut.Tp, ut.Tin, o.prepadded ? 0 : d.pad};
        int tf, te;
        res_tile_span(h, ut, &tf, &te);
        for (int t0 = tf; t0 < te; t0 += resnet_tile_frames()) rt[ti++] = ResTile{u, t0};
    }
    WRNN_TRY(ws_reserve(h, h->ws_meta, ml.bytes));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_meta.p, hm.data(), ml.bytes, cudaMemcpyHostToDevice, st));

    // ---- float64 tables ----
    const TableLayout tl = table_layout(h);
    std::vector<double> tab(tl.n, 0.0);
    if (batched) wrnn_xfade_tables(overlap, tab.data() + tl.fin, tab.data() + tl.fout);
    host_linspace(1.0, 0.0, 20 * (int64_t)d.hop, tab.data() + tl.fo);          // :256
    if (d.mode == WRNN_MODE_RAW) {
        if ((int)h->lut_override.size() == d.NC) memcpy(tab.data() + tl.lut, h->lut_override.data(), d.NC * sizeof(double));
        else wrnn_mulaw_lut(d.NC, tab.data() + tl.lut);
    }
    WRNN_TRY(ws_reserve(h, h->ws_tables, tl.n * sizeof(double)));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_tables.p, tab.data(), tl.n * sizeof(double), cudaMemcpyHostToDevice, st));

    // ---- injected noise ----
    h->have_noise = false;
    if (noise) {
        if (noise->steps != h->S || noise->folds != h->Btot)       // under a fold range the noise is that of the whole utterance
            return fail(h, WRNN_ERR_ARG, "noise is %lld x %lld but the call runs %lld steps x %lld folds",
                        (long long)noise->steps, (long long)noise->folds, (long long)h->S, (long long)h->Btot);
        const size_t SB = (size_t)h->S * h->Btot;
        if (d.mode == WRNN_MODE_MOL) {
            if (!noise->u1 || !noise->u2) return fail(h, WRNN_ERR_ARG, "MOL noise needs u1 and u2");
            const size_t nm = d.NC / 3;
            WRNN_TRY(ws_reserve(h, h->ws_noise, SB * (nm + 1) * sizeof(float)));
            WRNN_CUDA(h, cudaMemcpyAsync(h->ws_noise.p, noise->u1, SB * nm * sizeof(float), cudaMemcpyHostToDevice, st));
            WRNN_CUDA(h, cudaMemcpyAsync((float *)h->ws_noise.p + SB * nm, noise->u2, SB * sizeof(float),
                                         cudaMemcpyHostToDevice, st));
        } else {
            if (!noise->q) return fail(h, WRNN_ERR_ARG, "RAW noise needs q");
            WRNN_TRY(ws_reserve(h, h->ws_noise, SB * d.NC * sizeof(float)));
            WRNN_CUDA(h, cudaMemcpyAsync(h->ws_noise.p, noise->q, SB * d.NC * sizeof(float), cudaMemcpyHostToDevice, st));
        }
        h->have_noise = true;
    }
    if (o.teacher) {
        WRNN_TRY(ws_reserve(h, h->ws_teacher, (size_t)h->B * h->S * sizeof(float)));
        WRNN_CUDA(h, cudaMemcpyAsync(h->ws_teacher.p, o.teacher, (size_t)h->B * h->S * sizeof(float),
                                     cudaMemcpyHostToDevice, st));
    }
    // ---- device workspaces for run ----
    WRNN_TRY(ws_reserve(h, h->ws_res, (size_t)res_off * d.R * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_samples, (size_t)h->B * h->S * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_picks, (size_t)h->B * h->S * sizeof(int32_t)));
    WRNN_TRY(ws_reserve(h, h->ws_out, (size_t)std::max<int64_t>(h->total_out, 1) * sizeof(double)));
    if (o.want_logits) WRNN_TRY(ws_reserve(h, h->ws_logits, (size_t)h->B * h->S * d.NC * sizeof(float)));
    h->staged = true;
    return 0;
}

int do_run(wrnn_handle *h, uint64_t seed, cudaStream_t st, const StageOpts &o) {
    if (!h->staged) return fail(h, WRNN_ERR_STATE, "nothing staged: call wrnn_stage first");
    const Dims &d = h->d;
    WRNN_DEVICE(h);
    h->launches = 0;
    const int n_utt = (int)h->utts.size();
    const int64_t n_tiles = count_res_tiles(h);
    const MetaLayout ml = meta_layout(h->B, n_utt, n_tiles);
    const TableLayout tl = table_layout(h);
    float *mel_raw = (float *)h->ws_mel.p;
    size_t raw_floats = 0;
    for (auto &ut : h->utts) raw_floats += (size_t)d.F * ut.Tin;
    float *melT = mel_raw + align_up(raw_floats, 64);
    const UttMeta *d_utt = (const UttMeta *)((char *)h->ws_meta.p + ml.off_utt);
    WRNN_CUDA(h, cudaEventRecord(h->ev[0], st));
    // conditioning of every utterance of the call (fatchord_version.py:185-190) in three launches: pad + transpose,
    // MelResNet, and the fold rows [mel (polyphase upsampling) | aux (x hop repeat by index)]
    WRNN_TRY(launch_pad_mel_batch(h, mel_raw, d_utt, n_utt, h->total_frames, melT, st));
    WRNN_TRY(launch_resnet_batch(h, melT, d_utt, (const ResTile *)((char *)h->ws_meta.p + ml.off_tile), (int)n_tiles,
                                 (float *)h->ws_res.p, st));
    const bool chunked = o.chunk_folds > 0;
    const int kernel = loop_kernel_choice(h, chunked ? std::min<int64_t>(h->B, o.chunk_folds) : h->B);
    const FoldMeta *d_fold = (const FoldMeta *)((char *)h->ws_meta.p + ml.off_fold);
    if (kernel != 4) {      // the tensor-core kernel lays the features out as its own MMA operand (loop_mma.cu): no fp32 fold batch
        WRNN_TRY(ws_reserve(h, h->ws_cond, (size_t)h->B * h->S * (d.F + d.R) * sizeof(float)));
        WRNN_TRY(launch_fold_cond(h, melT, (const float *)h->ws_res.p, d_fold, h->B, h->S, (float *)h->ws_cond.p, st));
    }
    WRNN_CUDA(h, cudaEventRecord(h->ev[1], st));
    // per-sample loop (:201-241)
    LoopArgs a;
    a.cond = kernel != 4 ? (const float *)h->ws_cond.p : nullptr;
    a.melT = melT;
    a.res_out = (const float *)h->ws_res.p;
    a.fold_meta = d_fold;
    a.B = h->B;
    a.S = h->S;
    a.fold_len = (const int32_t *)((char *)h->ws_meta.p + ml.off_len);
    a.u1 = a.u2 = a.q = nullptr;
    if (h->have_noise) {
        if (d.mode == WRNN_MODE_MOL) {
            a.u1 = (const float *)h->ws_noise.p;
            a.u2 = a.u1 + (size_t)h->S * h->Btot * (d.NC / 3);
        } else {
            a.q = (const float *)h->ws_noise.p;
        }
    }
    a.teacher = o.teacher ? (const float *)h->ws_teacher.p : nullptr;
    a.logits_out = o.want_logits ? (float *)h->ws_logits.p : nullptr;
    a.samples = (float *)h->ws_samples.p;
    a.picks = (int32_t *)h->ws_picks.p;
    a.seed = seed;
    a.Btot = h->Btot;
    a.b0 = h->fold_lo;
    // barrier counter (bytes 0..63) is cleared per run; the abort flag (byte 64) is STICKY: it stays set across runs
    // until a fetch has reported it (check_abort), so a benchmark that runs N times and fetches once cannot count an
    // aborted (early-exited) iteration as throughput
    const bool fresh_sync = h->ws_sync.p == nullptr;
    WRNN_TRY(ws_reserve(h, h->ws_sync, 128));
    WRNN_CUDA(h, cudaMemsetAsync(h->ws_sync.p, 0, fresh_sync ? 128 : 64, st));
    const double *tab = (const double *)h->ws_tables.p;
    if (kernel != 1 || chunked) {
        // these kernels take fold groups of <= 20 (folds are independent: any grouping is exact)
        // (the tensor-core kernel takes up to 1 152 folds per launch: one (row tile, fold tile) job per CTA)
        const int64_t cap0 = kernel == 4 ? mma_max_folds(h) : kernel == 3 ? res_max_folds() : kernel == 2 ? pipe_max_folds() : h->B,
                      cap = chunked ? std::min(cap0, o.chunk_folds) : cap0, ng = (h->B + cap - 1) / cap,
                      per = chunked ? cap : (h->B + ng - 1) / ng;
        int64_t delivered = 0;          // samples of the waveform already final and on their way to the host
        const size_t CW = (size_t)(d.F + d.R);
        for (int64_t b0 = 0; b0 < h->B; b0 += per) {
            LoopArgs g = a;
            g.B = std::min<int64_t>(per, h->B - b0);
            g.b0 = h->fold_lo + b0;
            g.cond = a.cond ? a.cond + (size_t)b0 * h->S * CW : nullptr;
            g.fold_meta = d_fold + b0;
            g.fold_len = a.fold_len + b0;
            g.samples = a.samples + (size_t)b0 * h->S;
            g.picks = a.picks + (size_t)b0 * h->S;
            if (a.teacher) g.teacher = a.teacher + (size_t)b0 * h->S;
            if (a.logits_out) g.logits_out = a.logits_out + (size_t)b0 * h->S * d.NC;
            WRNN_TRY(kernel == 4 ? mma_run(h, g, st) : kernel == 3 ? res_run(h, g, st) : kernel == 2 ? pipe_run(h, g, st) : fma_run(h, g, st));
            if (chunked) {
                // fatchord_version.py:374-405 adds at most fold b - 1 and fold b into sample i, b = i / (target + overlap):
                // with folds [0, done) through, everything below done * (target + overlap) is final (the fade-out of the
                // last 20 hops, :255-258, lies in the last chunk)
                const auto &ut = h->utts[0];
                const int64_t done = b0 + g.B, stride = h->target + h->overlap;
                const int64_t end = done >= h->B ? ut.wave_len : std::min<int64_t>(done * stride, ut.wave_len);
                if (end > delivered) {
                    WRNN_TRY(launch_tail(h, a.samples, a.picks, (const TailMeta *)((char *)h->ws_meta.p + ml.off_tail), 1, end, h->S,
                                         h->batched, h->target, h->overlap, tab + tl.fin, tab + tl.fout, tab + tl.fo,
                                         20 * (int64_t)d.hop, h->mu_law ? tab + tl.lut : nullptr, (double *)h->ws_out.p, st, delivered));
                    WRNN_CUDA(h, cudaMemcpyAsync(h->h_pinned_out + delivered, (double *)h->ws_out.p + delivered,
                                                 (size_t)(end - delivered) * sizeof(double), cudaMemcpyDeviceToHost, st));
                }
                cudaEvent_t ev;
                WRNN_CUDA(h, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
                WRNN_CUDA(h, cudaEventRecord(ev, st));
                o.chunk_events->push_back(ev);
                o.chunk_end->push_back(end);
                delivered = end;
            }
        }
    } else {
        WRNN_TRY(fma_run(h, a, st));
    }
    WRNN_CUDA(h, cudaEventRecord(h->ev[2], st));
    // float64 tail (:243-258)
    if (!o.want_logits && h->Btot == h->B && !chunked) {      // under a fold range the float64 tail belongs to whoever gathers the shards
        int64_t max_wave = 0;
        for (auto &ut : h->utts) max_wave = std::max(max_wave, ut.wave_len);
        WRNN_TRY(launch_tail(h, a.samples, a.picks, (const TailMeta *)((char *)h->ws_meta.p + ml.off_tail), n_utt,
                             max_wave, h->S, h->batched, h->target, h->overlap, tab + tl.fin, tab + tl.fout,
                             tab + tl.fo, 20 * (int64_t)d.hop, h->mu_law ? tab + tl.lut : nullptr,
                             (double *)h->ws_out.p, st));
    }
    WRNN_CUDA(h, cudaEventRecord(h->ev[3], st));
    return 0;
}

int check_abort(wrnn_handle *h, cudaStream_t st) {
    int flag = 0;
    if (h->ws_sync.p) {
        WRNN_CUDA(h, cudaMemcpyAsync(&flag, (char *)h->ws_sync.p + 64, sizeof(int), cudaMemcpyDeviceToHost, st));
        WRNN_CUDA(h, cudaStreamSynchronize(st));
    }
    if (flag) {
        cudaMemsetAsync((char *)h->ws_sync.p + 64, 0, 4, st);      // reported once, then re-armed
        return fail(h, WRNN_ERR_CUDA, "per-sample loop kernel aborted: exchange watchdog fired (code 0x%x)", flag);
    }
    return 0;
}

}  // namespace wrnn

using namespace wrnn;

extern "C" {

int wrnn_stage(wrnn_handle *h, const float *const *mels, const int32_t *T, int32_t n_utt, int32_t batched,
               int32_t target, int32_t overlap, int32_t mu_law, const wrnn_noise *noise, void *cuda_stream) {
    if (!h) return WRNN_ERR_ARG;
    return do_stage(h, mels, T, n_utt, batched, target, overlap, mu_law, noise, (cudaStream_t)cuda_stream, StageOpts());
}

int wrnn_run_staged(wrnn_handle *h, uint64_t seed, void *cuda_stream) {
    if (!h) return WRNN_ERR_ARG;
    return do_run(h, seed, (cudaStream_t)cuda_stream, StageOpts());
}

int wrnn_fetch(wrnn_handle *h, double *const *out, int64_t *out_len, void *cuda_stream) {
    if (!h || !out) return fail(h, WRNN_ERR_ARG, "null argument");
    if (!h->staged) return fail(h, WRNN_ERR_STATE, "nothing staged");
    if (h->Btot != h->B)
        return fail(h, WRNN_ERR_STATE, "a fold range is staged: the waveform exists only after the shards are gathered "
                                       "(wrnn_fetch_samples, then the host cross-fade: wavernn_b200/shard.py)");
    WRNN_DEVICE(h);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const size_t bytes = (size_t)h->total_out * sizeof(double);
    if (h->h_pinned_out_bytes < bytes) {
        if (h->h_pinned_out) cudaFreeHost(h->h_pinned_out);
        h->h_pinned_out = nullptr;
        WRNN_CUDA(h, cudaMallocHost((void **)&h->h_pinned_out, bytes + bytes / 4 + 64));
        h->h_pinned_out_bytes = bytes + bytes / 4 + 64;
    }
    WRNN_CUDA(h, cudaMemcpyAsync(h->h_pinned_out, h->ws_out.p, bytes, cudaMemcpyDeviceToHost, st));
    WRNN_TRY(check_abort(h, st));      // synchronises the stream
    WRNN_CUDA(h, cudaGetLastError());
    // pinned staging -> the caller's arrays.  A batch call returns tens of MB into freshly allocated (untouched) host
    // memory: one thread spends ~20 ms of a 420 ms config-5 call on page faults and copying, four spend a quarter of that
    const size_t n_utt = h->utts.size();
    auto copy_range = [&](size_t u0, size_t u1) {
        for (size_t u = u0; u < u1; u++) {
            const auto &ut = h->utts[u];
            memcpy(out[u], h->h_pinned_out + ut.out_off, (size_t)ut.wave_len * sizeof(double));
        }
    };
    const size_t workers = (bytes >= (size_t)8 << 20 && n_utt >= 8) ? 4 : 1;
    if (workers > 1) {
        std::vector<std::thread> th;
        for (size_t k = 1; k < workers; k++) th.emplace_back(copy_range, n_utt * k / workers, n_utt * (k + 1) / workers);
        copy_range(0, n_utt / workers);
        for (auto &x : th) x.join();
    } else {
        copy_range(0, n_utt);
    }
    if (out_len)
        for (size_t u = 0; u < n_utt; u++) out_len[u] = h->utts[u].wave_len;
    return 0;
}

int wrnn_generate_stream(wrnn_handle *h, const float *mel, int32_t T, int32_t target, int32_t overlap, int32_t mu_law,
                         const wrnn_noise *noise, uint64_t seed, int32_t chunk_folds, wrnn_chunk_fn on_chunk, void *user,
                         double *out, int64_t *out_len, void *cuda_stream) {
    if (!h || !mel || chunk_folds < 0) return fail(h, WRNN_ERR_ARG, "bad argument");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const float *mels[1] = {mel};
    WRNN_TRY(do_stage(h, mels, &T, 1, 1, target, overlap, mu_law, noise, st, StageOpts()));
    if (h->Btot != h->B) return fail(h, WRNN_ERR_STATE, "a fold range is staged: streaming delivers a whole utterance");
    WRNN_DEVICE(h);
    const size_t bytes = (size_t)h->total_out * sizeof(double);
    if (h->h_pinned_out_bytes < bytes) {
        if (h->h_pinned_out) cudaFreeHost(h->h_pinned_out);
        h->h_pinned_out = nullptr;
        WRNN_CUDA(h, cudaMallocHost((void **)&h->h_pinned_out, bytes + bytes / 4 + 64));
        h->h_pinned_out_bytes = bytes + bytes / 4 + 64;
    }
    std::vector<cudaEvent_t> events;
    std::vector<int64_t> ends;
    StageOpts o;
    o.chunk_folds = chunk_folds > 0 ? chunk_folds : res_max_folds();
    o.chunk_events = &events;
    o.chunk_end = &ends;
    int rc = do_run(h, seed, st, o);          // enqueues every group, its tail and its copy; returns at once
    int64_t from = 0;
    for (size_t c = 0; c < events.size(); c++) {
        if (rc == 0) {
            const cudaError_t e = cudaEventSynchronize(events[c]);
            if (e != cudaSuccess) rc = fail(h, WRNN_ERR_CUDA, "wrnn_generate_stream: %s", cudaGetErrorString(e));
        }
        if (rc == 0 && ends[c] > from) {
            if (out) memcpy(out + from, h->h_pinned_out + from, (size_t)(ends[c] - from) * sizeof(double));
            if (on_chunk) on_chunk(user, from, h->h_pinned_out + from, ends[c] - from, c + 1 == events.size());
            from = ends[c];
        }
        cudaEventDestroy(events[c]);
    }
    if (rc) return rc;
    WRNN_TRY(check_abort(h, st));
    if (out_len) *out_len = h->utts[0].wave_len;
    return 0;
}

int wrnn_generate(wrnn_handle *h, const float *const *mels, const int32_t *T, int32_t n_utt, int32_t batched,
                  int32_t target, int32_t overlap, int32_t mu_law, const wrnn_noise *noise, uint64_t seed,
                  double *const *out, int64_t *out_len, void *cuda_stream) {
    if (!h) return WRNN_ERR_ARG;
    WRNN_TRY(wrnn_stage(h, mels, T, n_utt, batched, target, overlap, mu_law, noise, cuda_stream));
    WRNN_TRY(wrnn_run_staged(h, seed, cuda_stream));
    return wrnn_fetch(h, out, out_len, cuda_stream);
}

int wrnn_last_timing(wrnn_handle *h, float *ms_cond, float *ms_loop, float *ms_tail, int64_t *steps,
                     int64_t *folds, int64_t *kernel_launches) {
    if (!h) return WRNN_ERR_ARG;
    WRNN_DEVICE(h);
    float a = 0, b = 0, c = 0;
    WRNN_CUDA(h, cudaEventSynchronize(h->ev[3]));
    WRNN_CUDA(h, cudaEventElapsedTime(&a, h->ev[0], h->ev[1]));
    WRNN_CUDA(h, cudaEventElapsedTime(&b, h->ev[1], h->ev[2]));
    WRNN_CUDA(h, cudaEventElapsedTime(&c, h->ev[2], h->ev[3]));
    if (ms_cond) *ms_cond = a;
    if (ms_loop) *ms_loop = b;
    if (ms_tail) *ms_tail = c;
    if (steps) *steps = h->S;
    if (folds) *folds = h->B;
    if (kernel_launches) *kernel_launches = h->launches;
    return 0;
}

int wrnn_set_fold_range(wrnn_handle *h, int64_t lo, int64_t hi) {
    if (!h) return WRNN_ERR_ARG;
    if (lo < 0 || hi < lo) return fail(h, WRNN_ERR_ARG, "bad fold range");
    h->range_lo = lo;
    h->range_hi = hi;
    h->staged = false;
    return 0;
}

int wrnn_fetch_samples(wrnn_handle *h, float *samples, void *cuda_stream) {
    if (!h || !samples) return fail(h, WRNN_ERR_ARG, "null argument");
    if (!h->staged) return fail(h, WRNN_ERR_STATE, "nothing staged");
    WRNN_DEVICE(h);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    WRNN_CUDA(h, cudaMemcpyAsync(samples, h->ws_samples.p, (size_t)h->B * h->S * sizeof(float), cudaMemcpyDeviceToHost, st));
    WRNN_TRY(check_abort(h, st));      // synchronises the stream; a shard must not hand on samples of an aborted loop
    return 0;
}

int wrnn_forward_logits(wrnn_handle *h, const float *x, const float *mels, int32_t n_seq, int64_t S, float *logits) {
    if (!h || !x || !mels || !logits || n_seq < 1) return fail(h, WRNN_ERR_ARG, "null argument");
    const Dims &d = h->d;
    WRNN_DEVICE(h);
    if (S < 1 || S % d.hop) return fail(h, WRNN_ERR_ARG, "S must be a positive multiple of hop_length");
    const int Tp = (int)(S / d.hop) + 2 * d.pad;
    std::vector<const float *> mp(n_seq);
    std::vector<int32_t> Ts(n_seq, Tp);
    for (int i = 0; i < n_seq; i++) mp[i] = mels + (size_t)i * d.F * Tp;
    StageOpts o;
    o.prepadded = true;
    o.teacher = x;
    o.want_logits = true;
    WRNN_TRY(do_stage(h, mp.data(), Ts.data(), n_seq, 0, 0, 0, 0, nullptr, 0, o));
    WRNN_TRY(do_run(h, 0, 0, o));
    WRNN_CUDA(h, cudaMemcpyAsync(logits, h->ws_logits.p, (size_t)n_seq * S * d.NC * sizeof(float), cudaMemcpyDeviceToHost, 0));
    WRNN_TRY(check_abort(h, 0));
    h->staged = false;
    return 0;
}

// ------------------------------------------------------------------ component entry points
int wrnn_upsample(wrnn_handle *h, const float *mel, int32_t T, float *mels_out, float *aux_out) {
    if (!h || !mel || !mels_out || !aux_out || T < 1) return fail(h, WRNN_ERR_ARG, "bad argument");
    const Dims &d = h->d;
    WRNN_DEVICE(h);
    WRNN_TRY(ensure_packed_public(h));
    const int Tp = T + 2 * d.pad;
    const int64_t L = (int64_t)T * d.hop;
    const int n_tiles = (T + resnet_tile_frames() - 1) / resnet_tile_frames();
    const size_t raw = align_up((size_t)d.F * T, 64), padf = (size_t)d.F * Tp;
    WRNN_TRY(ws_reserve(h, h->ws_mel, (raw + padf) * sizeof(float) + 512));
    WRNN_TRY(ws_reserve(h, h->ws_up, (size_t)L * d.F * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_res, (size_t)T * d.R * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_cond, (size_t)L * d.R * sizeof(float)));
    const MetaLayout ml = meta_layout(1, 1, n_tiles);
    std::vector<char> hm(ml.bytes, 0);
    *(UttMeta *)(hm.data() + ml.off_utt) = UttMeta{0, 0, 0, T, Tp, T, d.pad};
    for (int i = 0; i < n_tiles; i++) ((ResTile *)(hm.data() + ml.off_tile))[i] = ResTile{0, i * resnet_tile_frames()};
    WRNN_TRY(ws_reserve(h, h->ws_meta, ml.bytes));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_meta.p, hm.data(), ml.bytes, cudaMemcpyHostToDevice, 0));
    float *mraw = (float *)h->ws_mel.p, *melT = mraw + raw;
    const UttMeta *d_utt = (const UttMeta *)((char *)h->ws_meta.p + ml.off_utt);
    WRNN_CUDA(h, cudaMemcpyAsync(mraw, mel, (size_t)d.F * T * sizeof(float), cudaMemcpyHostToDevice, 0));
    WRNN_TRY(launch_pad_mel_batch(h, mraw, d_utt, 1, Tp, melT, 0));
    WRNN_TRY(launch_resnet_batch(h, melT, d_utt, (const ResTile *)((char *)h->ws_meta.p + ml.off_tile), n_tiles, (float *)h->ws_res.p, 0));
    WRNN_TRY(launch_upsample_flat(h, melT, Tp, L, (float *)h->ws_up.p, 0));
    WRNN_TRY(launch_repeat_aux(h, (const float *)h->ws_res.p, T, (float *)h->ws_cond.p, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(mels_out, h->ws_up.p, (size_t)L * d.F * sizeof(float), cudaMemcpyDeviceToHost, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(aux_out, h->ws_cond.p, (size_t)L * d.R * sizeof(float), cudaMemcpyDeviceToHost, 0));
    WRNN_CUDA(h, cudaStreamSynchronize(0));
    h->staged = false;
    return 0;
}

int wrnn_fold(wrnn_handle *h, const float *x, int64_t L, int32_t F, int64_t target, int64_t overlap, float *out) {
    if (!h || !x || !out || L < 1 || F < 1 || target < 1 || overlap < 0 || L <= overlap)
        return fail(h, WRNN_ERR_ARG, "bad argument");
    WRNN_DEVICE(h);
    const int64_t folds = wrnn_num_folds(L, target, overlap), len = target + 2 * overlap;
    WRNN_TRY(ws_reserve(h, h->ws_tmp0, (size_t)L * F * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_tmp1, (size_t)folds * len * F * sizeof(float)));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_tmp0.p, x, (size_t)L * F * sizeof(float), cudaMemcpyHostToDevice, 0));
    WRNN_TRY(launch_fold(h, (const float *)h->ws_tmp0.p, L, F, target, overlap, folds, (float *)h->ws_tmp1.p, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(out, h->ws_tmp1.p, (size_t)folds * len * F * sizeof(float), cudaMemcpyDeviceToHost, 0));
    WRNN_CUDA(h, cudaStreamSynchronize(0));
    return 0;
}

int wrnn_unfold(wrnn_handle *h, const double *y, int64_t folds, int64_t target, int64_t overlap, double *out) {
    if (!h || !y || !out || folds < 1 || target < 0 || overlap < 0) return fail(h, WRNN_ERR_ARG, "bad argument");
    WRNN_DEVICE(h);
    const int64_t len = target + 2 * overlap, total = folds * (target + overlap) + overlap;
    const size_t ov = (size_t)std::max<int64_t>(overlap, 1);
    std::vector<double> tab(2 * ov, 0.0);
    wrnn_xfade_tables(overlap, tab.data(), tab.data() + ov);
    WRNN_TRY(ws_reserve(h, h->ws_tmp0, (size_t)folds * len * sizeof(double)));
    WRNN_TRY(ws_reserve(h, h->ws_tmp1, (size_t)total * sizeof(double)));
    WRNN_TRY(ws_reserve(h, h->ws_tables, 2 * ov * sizeof(double)));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_tmp0.p, y, (size_t)folds * len * sizeof(double), cudaMemcpyHostToDevice, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_tables.p, tab.data(), 2 * ov * sizeof(double), cudaMemcpyHostToDevice, 0));
    WRNN_TRY(launch_unfold(h, (const double *)h->ws_tmp0.p, folds, target, overlap, (const double *)h->ws_tables.p,
                           (const double *)h->ws_tables.p + ov, (double *)h->ws_tmp1.p, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(out, h->ws_tmp1.p, (size_t)total * sizeof(double), cudaMemcpyDeviceToHost, 0));
    WRNN_CUDA(h, cudaStreamSynchronize(0));
    h->staged = false;
    return 0;
}

int wrnn_decode_mu_law(wrnn_handle *h, const float *y, int64_t n, int32_t mu, double *out) {
    if (!h || !y || !out || n < 1 || mu < 2) return fail(h, WRNN_ERR_ARG, "bad argument");
    WRNN_DEVICE(h);
    std::vector<double> lut(mu);
    if ((int)h->lut_override.size() == mu) lut = h->lut_override;
    else wrnn_mulaw_lut(mu, lut.data());
    WRNN_TRY(ws_reserve(h, h->ws_tmp0, (size_t)n * sizeof(float)));
    WRNN_TRY(ws_reserve(h, h->ws_tmp1, (size_t)n * sizeof(double)));
    WRNN_TRY(ws_reserve(h, h->ws_tables, (size_t)mu * sizeof(double)));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_tmp0.p, y, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(h->ws_tables.p, lut.data(), (size_t)mu * sizeof(double), cudaMemcpyHostToDevice, 0));
    WRNN_TRY(launch_mulaw(h, (const float *)h->ws_tmp0.p, n, mu, (const double *)h->ws_tables.p, (double *)h->ws_tmp1.p, 0));
    WRNN_CUDA(h, cudaMemcpyAsync(out, h->ws_tmp1.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, 0));
    WRNN_CUDA(h, cudaStreamSynchronize(0));
    h->staged = false;
    return 0;
}

}  // extern "C"
