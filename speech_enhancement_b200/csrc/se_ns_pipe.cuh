// Pipelined NoiseSup kernel for SMALL batches: one CTA per utterance, the work of one block spread over 5-8 warps.
// Included by se_ns.cu (fast arithmetic only).
//
// The batch kernel (se_ns_kernel.cuh) gives an utterance one warp, so a lone 10 s utterance is 1 000 strictly serial
// block iterations of ~5 us (8 kHz) / ~6.5 us (16 kHz): twice a CPU core, and the ramp of every host-buffer chunk.
// The recursion is serial in TIME, but inside one block iteration the parts form a chain that can be software-pipelined
// over blocks:
//
//   warp 0  input -> ring1, zero-frame gate, frame energy, window + FFT + PSD of stage 0   (input only: runs ahead)
//   warp 1  stage-0 recursion: VAD, FilterCalc, mel / Gamma bank, denoised-energy history, IDCT   -> 9 filter taps
//   warp 2  17-tap FIR of stage 0 over its own ring1 -> ring2, window + FFT + PSD of stage 1
//   warp 3  stage-1 recursion: noise update, FilterCalc, bank, gain factorisation, IDCT           -> 9 filter taps
//   warp 4  FIR of stage 1 over its own ring2, DC filter, every output (16 kHz profile: also the statistics of warp 7)
//   warp 5  (8 kHz profile) CompCeps, first half: ring3, pre-emphasis, Hamming, FFT, power spectrum
//   warp 6  (8 kHz profile) CompCeps, second half: mel triangles, log, DCT, the cepstral row
//   warp 7  (8 kHz profile) frame-dropping statistics of stage 0 (three reductions + the double-precision SpeechQ recursions)
//
// Every warp keeps PRIVATE copies of what it reads repeatedly (warps 0 and 2 each hold ring1, warps 2 and 4 each ring2),
// so the only shared data are six to eight small FIFOs of kPipeDepth entries in shared memory, handed over with mbarriers (full /
// empty per slot).  Every role consumes and produces exactly one entry per input block, in order.  The arithmetic is the
// batch kernel's own device functions (frame_psd, ns_stage_rest, fd_stats, apply_fir, dc_offset, comp_ceps), so the
// outputs are bit-identical to the batch kernel's; the launcher picks this kernel when the batch is too small to fill
// the GPU with one warp per utterance (se_ctx_set_ns_pipeline overrides).
#pragma once

namespace SE_NSPACE {

#ifndef SE_PIPE_DEPTH
#define SE_PIPE_DEPTH 4
#endif
constexpr int kPipeDepth = SE_PIPE_DEPTH;      // entries per FIFO (a power of two)
template <int PROF> struct PipeWarps { static constexpr int value = PROF == 8 ? 8 : 5; };
constexpr int kPipeFifos = 8;
// FIFO index: producer -> consumer(s)
enum { kFA = 0,   // warp 0 -> 1, 2 : input block, stage-0 PSD, frame energy
       kFB = 1,   // warp 1 -> 2 and 7 (8 kHz) / 4 (16 kHz): stage-0 filter taps, lane partials for the statistics, VAD state
       kFC = 2,   // warp 2 -> 3    : stage-1 PSD, denoised-energy history
       kFE = 3,   // warp 2 -> 4    : stage-0 output block, VAD state
       kFD = 4,   // warp 3 -> 4    : stage-1 filter taps
       kFF = 5,   // warp 4 -> 5    : denoised block (8 kHz profile)
       kFG = 6,   // warp 7 -> 4    : frame-dropping flags (8 kHz profile)
       kFH = 7 }; // warp 5 -> 6    : power spectrum and frame energy (8 kHz profile)

template <int PROF> struct __align__(16) PipeMem {
    using P = Prof<PROF>;
    struct __align__(16) EntA { float x[P::S]; float psd[P::NB * 32]; float frame_sq; int process, run, pad0; };
    struct __align__(16) EntB {
        float h[12], wsum[32], wsq[32], mw[32];
        int process, ran0, nb_speech, nb_frame0, flag_vad, pad0, pad1, pad2;
        float den_en0, den_en1, den_en2, pad3;
    };
    struct __align__(16) EntC { float psd[P::NB * 32]; int run, pad0; float den_en0, den_en1, den_en2, pad1, pad2, pad3; };
    struct __align__(16) EntE { float x[P::S]; int process, ran0, nb_speech, nb_frame0, flag_vad, pad0, pad1, pad2; };
    struct __align__(16) EntD { float h[12]; int run, pad0, pad1, pad2; };
    struct __align__(16) EntF { float x[P::S]; int produced, nb_out2, pad0, pad1; };
    struct __align__(16) EntG { int var, spec, mel, pad; };
    struct __align__(16) EntH { float pw[132]; float e; int compute, count, m; };
    EntA a[kPipeDepth];
    EntB b[kPipeDepth];
    EntC c[kPipeDepth];
    EntE e[kPipeDepth];
    EntD d[kPipeDepth];
    EntF f[kPipeDepth];
    EntG g[PROF == 8 ? kPipeDepth : 1];
    EntH h[PROF == 8 ? kPipeDepth : 1];
    unsigned long long full[kPipeFifos][kPipeDepth], empty[kPipeFifos][kPipeDepth];
};

// -DSE_PIPE_STATS (a measurement build, tools/build_variants.sh): cycles every role spends waiting for its input
// (full) and for room in its output (empty), summed over the launch; read back with se_debug_pipe_stats().
#ifdef SE_PIPE_STATS
__device__ unsigned long long g_pipe_stats[8][4];
#define SE_PIPE_T0 const long long t0__ = clock64()
#define SE_PIPE_ACC(i) pipe_wait_cycles[i] += clock64() - t0__
#else
#define SE_PIPE_T0
#define SE_PIPE_ACC(i)
#endif

// consumer / producer sides of one FIFO slot; entry k lives in slot k % depth, round k / depth
__device__ __forceinline__ void pipe_wait_full(unsigned long long *full, unsigned k, long long *pipe_wait_cycles)
{
    SE_PIPE_T0;
    se_bulk::mbar_wait_warp(&full[k & (kPipeDepth - 1)], (k / kPipeDepth) & 1u);
    SE_PIPE_ACC(0);
}
__device__ __forceinline__ void pipe_wait_empty(unsigned long long *empty, unsigned k, long long *pipe_wait_cycles)
{
    SE_PIPE_T0;
    se_bulk::mbar_wait_warp(&empty[k & (kPipeDepth - 1)], ((k / kPipeDepth) & 1u) ^ 1u);
    SE_PIPE_ACC(1);
}
// every lane's shared-memory accesses to the slot are done: one arrival by lane 0
__device__ __forceinline__ void pipe_signal(unsigned long long *bars, unsigned k, int lane)
{
    __syncwarp();
    if (lane == 0) se_bulk::mbar_arrive(&bars[k & (kPipeDepth - 1)]);
}

template <int PROF, typename IN_T>
__global__ void __launch_bounds__(PipeWarps<PROF>::value * 32, 2)
ns_pipe_kernel(const NsLaunch a, const NsDevTables tb, const CcDevTables cc)
{
    using P = Prof<PROF>;
    constexpr int NW = PipeWarps<PROF>::value;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BlockMem<PROF> &bm = *reinterpret_cast<BlockMem<PROF> *>(smem_raw);
    WarpMem<PROF> *wms = reinterpret_cast<WarpMem<PROF> *>(smem_raw + ((sizeof(BlockMem<PROF>) + 15) & ~size_t(15)));
    PipeMem<PROF> &pm = *reinterpret_cast<PipeMem<PROF> *>(wms + NW);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpMem<PROF> &wm = wms[warp];

    load_block_tables<PROF>(bm, tb, cc);
    if (threadIdx.x == 0) {
        const int consumers[kPipeFifos] = {2, 2, 1, 1, 1, 1, 1, 1};
        for (int f = 0; f < kPipeFifos; f++)
            for (int k = 0; k < kPipeDepth; k++) {
                se_bulk::mbar_init(&pm.full[f][k], 1);
                se_bulk::mbar_init(&pm.empty[f][k], consumers[f]);
            }
        se_bulk::mbar_init_fence();
    }
    __syncthreads();
    bank_tmem_setup<PROF>(bm, tb.bankT);
    FftTw tw;
    fft_tw_init(tw, lane);
    IdctRegs ic;
    idct_regs_init(ic, tb.idctT, tb.fir_window, lane);
    BankRegs br;
    if constexpr (PROF == 8) bank_regs_init(br, tb, cc, lane);
    else { br.mel_st = br.tri_st = 0; }
#ifdef SE_NS_CONST_SMEM
    if (warp == 0) {
        lane_consts_store(bm.lane_tw, bm.lane_ic, tw, ic, lane);
        if constexpr (PROF == 8) bank_regs_store(bm.lane_br, br, lane);
    }
    __syncthreads();
#endif
    if constexpr (WarpMem<PROF>::kSpecScratch)
        for (int i = lane; i < 132; i += 32) wm.W[i] = 0.0f;
    const float eps = tb.eps;
    const IN_T *all = reinterpret_cast<const IN_T *>(a.samples);
    unsigned kbase = 0;                      // FIFO entry index of block 0 of the current utterance (same on every warp)
    long long wait_cycles[2] = {0, 0};
    (void)wait_cycles;
#ifdef SE_PIPE_STATS
    const long long t_start = clock64();
#endif

    for (int q = blockIdx.x; q < a.n_utts; q += gridDim.x) {
        int u = a.utt_begin + q;
        if (a.order) u = a.order[u];
        const int64_t off0 = a.offsets[u] - a.sample_base;
        const int64_t n = a.offsets[u + 1] - a.offsets[u];
        const int nblk = (int)(n / P::S);
        const IN_T *in = all + off0 + a.in_shift;
        const int64_t blk0 = a.block_offsets ? a.block_offsets[u] - a.block_base : 0;
        const int64_t tr0 = a.trace_offsets ? a.trace_offsets[u] - a.trace_base : 0;
        const int64_t ceps0 = a.ceps_offsets ? a.ceps_offsets[u] - a.ceps_base : 0;
        const int64_t ceps_cap = a.ceps_offsets ? a.ceps_offsets[u + 1] - a.ceps_offsets[u] : 0;

        // every role starts the utterance from the reference's initial state (DoNoiseSupInit); a role only ever touches
        // its own part of it
        Scal s;
        scal_init(s);
        float noise[P::NB], den[P::NB];      // this role's stage: noiseSE / denSigSE of its bins
#pragma unroll
        for (int r = 0; r < P::NB; r++) {
            noise[r] = eps;
            den[r] = 0.0f;
            wm.prev[r * 32 + lane] = 0.0f;
            wm.prev[(P::NB + r) * 32 + lane] = 0.0f;
        }
        for (int i = lane; i < P::RB + 16; i += 32) { wm.ring1[i] = 0.0f; wm.ring2[i] = 0.0f; }
        if (P::CEPS)
            for (int i = lane; i < P::RB; i += 32) wm.ring3[i] = 0.0f;
        SE_COLD(cold_fd_store(wm.cold, s));
        SE_COLD(cold_gf_store(wm.cold, s));
        float *ring1 = wm.ring1 + 8, *ring2 = wm.ring2 + 8;
        __syncwarp();

        constexpr int SL = kPipeDepth - 1;
        if (warp == 0) {
            // ---- input, zero-frame gate, frame energy, stage-0 PSD -----------------------------------------------------
            IN_T nx[P::NR];
#pragma unroll
            for (int r = 0; r < P::NR; r++) {
                const int i = lane + 32 * r;
                nx[r] = (nblk > 0 && i < P::S) ? in[i] : IN_T(0);
            }
            for (int b = 0; b < nblk; b++) {
                float cur[P::NR];
                float ss = 0.0f;
                bool nzl = false;
#pragma unroll
                for (int r = 0; r < P::NR; r++) { cur[r] = to_f32(nx[r]); ss += cur[r] * cur[r]; nzl |= (cur[r] != 0.0f); }
                if (b + 1 < nblk) {
#pragma unroll
                    for (int r = 0; r < P::NR; r++) {
                        const int i = lane + 32 * r;
                        if (i < P::S) nx[r] = in[(int64_t)(b + 1) * P::S + i];
                    }
                }
                bool process = true;
                if (a.zero_check) {
                    bool nonzero;
                    if (sizeof(IN_T) == 2) nonzero = __any_sync(0xffffffffu, nzl);
                    else nonzero = !(warp_sum(ss) < 1.0f);
                    if (a.zero_check == 1) { process = nonzero || (s.onset != 0); if (process) s.onset = 1; }
                    else process = nonzero;
                }
                const unsigned k = kbase + b;
                pipe_wait_empty(pm.empty[kFA], k, wait_cycles);
                typename PipeMem<PROF>::EntA &e = pm.a[k & SL];
                bool run = false;
                if (process) {
                    const int slot = s.nb_in1 & 3;
#pragma unroll
                    for (int r = 0; r < P::NR; r++) {
                        const int i = lane + 32 * r;
                        if (i < P::S) { ring_store<PROF>(ring1, slot, i, cur[r]); e.x[i] = cur[r]; }
                    }
                    s.nb_in1++;
                    __syncwarp();
                    run = s.nb_in1 - s.nb_in2 > 2;
                    if (run) {
                        float psd[P::NB];
                        frame_psd<PROF>(wm, bm, tw, lane, ring1, (s.nb_in1 & 3) * P::S + P::ANA, psd);
#pragma unroll
                        for (int r = 0; r < P::NB; r++) e.psd[r * 32 + lane] = psd[r];
                        const float fsq = frame_energy<PROF>(ring1, ((s.nb_in1 + 1) & 3) * P::S, lane);
                        if (lane == 0) e.frame_sq = fsq;
                        s.nb_in2++;
                    }
                }
                if (lane == 0) { e.process = process ? 1 : 0; e.run = run ? 1 : 0; }
                pipe_signal(pm.full[kFA], k, lane);
            }
        } else if (warp == 1) {
            // ---- stage-0 recursion -> filter taps ---------------------------------------------------------------------
            for (int b = 0; b < nblk; b++) {
                const unsigned k = kbase + b;
                pipe_wait_full(pm.full[kFA], k, wait_cycles);
                const typename PipeMem<PROF>::EntA &e = pm.a[k & SL];
                const bool process = e.process != 0, run = e.run != 0;
                float psd[P::NB];
#pragma unroll
                for (int r = 0; r < P::NB; r++) psd[r] = run ? e.psd[r * 32 + lane] : 0.0f;
                StageSplit sp;
                sp.frame_sq = run ? e.frame_sq : 0.0f;
                pipe_signal(pm.empty[kFA], k, lane);
                if (process) s.nb_in1++;
                float o[P::CH];
                StageFlags fl = {0, 0, 0};
                if (run) ns_stage_rest<PROF, true>(wm, bm, ic, br, lane, s, noise, den, eps, 0, psd, o, fl, nullptr, nullptr, &sp);
                pipe_wait_empty(pm.empty[kFB], k, wait_cycles);
                typename PipeMem<PROF>::EntB &eb = pm.b[k & SL];
                if (run) {
                    eb.wsum[lane] = sp.wsum_part;
                    eb.wsq[lane] = sp.wsq_part;
                    eb.mw[lane] = sp.mw;
                    if (lane == 0) {
#pragma unroll
                        for (int t = 0; t < 9; t++) eb.h[t] = sp.h[t];
                    }
                }
                if (lane == 0) {
                    eb.process = process ? 1 : 0; eb.ran0 = run ? 1 : 0;
                    eb.nb_speech = s.nb_speech; eb.nb_frame0 = s.nb_frame0; eb.flag_vad = s.flag_vad;
                    eb.den_en0 = s.den_en0; eb.den_en1 = s.den_en1; eb.den_en2 = s.den_en2;
                }
                pipe_signal(pm.full[kFB], k, lane);
            }
        } else if (warp == 2) {
            // ---- stage-0 FIR over a private ring1 -> ring2, stage-1 PSD ---------------------------------------------------
            for (int b = 0; b < nblk; b++) {
                const unsigned k = kbase + b;
                pipe_wait_full(pm.full[kFA], k, wait_cycles);
                const typename PipeMem<PROF>::EntA &e = pm.a[k & SL];
                if (e.process != 0) {
                    const int slot = s.nb_in1 & 3;
#pragma unroll
                    for (int r = 0; r < P::NR; r++) {
                        const int i = lane + 32 * r;
                        if (i < P::S) ring_store<PROF>(ring1, slot, i, e.x[i]);
                    }
                    s.nb_in1++;
                }
                pipe_signal(pm.empty[kFA], k, lane);
                pipe_wait_full(pm.full[kFB], k, wait_cycles);
                const typename PipeMem<PROF>::EntB &eb = pm.b[k & SL];
                const bool ran0 = eb.ran0 != 0;
                float h[9];
#pragma unroll
                for (int t = 0; t < 9; t++) h[t] = eb.h[t];
                const float de0 = eb.den_en0, de1 = eb.den_en1, de2 = eb.den_en2;
                const int m_process = eb.process, m_nbs = eb.nb_speech, m_nbf = eb.nb_frame0, m_vad = eb.flag_vad;
                pipe_signal(pm.empty[kFB], k, lane);
                pipe_wait_empty(pm.empty[kFE], k, wait_cycles);
                typename PipeMem<PROF>::EntE &ee = pm.e[k & SL];
                if (ran0) {
                    float o[P::CH];
                    apply_fir<PROF>(ring1, ((s.nb_in1 + 1) & 3) * P::S, h, lane, o);
                    const int slot = s.nb_in2 & 3, i0 = lane * P::CH;
#pragma unroll
                    for (int t = 0; t < P::CH; t++)
                        if (i0 + t < P::S) { ring_store<PROF>(ring2, slot, i0 + t, o[t]); ee.x[i0 + t] = o[t]; }
                    s.nb_in2++;
                }
                if (lane == 0) {
                    ee.process = m_process; ee.ran0 = ran0 ? 1 : 0;
                    ee.nb_speech = m_nbs; ee.nb_frame0 = m_nbf; ee.flag_vad = m_vad;
                }
                pipe_signal(pm.full[kFE], k, lane);            // (its __syncwarp also orders the ring2 stores before the PSD)
                const bool run = ran0 && (s.nb_in2 - s.nb_out2 > 2);
                pipe_wait_empty(pm.empty[kFC], k, wait_cycles);
                typename PipeMem<PROF>::EntC &ec = pm.c[k & SL];
                if (run) {
                    float psd[P::NB];
                    frame_psd<PROF>(wm, bm, tw, lane, ring2, (s.nb_in2 & 3) * P::S + P::ANA, psd);
#pragma unroll
                    for (int r = 0; r < P::NB; r++) ec.psd[r * 32 + lane] = psd[r];
                    s.nb_out2++;
                }
                if (lane == 0) { ec.run = run ? 1 : 0; ec.den_en0 = de0; ec.den_en1 = de1; ec.den_en2 = de2; }
                pipe_signal(pm.full[kFC], k, lane);
            }
        } else if (warp == 3) {
            // ---- stage-1 recursion -> filter taps (and the gains row of the 16 kHz profile) --------------------------------
            for (int b = 0; b < nblk; b++) {
                const unsigned k = kbase + b;
                pipe_wait_full(pm.full[kFC], k, wait_cycles);
                const typename PipeMem<PROF>::EntC &ec = pm.c[k & SL];
                const bool run = ec.run != 0;
                float psd[P::NB];
#pragma unroll
                for (int r = 0; r < P::NB; r++) psd[r] = run ? ec.psd[r * 32 + lane] : 0.0f;
                s.den_en0 = ec.den_en0; s.den_en1 = ec.den_en1; s.den_en2 = ec.den_en2;
                pipe_signal(pm.empty[kFC], k, lane);
                if constexpr (PROF == 8) { wm.cold[8] = s.den_en0; wm.cold[9] = s.den_en1; wm.cold[10] = s.den_en2; }
                StageSplit sp;
                if (run) {
                    float o[P::CH];
                    StageFlags fl1 = {0, 0, 0};
                    float *grow = (PROF == 16 && a.gains) ? a.gains + (blk0 + s.n_rows) * kMel : nullptr;
                    ns_stage_rest<PROF, true>(wm, bm, ic, br, lane, s, noise, den, eps, 1, psd, o, fl1, grow, nullptr, &sp);
                    s.n_rows++;
                }
                pipe_wait_empty(pm.empty[kFD], k, wait_cycles);
                typename PipeMem<PROF>::EntD &ed = pm.d[k & SL];
                if (lane == 0) {
                    if (run) {
#pragma unroll
                        for (int t = 0; t < 9; t++) ed.h[t] = sp.h[t];
                    }
                    ed.run = run ? 1 : 0;
                }
                pipe_signal(pm.full[kFD], k, lane);
            }
            if (lane == 0 && a.n_rows) a.n_rows[u] = s.n_rows;
        } else if (warp == 4) {
            // ---- frame-dropping statistics, stage-1 FIR over a private ring2, DC filter, outputs -----------------------------
            for (int b = 0; b < nblk; b++) {
                const unsigned k = kbase + b;
                StageFlags fl = {0, 0, 0};
                if constexpr (PROF != 8) {                     // 16 kHz profile: the statistics run here (8 kHz: warp 7)
                    pipe_wait_full(pm.full[kFB], k, wait_cycles);
                    const typename PipeMem<PROF>::EntB &eb = pm.b[k & SL];
                    const bool r0 = eb.ran0 != 0;
                    s.nb_frame0 = eb.nb_frame0;
                    const float wsum = eb.wsum[lane], wsq = eb.wsq[lane], mw = eb.mw[lane];
                    pipe_signal(pm.empty[kFB], k, lane);
                    if (r0) fd_stats<PROF>(wm, s, wsum, wsq, mw, fl);
                }
                pipe_wait_full(pm.full[kFE], k, wait_cycles);
                const typename PipeMem<PROF>::EntE &ee = pm.e[k & SL];
                const bool process = ee.process != 0, ran0 = ee.ran0 != 0;
                s.nb_speech = ee.nb_speech; s.nb_frame0 = ee.nb_frame0; s.flag_vad = ee.flag_vad;
                if (ran0) {
                    const int slot = s.nb_in2 & 3, i0 = lane * P::CH;
#pragma unroll
                    for (int t = 0; t < P::CH; t++)
                        if (i0 + t < P::S) ring_store<PROF>(ring2, slot, i0 + t, ee.x[i0 + t]);
                    s.nb_in2++;
                }
                pipe_signal(pm.empty[kFE], k, lane);
                pipe_wait_full(pm.full[kFD], k, wait_cycles);
                const typename PipeMem<PROF>::EntD &ed = pm.d[k & SL];
                const bool run = ed.run != 0;
                float h[9];
#pragma unroll
                for (int t = 0; t < 9; t++) h[t] = ed.h[t];
                pipe_signal(pm.empty[kFD], k, lane);
                float o[P::CH];
                bool produced = false;
                if (run) {
                    apply_fir<PROF>(ring2, ((s.nb_in2 + 1) & 3) * P::S, h, lane, o);
                    s.nb_out2++;
                    produced = true;
                    dc_offset<PROF>(o, s, lane);
                }
                const int64_t obase = off0 + (int64_t)b * P::S;
                const int i0 = lane * P::CH;
                if constexpr (PROF == 8) {
                    // denoised shift -> CompCeps warp, then coalesced int16 / float stores (as the batch kernel)
                    pipe_wait_empty(pm.empty[kFF], k, wait_cycles);
                    typename PipeMem<PROF>::EntF &ef = pm.f[k & SL];
                    float *stage_out = wm.W;
#pragma unroll
                    for (int t = 0; t < P::CH; t++)
                        if (i0 + t < P::S) {
                            const float v = produced ? o[t] : 0.0f;
                            stage_out[i0 + t] = v;
                            ef.x[i0 + t] = v;
                        }
                    if (lane == 0) { ef.produced = produced ? 1 : 0; ef.nb_out2 = s.nb_out2; }
                    pipe_signal(pm.full[kFF], k, lane);        // (its __syncwarp also orders the stage_out reads below)
                    if (a.den_i16) {
                        for (int i = lane; i < P::S / 2; i += 32) {
                            const float2 f = *reinterpret_cast<const float2 *>(stage_out + 2 * i);
                            short2 qv;
                            qv.x = (short)__float2int_rz(f.x);
                            qv.y = (short)__float2int_rz(f.y);
                            if ((reinterpret_cast<uintptr_t>(a.den_i16 + obase) & 3) == 0) *reinterpret_cast<short2 *>(a.den_i16 + obase + 2 * i) = qv;
                            else { a.den_i16[obase + 2 * i] = qv.x; a.den_i16[obase + 2 * i + 1] = qv.y; }
                        }
                    }
                    if (a.den_f32)
                        for (int i = lane; i < P::S; i += 32) a.den_f32[obase + i] = stage_out[i];
                    {                                          // this block's frame-dropping flags, from warp 7
                        pipe_wait_full(pm.full[kFG], k, wait_cycles);
                        const typename PipeMem<PROF>::EntG &eg = pm.g[k & SL];
                        fl.var = eg.var; fl.spec = eg.spec; fl.mel = eg.mel;
                        pipe_signal(pm.empty[kFG], k, lane);
                    }
                    if (a.vad_trace && lane == 0) {
                        unsigned char t = 128;
                        if (process)
                            t = (unsigned char)((s.flag_vad ? 1 : 0) | (fl.var << 1) | (fl.spec << 2) | (fl.mel << 3) |
                                                ((s.nb_speech > 4 ? 1 : 0) << 4) | (ran0 ? 32 : 0) | (produced ? 64 : 0));
                        a.vad_trace[tr0 + b] = t;
                    }
                    __syncwarp();
                } else {
                    if (a.out_f32) {
#pragma unroll
                        for (int t = 0; t < P::CH; t++)
                            if (i0 + t < P::S) a.out_f32[obase + i0 + t] = produced ? o[t] : 0.0f;
                    }
                    if (a.flags && lane == 0) {
                        int32_t *f = a.flags + blk0 + b;
                        const int64_t T = a.total_blocks;
                        f[0 * T] = ran0 ? fl.var : 0;
                        f[1 * T] = ran0 ? fl.spec : 0;
                        f[2 * T] = ran0 ? fl.mel : 0;
                        f[3 * T] = ran0 ? (s.nb_speech > 4 ? 1 : 0) : 0;
                        f[4 * T] = (ran0 || produced) ? s.nb_frame0 : 0;
                    }
                }
            }
            // tail (len % S): the reference leaves it untouched; the batch API defines it as 0
            if (PROF == 8) {
                for (int64_t i = (int64_t)nblk * P::S + lane; i < n; i += 32) {
                    if (a.den_i16) a.den_i16[off0 + i] = 0;
                    if (a.den_f32) a.den_f32[off0 + i] = 0.0f;
                }
            } else if (a.out_f32) {
                for (int64_t i = (int64_t)nblk * P::S + lane; i < n; i += 32) a.out_f32[off0 + i] = 0.0f;
            }
        } else if (warp == 5) {
            // ---- CompCeps, first half (8 kHz profile): ring3, pre-emphasis, Hamming, FFT, power spectrum ------------------------
            if constexpr (PROF == 8) {
                for (int b = 0; b < nblk; b++) {
                    const unsigned k = kbase + b;
                    pipe_wait_full(pm.full[kFF], k, wait_cycles);
                    const typename PipeMem<PROF>::EntF &ef = pm.f[k & SL];
                    const bool produced = ef.produced != 0;
                    const int nb_out2 = ef.nb_out2;
                    if (produced) {
                        const int nb3 = ((nb_out2 - 1) & 3) * P::S;
                        for (int i = lane; i < P::S; i += 32) wm.ring3[nb3 + i] = ef.x[i];
                    }
                    pipe_signal(pm.empty[kFF], k, lane);
                    const bool count = produced && a.ceps && nb_out2 >= 3;       // third denoised shift onwards
                    const int m = nb_out2 - 3;
                    const bool compute = count && m < ceps_cap;
                    float e = 0.0f, d0, d1, d2;
                    if (compute) {
                        const float *d = wm.ring3;
                        const int base = (m & 3) * P::S;
                        comp_ceps_front(wm, bm, tw, lane, [&](int i) { return d[wrap<P::RB>(base + i)]; }, e, d0, d1, d2);
                    }
                    pipe_wait_empty(pm.empty[kFH], k, wait_cycles);
                    typename PipeMem<PROF>::EntH &eh = pm.h[k & SL];
                    if (compute) {
#pragma unroll
                        for (int r = 0; r < 4; r++) eh.pw[lane + 32 * r] = wm.W[lane + 32 * r];
                        if (lane == 0) eh.pw[128] = wm.W[128];
                    }
                    if (lane == 0) { eh.e = e; eh.compute = compute ? 1 : 0; eh.count = count ? 1 : 0; eh.m = m; }
                    pipe_signal(pm.full[kFH], k, lane);
                }
            }
        } else if (warp == 6) {
            // ---- CompCeps, second half (8 kHz profile): mel triangles, log, DCT ------------------------------------------------
            if constexpr (PROF == 8) {
                for (int b = 0; b < nblk; b++) {
                    const unsigned k = kbase + b;
                    pipe_wait_full(pm.full[kFH], k, wait_cycles);
                    const typename PipeMem<PROF>::EntH &eh = pm.h[k & SL];
                    const bool compute = eh.compute != 0, count = eh.count != 0;
                    const int m = eh.m;
                    const float e = eh.e;
                    if (compute) {
#pragma unroll
                        for (int r = 0; r < 4; r++) wm.W[lane + 32 * r] = eh.pw[lane + 32 * r];
                        if (lane == 0) wm.W[128] = eh.pw[128];
                    }
                    pipe_signal(pm.empty[kFH], k, lane);       // (its __syncwarp orders the wm.W stores before the triangles)
                    if (compute) {
                        const float out = comp_ceps_back(wm, bm, br, lane, e, 0.0f, 0.0f, 0.0f);
                        if (lane < 14) a.ceps[(ceps0 + m) * 14 + lane] = out;
                    }
                    if (count) s.n_ceps = m + 1;
                }
                if (lane == 0 && a.n_ceps) a.n_ceps[u] = s.n_ceps;
            }
        } else {
            // ---- frame-dropping statistics of stage 0 (8 kHz profile) -----------------------------------------------------------
            if constexpr (PROF == 8) {
                for (int b = 0; b < nblk; b++) {
                    const unsigned k = kbase + b;
                    pipe_wait_full(pm.full[kFB], k, wait_cycles);
                    const typename PipeMem<PROF>::EntB &eb = pm.b[k & SL];
                    const bool ran0 = eb.ran0 != 0;
                    s.nb_frame0 = eb.nb_frame0;
                    const float wsum = eb.wsum[lane], wsq = eb.wsq[lane], mw = eb.mw[lane];
                    pipe_signal(pm.empty[kFB], k, lane);
                    StageFlags fl = {0, 0, 0};
                    if (ran0) fd_stats<PROF>(wm, s, wsum, wsq, mw, fl);
                    pipe_wait_empty(pm.empty[kFG], k, wait_cycles);
                    typename PipeMem<PROF>::EntG &eg = pm.g[k & SL];
                    if (lane == 0) { eg.var = fl.var; eg.spec = fl.spec; eg.mel = fl.mel; }
                    pipe_signal(pm.full[kFG], k, lane);
                }
            }
        }
        kbase += (unsigned)nblk;
        __syncwarp();
    }
    bank_tmem_release<PROF>(bm);
#ifdef SE_PIPE_STATS
    if (lane == 0) {
        atomicAdd(&g_pipe_stats[warp][0], (unsigned long long)wait_cycles[0]);
        atomicAdd(&g_pipe_stats[warp][1], (unsigned long long)wait_cycles[1]);
        atomicAdd(&g_pipe_stats[warp][2], (unsigned long long)(clock64() - t_start));
        atomicAdd(&g_pipe_stats[warp][3], 1ull);
    }
#endif
}

template <int PROF> static size_t pipe_smem_bytes()
{
    return ((sizeof(BlockMem<PROF>) + 15) & ~size_t(15)) + PipeWarps<PROF>::value * sizeof(WarpMem<PROF>) + sizeof(PipeMem<PROF>);
}

// CTAs of the pipelined kernel the GPU holds at once (0: cannot run)
template <int PROF, typename IN_T> static int pipe_capacity(se_ctx *ctx)
{
    const size_t smem = pipe_smem_bytes<PROF>();
    auto kern = ns_pipe_kernel<PROF, IN_T>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, PipeWarps<PROF>::value * 32, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    return per_sm * ctx->sm_count;
}

template <int PROF, typename IN_T>
static int launch_ns_pipe(se_ctx *ctx, const NsLaunch &a, const NsDevTables &tb, int kid, int capacity)
{
    const size_t smem = pipe_smem_bytes<PROF>();
    int blocks = a.n_utts < capacity ? a.n_utts : capacity;
    prof_begin(ctx, kid);
    ns_pipe_kernel<PROF, IN_T><<<blocks, PipeWarps<PROF>::value * 32, smem, ctx->cur->stream>>>(a, tb, ctx->d_cc);
    prof_end(ctx, kid);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

// The pipelined kernel serves plain batch launches (no persistent state, feature chain or traces) whose utterances all
// get a CTA of their own at once; ctx->ns_pipeline: 0 = that rule, 1 = never, 2 = whenever it is able to.
template <int PROF, typename IN_T>
static bool try_ns_pipe(se_ctx *ctx, const NsLaunch &a, const NsDevTables &tb, int kid, int *rc)
{
    if (ctx->ns_pipeline == 1 || a.state || a.features || a.spec_trace || a.produced || a.n_utts <= 0) return false;
    const int cap = pipe_capacity<PROF, IN_T>(ctx);
    if (cap <= 0) return false;
    if (ctx->ns_pipeline != 2 && a.n_utts > cap) return false;
    *rc = launch_ns_pipe<PROF, IN_T>(ctx, a, tb, kid, cap);
    return true;
}

// Utterances a launch with these arguments would run on the pipelined kernel at once (0: it would not use it); lets the
// host-buffer pipeline cut a first chunk of exactly that size, whose results start flowing back after ~2 ms.
int ns_pipe_capacity(se_ctx *ctx, int profile, const NsLaunch &a)
{
    if (ctx->ns_pipeline == 1 || a.state || a.features || a.spec_trace || a.produced) return 0;
    if (profile == 8) return a.is_f32 ? pipe_capacity<8, float>(ctx) : pipe_capacity<8, int16_t>(ctx);
    return a.is_f32 ? pipe_capacity<16, float>(ctx) : pipe_capacity<16, int16_t>(ctx);
}

}  // namespace SE_NSPACE

#ifdef SE_PIPE_STATS
extern "C" int se_debug_pipe_stats(unsigned long long *out32, int reset)
{
    if (cudaMemcpyFromSymbol(out32, SE_NSPACE::g_pipe_stats, sizeof(unsigned long long) * 32) != cudaSuccess) return -1;
    if (reset) {
        unsigned long long z[32] = {0};
        cudaMemcpyToSymbol(SE_NSPACE::g_pipe_stats, z, sizeof z);
    }
    return 0;
}
#endif
