// Persistent fused NoiseSup kernel + host launchers.  Included by se_ns.cu.
#pragma once

#include "se_bulk.cuh"

namespace SE_NSPACE {

#ifdef SE_NS_NO_BULK
constexpr bool kBulk = false;
#else
constexpr bool kBulk = true;
#endif

// warps (= concurrent utterances) per CTA; 2 CTAs per SM.  NoiseSup and the 8 kHz feature chain run 8 warps at 128
// registers; the 16 kHz-mode variant keeps more state live per warp and trades occupancy (6 warps per CTA, 12 per SM)
// for 157 registers per thread.
#ifndef SE_NS_WARPS
#define SE_NS_WARPS 8
#endif
// FEAT: 0 = NoiseSup (+ raw CompCeps), 1 = 8 kHz feature chain, 2 = 16 kHz feature chain (se_ns_hp16k.cuh)
#ifndef SE_NS_WARPS_F1
#define SE_NS_WARPS_F1 8
#endif
#ifndef SE_NS_WARPS_F2
#define SE_NS_WARPS_F2 6
#endif
template <int FEAT> struct NsWarps { static constexpr int value = FEAT == 2 ? SE_NS_WARPS_F2 : (FEAT ? SE_NS_WARPS_F1 : SE_NS_WARPS); };

template <int PROF> __device__ __forceinline__ void load_block_tables(BlockMem<PROF> &bm, const NsDevTables &tb,
                                                                     const CcDevTables &cc)
{
    using P = Prof<PROF>;
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < P::FL; i += nt) bm.win[i] = tb.sig_window[i];
    for (int i = tid; i < 129; i += nt) bm.twh[i] = reinterpret_cast<const float2 *>(tb.tw_half)[i];
    for (int i = tid; i < (int)(sizeof(bm.bankT) / sizeof(float)); i += nt) bm.bankT[i] = tb.bankT[i];
    for (int i = tid; i < 32; i += nt) { bm.bank_start[i] = tb.bank_start[i]; bm.bank_len[i] = tb.bank_len[i]; }
#ifdef SE_NS_EXACT
    for (int i = tid; i < 25 * 16; i += nt) bm.idctT[i] = tb.idctT[i];
    for (int i = tid; i < 9; i += nt) bm.firwin[i] = tb.fir_window[i];
    if (tid == 0) { bm.rf.ops = tb.rf_ops; bm.rf.lvl = tb.rf_lvl; bm.rf.tw = tb.rf_tw; bm.rf.m = tb.rf_m; }
#endif
    if constexpr (P::CEPS) {
        for (int i = tid; i < 200; i += nt) bm.ham[i] = cc.hamming[i];
        for (int i = tid; i < (int)(sizeof(bm.triT) / sizeof(float)); i += nt) bm.triT[i] = cc.triT[i];
        for (int i = tid; i < 32; i += nt) { bm.tri_start[i] = cc.tri_start[i]; bm.tri_len[i] = cc.tri_len[i]; }
        for (int i = tid; i < 23 * 16; i += nt) bm.dctT[i] = cc.dctT[i];
    }
}

template <typename T> __device__ __forceinline__ float to_f32(T v) { return (float)v; }

// FEAT (P8 only): a.ceps receives 15-column feature rows (WaveProc -> CompCeps -> PostProc -> VAD,
// se_ns_feat.cuh) instead of 14-column raw cepstra; a.n_ceps counts released rows.
template <int PROF, typename IN_T, int FEAT>
#ifndef SE_NS_MINBLOCKS
#define SE_NS_MINBLOCKS 2
#endif
__global__ void __launch_bounds__(NsWarps<FEAT>::value * 32, (FEAT ? 2 : SE_NS_MINBLOCKS))
ns_stream_kernel(const NsLaunch a, const NsDevTables tb, const CcDevTables cc)
{
    using P = Prof<PROF>;
    static_assert(!FEAT || PROF == 8, "the feature chain exists for the 8 kHz profile only");
    constexpr bool HP16 = FEAT == 2;
    constexpr int SIN = HP16 ? 160 : P::S;      // input samples per call (16 kHz mode: two per low-band sample)
    constexpr int NRI = HP16 ? 5 : P::NR;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BlockMem<PROF> &bm = *reinterpret_cast<BlockMem<PROF> *>(smem_raw);
    WarpMem<PROF> *wms = reinterpret_cast<WarpMem<PROF> *>(smem_raw + ((sizeof(BlockMem<PROF>) + 15) & ~size_t(15)));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpMem<PROF> &wm = wms[warp];
    FeatMem *fms = reinterpret_cast<FeatMem *>(wms + NsWarps<FEAT>::value);
    FeatMem &fm = fms[FEAT ? warp : 0];
    Hp16Mem *hms = reinterpret_cast<Hp16Mem *>(fms + (FEAT ? NsWarps<FEAT>::value : 0));
    Hp16Mem &hm = hms[HP16 ? warp : 0];
    Hp16Block &hb = *reinterpret_cast<Hp16Block *>(hms + (HP16 ? NsWarps<FEAT>::value : 0));

    load_block_tables<PROF>(bm, tb, cc);
    if constexpr (HP16) {
        for (int i = threadIdx.x; i < 3 * 64; i += blockDim.x) hb.hpfb[i] = cc.hpfb[i];
        for (int i = threadIdx.x; i < 26 * 16; i += blockDim.x) hb.dct26T[i] = cc.dct26T[i];
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
        for (int i = lane; i < 132; i += 32) wm.W[i] = 0.0f;   // zero-weight taps may read past the spectrum
    const float eps = tb.eps;
    const IN_T *all = reinterpret_cast<const IN_T *>(a.samples);
    unsigned in_phase = 0;                 // bit s: parity the next wait on staging slot s uses (flips per completed wait)
#ifndef SE_NS_NO_BULK
    if (lane == 0) {
        for (int k = 0; k < kInSlots; k++) se_bulk::mbar_init(&wm.in_bar[k], 1);
        se_bulk::mbar_init_fence();
    }
    __syncwarp();
#endif

    for (;;) {
        int u = 0;
        if (lane == 0) u = atomicAdd(a.work_counter, 1);
        u = __shfl_sync(0xffffffffu, u, 0);
        if (u >= a.n_utts) break;
        u += a.utt_begin;
        if (a.order) u = a.order[u];

        const int64_t off0 = a.offsets[u] - a.sample_base;
        const int64_t n = a.offsets[u + 1] - a.offsets[u];
        const int nblk = (int)(n / SIN);
        const IN_T *in = all + off0 + a.in_shift;

        // ---- stream state -------------------------------------------------------------------
        Scal s;
        float noise1[P::NB], den1[P::NB], noise2[P::NB], den2[P::NB];
        float *ring1 = wm.ring1 + 8, *ring2 = wm.ring2 + 8;
        NsState *st = a.state ? a.state + u : nullptr;
        if (st && a.load_state) {
            s.nb_in1 = st->nb_in1; s.nb_in2 = st->nb_in2; s.nb_out2 = st->nb_out2;
            s.nb_frame0 = st->nb_frame0; s.nb_frame1 = st->nb_frame1;
            s.flag_vad = st->flag_vad; s.hang_over = st->hang_over; s.nb_speech = st->nb_speech;
            s.onset = st->onset; s.n_ceps = st->n_ceps;
            s.n_rows = 0;     // rows are counted per push (gains rows of this call start at 0)
            s.mean_en = st->mean_en; s.dc_x = st->dc_x; s.dc_y = st->dc_y;
            s.den_en0 = st->den_en0; s.den_en1 = st->den_en1; s.den_en2 = st->den_en2;
            s.low_snr = st->low_snr; s.alfa_gf = st->alfa_gf;
            s.fd_mel_mean = st->fd_mel_mean; s.fd_var_mean = st->fd_var_mean; s.fd_acc_test = st->fd_acc_test;
            s.fd_spec_mean = st->fd_spec_mean; s.fd_mel_val0 = st->fd_mel_val0; s.fd_mel_val1 = st->fd_mel_val1;
            s.fd_spec_val = st->fd_spec_val; s.fd_in_vadq = st->fd_in_vadq;
#pragma unroll
            for (int r = 0; r < P::NB; r++) {
                const int b = lane + 32 * r;
                const bool v = b < P::SPEC;
                noise1[r] = v ? st->noise1[b] : eps; den1[r] = v ? st->den1[b] : 0.0f;
                noise2[r] = v ? st->noise2[b] : eps; den2[r] = v ? st->den2[b] : 0.0f;
                wm.prev[r * 32 + lane] = v ? st->prev_psd1[b] : 0.0f;
                wm.prev[(P::NB + r) * 32 + lane] = v ? st->prev_psd2[b] : 0.0f;
            }
            for (int i = lane - 8; i < P::RB + 8; i += 32) {      // core + both guards
                const int src = wrap<P::RB>(i);
                ring1[i] = st->ring1[src];
                ring2[i] = st->ring2[src];
            }
            if (P::CEPS)
                for (int i = lane; i < P::RB; i += 32) wm.ring3[i] = st->ring3[i];
        } else {
            scal_init(s);
#pragma unroll
            for (int r = 0; r < P::NB; r++) {
                noise1[r] = noise2[r] = eps;     // NoiseSup.c:948-953
                den1[r] = den2[r] = 0.0f;
                wm.prev[r * 32 + lane] = 0.0f;
                wm.prev[(P::NB + r) * 32 + lane] = 0.0f;
            }
            for (int i = lane; i < P::RB + 16; i += 32) {
                wm.ring1[i] = 0.0f;
                wm.ring2[i] = 0.0f;
            }
            if (P::CEPS)
                for (int i = lane; i < P::RB; i += 32) wm.ring3[i] = 0.0f;
        }
        SE_COLD(cold_fd_store(wm.cold, s));            // after scal_init / the state load above
        SE_COLD(cold_gf_store(wm.cold, s));
        FeatState fs;
        float fb_last = 0.0f;                             // lane i < 15: FeatureBuffer[i] as the last call left it
        if constexpr (FEAT) feat_state_init(fs, fm, lane);
        Hp16State hs;
        if constexpr (HP16) hp16_state_init(hs, hm, lane);
        __syncwarp();
        // cepstral rows are numbered per launch: a resumed stream continues at row 0 of this push
        const int ceps_row0 = (st && a.load_state && s.nb_out2 > 2) ? s.nb_out2 - 2 : 0;

        const int64_t blk0 = a.block_offsets ? a.block_offsets[u] - a.block_base : 0;
        const int64_t tr0 = a.trace_offsets ? a.trace_offsets[u] - a.trace_base : 0;
        const int64_t ceps0 = a.ceps_offsets ? a.ceps_offsets[u] - a.ceps_base : 0;
        const int64_t ceps_cap = a.ceps_offsets ? a.ceps_offsets[u + 1] - a.ceps_offsets[u] : 0;

        // ---- input ------------------------------------------------------------------------------
        // int16 input: groups of kGroup blocks (640 bytes) are posted two groups ahead as ONE bulk-async copy each into
        // this warp's two staging slots (se_bulk.cuh).  A copy starts at the 16-byte boundary below the group's first
        // sample (`mis` samples earlier) and is kSpan bytes long; groups whose span would leave [bulk_lo, bulk_hi) -- at
        // most the first and the last group of a buffer -- and the ragged last group of an utterance are staged with
        // ordinary loads and stores instead, behind the same barrier, so the consumer is the same either way.
        // float input (persistent streams, the plugin's float entry): software prefetch of the next block in registers.
#ifdef SE_NS_BULK_ALL
        constexpr bool BULK = kBulk && sizeof(IN_T) == 2;
#else
        // measured (B200, profiles/r02_kernel_ab.md): 160-sample blocks gain 1.3 % from it, the 80-sample blocks of the
        // 8 kHz profile lose 1.7 % (their register prefetch converts at consumption and is cheaper than wait + LDS)
        constexpr bool BULK = kBulk && sizeof(IN_T) == 2 && SIN == 160;
#endif
        constexpr int kGroup = kInGroupBytes / (SIN * 2);
        constexpr int kSpan = kInGroupBytes + 16;
        static_assert(kGroup * SIN * 2 == kInGroupBytes && kSpan == kInSlotBytes && kInSlots == 2, "staging slot layout");
        const uintptr_t in_addr = reinterpret_cast<uintptr_t>(in);
        const int mis = (int)(in_addr & 15) >> 1;
        const char *span0 = reinterpret_cast<const char *>(in_addr & ~uintptr_t(15));
        const int ngrp = (nblk + kGroup - 1) / kGroup;
        using NX_T = typename std::conditional<PROF == 8, IN_T, float>::type;
        NX_T nx[BULK ? 1 : NRI];
        (void)mis; (void)span0; (void)ngrp; (void)in_phase; (void)nx;       // unused in the instantiations without bulk input
#ifndef SE_NS_NO_BULK
        auto post_group = [&](int g) {
            const int slot = g & 1;
            const char *lo = span0 + (int64_t)g * kInGroupBytes;
            const bool whole = (g + 1) * kGroup <= nblk && lo >= a.bulk_lo && lo + kSpan <= a.bulk_hi;
            if (whole) {
                if (lane == 0) se_bulk::bulk_g2s(wm.in_stage[slot], lo, kSpan, &wm.in_bar[slot]);
            } else {
                int16_t *sp = reinterpret_cast<int16_t *>(wm.in_stage[slot]) + mis;
                const int cnt = (nblk - g * kGroup < kGroup ? nblk - g * kGroup : kGroup) * SIN;
                for (int i = lane; i < cnt; i += 32) sp[i] = (int16_t)in[(int64_t)g * (kGroup * SIN) + i];
                se_bulk::proxy_fence();
                __syncwarp();
                if (lane == 0) se_bulk::mbar_arrive(&wm.in_bar[slot]);
            }
        };
#endif
        if constexpr (BULK) {
#ifndef SE_NS_NO_BULK
            if (ngrp > 0) post_group(0);
            if (ngrp > 1) post_group(1);
#endif
        } else {
            // P8 keeps the prefetched samples in the input type and converts them one block later, when they are
            // consumed (converting right after the load makes the warp wait for the load at once): -3 % on P8.
            // P16 measured 1 % slower that way, so it converts at the load.
#pragma unroll
            for (int r = 0; r < NRI; r++) {
                const int i = lane + 32 * r;
                nx[r] = (nblk > 0 && i < SIN) ? NX_T(in[i]) : NX_T(0);
            }
        }

        for (int b = 0; b < nblk; b++) {
            float cur[NRI];
            float ss = 0.0f;
            bool nzl = false;
            if constexpr (BULK) {
#ifndef SE_NS_NO_BULK
                const int g = b / kGroup, k = b - g * kGroup, slot = g & 1;
                if (k == 0) {                                  // first block of a group: its copy has landed?
                    se_bulk::mbar_wait_warp(&wm.in_bar[slot], (in_phase >> slot) & 1u);
                    in_phase ^= 1u << slot;
                }
                const int16_t *sp = reinterpret_cast<const int16_t *>(wm.in_stage[slot]) + mis + k * SIN;
#pragma unroll
                for (int r = 0; r < NRI; r++) {
                    const int i = lane + 32 * r;
                    cur[r] = (i < SIN) ? (float)sp[i] : 0.0f;
                }
                if (k == kGroup - 1 && g + 2 < ngrp) {
                    __syncwarp();                              // every lane has read the slot: it can be refilled
                    post_group(g + 2);
                }
#endif
#pragma unroll
                for (int r = 0; r < NRI; r++) { ss += cur[r] * cur[r]; nzl |= (cur[r] != 0.0f); }
            } else {
#pragma unroll
                for (int r = 0; r < NRI; r++) { cur[r] = to_f32(nx[r]); ss += cur[r] * cur[r]; nzl |= (cur[r] != 0.0f); }
                if (b + 1 < nblk) {
#pragma unroll
                    for (int r = 0; r < NRI; r++) {
                        const int i = lane + 32 * r;
                        if (i < SIN) nx[r] = NX_T(in[(int64_t)(b + 1) * SIN + i]);
                    }
                }
            }
            // zero-frame gate: DoAdvProcess (ParmInterface.c:244-250, leading frames only) /
            // func_Wiener (NoiseSup.cpp:1159-1170, anywhere).  (int)FrameCheck != 0  <=>  energy >= 1.
            bool process = true;
            if (a.zero_check) {
                bool nonzero;
                if (sizeof(IN_T) == 2) nonzero = __any_sync(0xffffffffu, nzl);
                else nonzero = !(warp_sum(ss) < 1.0f);
                if (a.zero_check == 1) { process = nonzero || (s.onset != 0); if (process) s.onset = 1; }
                else process = nonzero;
            }

            const int64_t obase = (HP16 ? tr0 : off0) + (int64_t)b * P::S;   // 16 kHz mode: low-band output, 80 per call
            float o[P::CH];
            float hp_code = 0.0f, hp_fbss = 0.0f;
            bool produced = false, ran0 = false;
            StageFlags fl = {0, 0, 0};
            if (process) {
                const int slot = s.nb_in1 & 3;
                int hslot = 0;
                if constexpr (HP16) {
                    // Do16kProcessing (ParmInterface.c:259-260): the low band feeds NoiseSup, the high band its ring
                    hslot = s.nb_in1 % 6;
#pragma unroll
                    for (int r = 0; r < NRI; r++) hm.line[117 + lane + 32 * r] = cur[r];
                    __syncwarp();
                    qmf_split_warp(hm, ring1, slot, hslot, lane);
                } else {
#pragma unroll
                    for (int r = 0; r < P::NR; r++) {
                        const int i = lane + 32 * r;
                        if (i < P::S) ring_store<PROF>(ring1, slot, i, cur[r]);
                    }
                }
                s.nb_in1++;
                __syncwarp();
                float *trace = a.spec_trace ? a.spec_trace + ((PROF == 8 ? tr0 : blk0) + b) * 8 * P::SPEC : nullptr;
#pragma unroll 1
                for (int fstage = 0; fstage < 2; fstage++) {
                    // stage 0 runs from the 3rd block on, stage 1 from the 3rd stage-0 output on
                    // (NS_NB_FRAMES_LATENCY, NoiseSup.c:1144-1196); stage 1 sees stage 0's update of nb_in2
                    const bool run = fstage == 0 ? (s.nb_in1 - s.nb_in2 > 2) : (s.nb_in2 - s.nb_out2 > 2);
                    if (!run) continue;
                    float noise[P::NB], den[P::NB];
#pragma unroll
                    for (int r = 0; r < P::NB; r++) {
                        noise[r] = fstage ? noise2[r] : noise1[r];
                        den[r] = fstage ? den2[r] : den1[r];
                    }
                    float *grow = (PROF == 16 && fstage == 1 && a.gains) ? a.gains + (blk0 + s.n_rows) * kMel : nullptr;
                    float psd_hi[2];
                    if constexpr (HP16) {
                        // stage-1 hook first (NoiseSup.c:1307-1327 runs it after ApplyWF; nothing it reads changes in
                        // between, and fewer registers are live here).  nbFrame[1] is incremented by this stage's VAD.
                        if (fstage == 1) hp16_stage1(wm, bm, hb, hm, tw, lane, hs, hslot, s.nb_frame1 + 1, hp_code, hp_fbss);
                    }
                    ns_stage<PROF>(wm, bm, tw, ic, br, lane, s, noise, den, eps, fstage, o, fl, grow, trace,
                                   (HP16 && fstage == 0) ? psd_hi : nullptr);
                    if constexpr (HP16) {
                        if (fstage == 0) hp16_lowband_log(hs, psd_hi[0], psd_hi[1], lane);
                    }
#pragma unroll
                    for (int r = 0; r < P::NB; r++) {
                        if (fstage) { noise2[r] = noise[r]; den2[r] = den[r]; }
                        else { noise1[r] = noise[r]; den1[r] = den[r]; }
                    }
                    if (fstage == 0) {
                        ran0 = true;
                    } else {
                        s.n_rows++;
                        produced = true;
#ifdef SE_NS_EXACT
                        dc_offset<PROF>(o, s, lane, wm.xf);
#else
                        dc_offset<PROF>(o, s, lane);
#endif
                        if constexpr (HP16) hp16_push(hs, hp_code, hp_fbss, lane);
                    }
                }
            }

            // ---- outputs of this shift --------------------------------------------------------
            const int i0 = lane * P::CH;
            if constexpr (PROF == 8) {
                // float denoised -> ring3 (CompCeps history), then coalesced int16 / float stores
                float *stage_out = wm.W;     // 80 floats of scratch (W is dead here)
                if (produced) {
                    const int nb3 = ((s.nb_out2 - 1) & 3) * P::S;
#pragma unroll
                    for (int t = 0; t < P::CH; t++)
                        if (i0 + t < P::S) { wm.ring3[nb3 + i0 + t] = o[t]; stage_out[i0 + t] = o[t]; }
                } else {
#pragma unroll
                    for (int t = 0; t < P::CH; t++)
                        if (i0 + t < P::S) stage_out[i0 + t] = 0.0f;
                }
                __syncwarp();
                if (a.den_i16) {
                    // (FILE_TYPE) ptDenoised[i], ParmInterface.c:266: C cast = truncation toward zero
                    for (int i = lane; i < P::S / 2; i += 32) {
                        const float2 f = *reinterpret_cast<const float2 *>(stage_out + 2 * i);
                        short2 q;
                        q.x = (short)__float2int_rz(f.x);
                        q.y = (short)__float2int_rz(f.y);
                        if ((reinterpret_cast<uintptr_t>(a.den_i16 + obase) & 3) == 0) *reinterpret_cast<short2 *>(a.den_i16 + obase + 2 * i) = q;
                        else { a.den_i16[obase + 2 * i] = q.x; a.den_i16[obase + 2 * i + 1] = q.y; }
                    }
                }
                if (a.den_f32)
                    for (int i = lane; i < P::S; i += 32) a.den_f32[obase + i] = stage_out[i];
                if constexpr (FEAT) {
                    bool emit = false;
                    float c = 0.0f;
                    if (!process) {
                        emit = true;                     // null feature vector, ParmInterface.c:314-327
                        if (lane < 14) fb_last = 0.0f;
                    } else if (produced && s.nb_out2 >= 3) {
                        const int base = ((s.nb_out2 - 3) & 3) * P::S;
                        for (int i = lane; i < 201; i += 32) fm.frame[i] = wm.ring3[wrap<P::RB>(base + i - 1)];
                        __syncwarp();
                        wave_proc_warp(fm, lane);
                        if constexpr (HP16)
                            c = comp_ceps_core<true>(wm, bm, tw, br, lane, [&](int i) { return fm.frame[1 + i]; }, hb.dct26T,
                                                     hs.codeb, hs.hpb);
                        else
                            c = comp_ceps_core(wm, bm, tw, br, lane, [&](int i) { return fm.frame[1 + i]; });
                        c = post_proc_warp(c, fs, lane);
                        const bool speech = fl.spec || fl.mel || fl.var || (s.nb_speech > 4);
                        emit = vad_proc_warp(c, fs, fm, lane, speech, s.nb_frame0);
                        if (emit || lane < 14) fb_last = c;       // what the caller's FeatureBuffer holds after this call
                    }
                    if (emit) {
                        if (a.ceps && fs.rows < ceps_cap && lane < kFeatCols) a.ceps[(ceps0 + fs.rows) * kFeatCols + lane] = c;
                        fs.rows++;
                    }
                } else if (produced && a.ceps && s.nb_out2 >= 3) {
                    // third denoised shift onwards: frame m = nb_out2-3 (ParmInterface.c:275-293)
                    const int m = s.nb_out2 - 3;
                    if (m - ceps_row0 < ceps_cap)
                        comp_ceps(wm, bm, tw, br, lane, (m & 3) * P::S, a.ceps + (ceps0 + m - ceps_row0) * 14);
                    s.n_ceps = m + 1;
                }
                if (a.vad_trace && lane == 0) {
                    unsigned char t = 0;
                    if (process) {
                        t = (unsigned char)((s.flag_vad ? 1 : 0) | (fl.var << 1) | (fl.spec << 2) | (fl.mel << 3) |
                                            ((s.nb_speech > 4 ? 1 : 0) << 4) | (ran0 ? 32 : 0) | (produced ? 64 : 0));
                    } else {
                        t = 128;
                    }
                    a.vad_trace[tr0 + b] = t;
                }
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
            if (a.produced && lane == 0)
                a.produced[(PROF == 8 ? tr0 : blk0) + b] = (produced ? 1 : 0) | (ran0 ? 2 : 0) | (process ? 0 : 4);
            __syncwarp();
        }

        if constexpr (FEAT) {
            if (a.flush) {                                // FlushAdvProcess, ParmInterface.c:347-355: 6 more rows
                int fc = s.nb_frame0;                     // FEParamsX.FrameCounter as DoNoiseSup last published it
                for (int k = 0; k < 6; k++) {
                    vad_flush_warp(fb_last, fs, fm, lane, fc);
                    if (a.ceps && fs.rows < ceps_cap && lane < kFeatCols) a.ceps[(ceps0 + fs.rows) * kFeatCols + lane] = fb_last;
                    fs.rows++;
                }
            }
        }

        // tail (len % S): the reference leaves it untouched; the batch API defines it as 0
        if (PROF == 8) {
            const int64_t den_n = HP16 ? (a.trace_offsets ? a.trace_offsets[u + 1] - a.trace_offsets[u] : 0) : n;
            const int64_t den0 = HP16 ? tr0 : off0;
            for (int64_t i = (int64_t)nblk * P::S + lane; i < den_n; i += 32) {
                if (a.den_i16) a.den_i16[den0 + i] = 0;
                if (a.den_f32) a.den_f32[den0 + i] = 0.0f;
            }
        } else if (a.out_f32) {
            for (int64_t i = (int64_t)nblk * P::S + lane; i < n; i += 32) a.out_f32[off0 + i] = 0.0f;
        }
        if (lane == 0) {
            if (a.n_ceps) {
                if constexpr (FEAT) a.n_ceps[u] = fs.rows;
                else a.n_ceps[u] = s.n_ceps - ceps_row0;
            }
            if (a.n_rows) a.n_rows[u] = s.n_rows;
        }

        if (st) {
            SE_COLD(cold_fd_load(s, wm.cold));
            SE_COLD(cold_gf_load(s, wm.cold));
            if (lane == 0) {
                st->nb_in1 = s.nb_in1; st->nb_in2 = s.nb_in2; st->nb_out2 = s.nb_out2;
                st->nb_frame0 = s.nb_frame0; st->nb_frame1 = s.nb_frame1;
                st->flag_vad = s.flag_vad; st->hang_over = s.hang_over; st->nb_speech = s.nb_speech;
                st->onset = s.onset; st->n_ceps = s.n_ceps; st->n_rows = s.n_rows;
                st->mean_en = s.mean_en; st->dc_x = s.dc_x; st->dc_y = s.dc_y;
                st->den_en0 = s.den_en0; st->den_en1 = s.den_en1; st->den_en2 = s.den_en2;
                st->low_snr = s.low_snr; st->alfa_gf = s.alfa_gf;
                st->fd_mel_mean = s.fd_mel_mean; st->fd_var_mean = s.fd_var_mean; st->fd_acc_test = s.fd_acc_test;
                st->fd_spec_mean = s.fd_spec_mean; st->fd_mel_val0 = s.fd_mel_val0; st->fd_mel_val1 = s.fd_mel_val1;
                st->fd_spec_val = s.fd_spec_val; st->fd_in_vadq = s.fd_in_vadq;
            }
#pragma unroll
            for (int r = 0; r < P::NB; r++) {
                const int b = lane + 32 * r;
                if (b < P::SPEC) {
                    st->noise1[b] = noise1[r]; st->den1[b] = den1[r]; st->prev_psd1[b] = wm.prev[r * 32 + lane];
                    st->noise2[b] = noise2[r]; st->den2[b] = den2[r]; st->prev_psd2[b] = wm.prev[(P::NB + r) * 32 + lane];
                }
            }
            for (int i = lane; i < P::RB; i += 32) {
                st->ring1[i] = ring1[i];
                st->ring2[i] = ring2[i];
                if (P::CEPS) st->ring3[i] = wm.ring3[i];
            }
        }
        __syncwarp();
    }
    bank_tmem_release<PROF>(bm);
}

// ------------------------------------------------------------------------------------------------
// test kernel: batched rfft in the reference's output layout (one warp per transform)
// ------------------------------------------------------------------------------------------------
#ifdef SE_NS_EXACT
// exact mode: the reference's own butterfly schedule (rfft_exact), one warp per transform
__global__ void rfft_kernel(float *x, int n, int count, const RfPlanDev pl)
{
    __shared__ float xs[4][512];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * 4 + warp;
    if (idx >= count) return;
    float *p = x + (int64_t)idx * n;
    const int bits = (n == 256) ? 8 : 9;
    for (int i = lane; i < n; i += 32) xs[warp][__brev((unsigned)i) >> (32 - bits)] = p[i];
    __syncwarp();
    rfft_exact(xs[warp], pl, lane);
    for (int i = lane; i < n; i += 32) p[i] = xs[warp][i];
}
#else
__global__ void rfft_kernel(float *x, int n, int count, const float *tw_half)
{
    __shared__ float2 Zs[4][kZ], Z2s[4][kZ];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * 4 + warp;
    if (idx >= count) return;
    float *p = x + (int64_t)idx * n;
    const float2 *twh = reinterpret_cast<const float2 *>(tw_half);
    FftTw tw;
    fft_tw_init(tw, lane);
    float2 *Z = Zs[warp], *Z2 = Z2s[warp];
    if (n == 256) {
        float2 v[1][4];
        for (int j = 0; j < 4; j++) v[0][j] = reinterpret_cast<const float2 *>(p)[lane + 32 * j];
        __syncwarp();
        float2 *const zs[1] = {Z};
        fft128_spectrum<1>(v, zs, tw, lane);
        __syncwarp();
        for (int k = lane; k <= 128; k += 32) {
            const float2 X = rsplit256(Z, k, twh[k]);
            p[k] = X.x;
            if (k > 0 && k < 128) p[256 - k] = X.y;
        }
    } else {   // n == 512, m == 8 : the plugin's truncated schedule
        float2 v[2][4];
        for (int j = 0; j < 4; j++) {
            const float4 q = reinterpret_cast<const float4 *>(p)[lane + 32 * j];
            v[0][j] = make_float2(q.x, q.z);
            v[1][j] = make_float2(q.y, q.w);
        }
        __syncwarp();
        float2 *const zs[2] = {Z, Z2};
        fft128_spectrum<2>(v, zs, tw, lane);
        __syncwarp();
        for (int k = lane; k <= 128; k += 32) {
            const float2 A = rsplit256(Z, k, twh[k]);
            p[k] = A.x;
            if (k > 0 && k < 128) p[256 - k] = A.y;
        }
        for (int k = lane; k <= 64; k += 32) {
            float2 B, C;
            rsplit_pair(Z2, k, B, C);
            p[256 + k] = B.x;
            p[384 + k] = C.x;
            if (k > 0 && k < 64) { p[384 - k] = B.y; p[512 - k] = C.y; }
        }
    }
}

#endif

// module-level kernel: Do16kProcessing over whole signals (one warp per signal, 160 samples per step)
__global__ void qmf_split_kernel(const int16_t *x, int64_t n, int count, float *lp, float *hp)
{
    __shared__ Hp16Mem hms[4];
    __shared__ float rings[4][Prof<8>::RB + 16];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * 4 + warp;
    if (idx >= count) return;
    Hp16Mem &hm = hms[warp];
    float *ring = rings[warp] + 8;
    for (int i = lane; i < 280; i += 32) hm.line[i] = 0.0f;
    __syncwarp();
    const int16_t *in = x + (int64_t)idx * n;
    for (int64_t b = 0; b < n / 160; b++) {
        for (int i = lane; i < 160; i += 32) hm.line[117 + i] = (float)in[b * 160 + i];
        __syncwarp();
        qmf_split_warp(hm, ring, 1, 0, lane);
        for (int i = lane; i < 80; i += 32) {
            lp[(int64_t)idx * (n / 2) + b * 80 + i] = ring[80 + i];
            hp[(int64_t)idx * (n / 2) + b * 80 + i] = hm.hpring[i];
        }
        __syncwarp();
    }
}

// test kernel: DoWaveProc on independent frames (one warp per frame)
__global__ void wave_proc_kernel(float *frames, int count)
{
    __shared__ FeatMem fms[4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int idx = blockIdx.x * 4 + warp;
    if (idx >= count) return;
    FeatMem &fm = fms[warp];
    float *p = frames + (int64_t)idx * 200;
    for (int i = lane; i < 200; i += 32) fm.frame[1 + i] = p[i];
    if (lane == 0) fm.frame[0] = 0.0f;
    __syncwarp();
    wave_proc_warp(fm, lane);
    for (int i = lane; i < 200; i += 32) p[i] = fm.frame[1 + i];
}

// test / module-level kernel: DoCompCeps on independent 201-sample frames (frame[0] = sample -1), one warp each
__global__ void __launch_bounds__(128) comp_ceps_kernel(const float *frames, float *coef, int count, const NsDevTables tb,
                                                        const CcDevTables cc)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BlockMem<8> &bm = *reinterpret_cast<BlockMem<8> *>(smem_raw);
    WarpMem<8> *wms = reinterpret_cast<WarpMem<8> *>(smem_raw + ((sizeof(BlockMem<8>) + 15) & ~size_t(15)));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpMem<8> &wm = wms[warp];
    load_block_tables<8>(bm, tb, cc);
    __syncthreads();
    FftTw tw;
    fft_tw_init(tw, lane);
    BankRegs br;
    bank_regs_init(br, tb, cc, lane);
#ifdef SE_NS_CONST_SMEM
    {
        IdctRegs ic0 = {};
        if (warp == 0) { lane_consts_store(bm.lane_tw, bm.lane_ic, tw, ic0, lane); bank_regs_store(bm.lane_br, br, lane); }
        __syncthreads();
    }
#endif
    for (int i = lane; i < 132; i += 32) wm.W[i] = 0.0f;
    __syncwarp();
    for (int idx = blockIdx.x * 4 + warp; idx < count; idx += gridDim.x * 4) {
        const float *f = frames + (int64_t)idx * 201;
        const float out = comp_ceps_core(wm, bm, tw, br, lane, [&](int i) { return f[1 + i]; });
        if (lane < 14) coef[(int64_t)idx * 14 + lane] = out;
        __syncwarp();
    }
}

}  // namespace SE_NSPACE
#ifndef SE_NS_EXACT
#include "se_ns_pipe.cuh"
#endif
namespace SE_NSPACE {

// ------------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------------
template <int PROF, int FEAT> static size_t ns_smem_bytes()
{
    return ((sizeof(BlockMem<PROF>) + 15) & ~size_t(15)) + NsWarps<FEAT>::value * sizeof(WarpMem<PROF>) +
           (FEAT ? NsWarps<FEAT>::value * sizeof(FeatMem) : 0) +
           (FEAT == 2 ? NsWarps<FEAT>::value * sizeof(Hp16Mem) + sizeof(Hp16Block) : 0);
}

template <int PROF, typename IN_T, int FEAT = 0>
static int launch_ns(se_ctx *ctx, const NsLaunch &a, const NsDevTables &tb, int kid)
{
    if (a.n_utts <= 0) return SE_OK;
#ifndef SE_NS_EXACT
    if constexpr (FEAT == 0) {      // few utterances: one CTA per utterance, a block's work pipelined over its warps
        int rc = SE_OK;
        if (try_ns_pipe<PROF, IN_T>(ctx, a, tb, kid, &rc)) return rc;
    }
#endif
    const size_t smem = ns_smem_bytes<PROF, FEAT>();
    auto kern = ns_stream_kernel<PROF, IN_T, FEAT>;
    SE_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    constexpr int warps = NsWarps<FEAT>::value;
    SE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, warps * 32, smem));
    if (per_sm < 1) per_sm = 1;
    int blocks = (a.n_utts + warps - 1) / warps;
    const int cap = ctx->sm_count * per_sm;
    if (blocks > cap) blocks = cap;
    SE_CUDA(cudaMemsetAsync(a.work_counter, 0, sizeof(int), ctx->cur->stream));
    prof_begin(ctx, kid);
    kern<<<blocks, warps * 32, smem, ctx->cur->stream>>>(a, tb, ctx->d_cc);
    prof_end(ctx, kid);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_ns_p8(se_ctx *ctx, const NsLaunch &a)
{
    if (a.features) {
        if (a.is_f32 || a.state) { set_error("feature chain: int16 batch input only"); return SE_EUNSUPPORTED; }
        if (a.features == 2) return launch_ns<8, int16_t, 2>(ctx, a, ctx->d_p8, SE_K_NS_P8);
        return launch_ns<8, int16_t, 1>(ctx, a, ctx->d_p8, SE_K_NS_P8);
    }
    return a.is_f32 ? launch_ns<8, float>(ctx, a, ctx->d_p8, SE_K_NS_P8)
                    : launch_ns<8, int16_t>(ctx, a, ctx->d_p8, SE_K_NS_P8);
}

int launch_ns_p16(se_ctx *ctx, const NsLaunch &a)
{
    return a.is_f32 ? launch_ns<16, float>(ctx, a, ctx->d_p16, SE_K_NS_P16)
                    : launch_ns<16, int16_t>(ctx, a, ctx->d_p16, SE_K_NS_P16);
}

int launch_qmf_split(se_ctx *ctx, const int16_t *x, int64_t n, int count, float *lp, float *hp)
{
    if (count <= 0 || n <= 0) return SE_OK;
    qmf_split_kernel<<<(count + 3) / 4, 128, 0, ctx->cur->stream>>>(x, n, count, lp, hp);
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_wave_proc(se_ctx *ctx, float *frames, int count)
{
    if (count <= 0) return SE_OK;
    wave_proc_kernel<<<(count + 3) / 4, 128, 0, ctx->cur->stream>>>(frames, count);
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_comp_ceps(se_ctx *ctx, const float *frames, float *coef, int count)
{
    if (count <= 0) return SE_OK;
    const size_t smem = ((sizeof(BlockMem<8>) + 15) & ~size_t(15)) + 4 * sizeof(WarpMem<8>);
    SE_CUDA(cudaFuncSetAttribute(comp_ceps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int blocks = (count + 3) / 4;
    if (blocks > ctx->sm_count * 4) blocks = ctx->sm_count * 4;
    comp_ceps_kernel<<<blocks, 128, smem, ctx->cur->stream>>>(frames, coef, count, ctx->d_p8, ctx->d_cc);
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_rfft(se_ctx *ctx, float *x, int n, int m, int count)
{
    if (!((n == 256 && m == 8) || (n == 512 && m == 8))) {
        set_error("se_rfft_batch: only (256,8) and (512,8) are on the hot path");
        return SE_EUNSUPPORTED;
    }
    if (count <= 0) return SE_OK;
#ifdef SE_NS_EXACT
    const NsDevTables &t = (n == 256) ? ctx->d_p8 : ctx->d_p16;
    RfPlanDev pl;
    pl.ops = t.rf_ops; pl.lvl = t.rf_lvl; pl.tw = t.rf_tw; pl.m = t.rf_m;
    rfft_kernel<<<(count + 3) / 4, 128, 0, ctx->cur->stream>>>(x, n, count, pl);
#else
    rfft_kernel<<<(count + 3) / 4, 128, 0, ctx->cur->stream>>>(x, n, count, ctx->d_p8.tw_half);
#endif
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

}  // namespace SE_NSPACE
