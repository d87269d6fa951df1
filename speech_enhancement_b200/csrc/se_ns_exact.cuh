// The "exact" arithmetic mode of the fused NoiseSup kernel (se_ctx_set_exact): the same warp-per-utterance kernel
// (se_ns_kernel.cuh) with every value computed by the reference's own operations in the reference's own order --
// its split-radix rfft schedule (etsi/cpp/rfft.c:45-180), sequential float sums, IEEE division / square root and the
// double-promoted sub-expressions of the C sources.  This translation unit is compiled with
// -fmad=false (no FMA contraction), so plain C expressions evaluate as gcc -O2 -ffp-contract=off evaluates the
// reference; the only remaining freedom is libm (log / log10 / exp in double, <= 1 ulp in both libraries before the
// result is rounded to float).
//
// Parallelism: butterflies of one FFT level, PSD bins, filter-bank bands, IDCT taps, FIR output samples and DCT
// coefficients are independent and are spread over the lanes; every SUM is evaluated by one lane (or redundantly by
// all lanes when the result is warp-uniform) in the reference's order.  Replaces frame_psd / ns_stage / dc_offset /
// comp_ceps_core of se_ns_stage.cuh; included from there when SE_NS_EXACT is defined.
#pragma once

namespace SE_NSPACE {

// rfft(x, n, m) on x[] in shared memory, bit-reversed on entry (the digit-reverse counter of rfft.c:56-75 is the
// plain bit reversal of the index).  One level at a time; the ops of a level touch disjoint elements.
__device__ __forceinline__ void rfft_exact(float *x, const RfPlanDev &pl, int lane)
{
    const int *lvl = pl.lvl;
    for (int op = lvl[0] + lane; op < lvl[1]; op += 32) {          // length-two butterflies, rfft.c:77-92
        const int i0 = (int)pl.ops[op];
        const float a0 = x[i0], a1 = x[i0 + 1];
        x[i0] = a0 + a1;
        x[i0 + 1] = a0 - a1;
    }
    __syncwarp();
    for (int k = 1; k < pl.m; k++) {                                // L-shaped butterflies, rfft.c:94-179
        const int n2 = 2 << k, n4 = n2 >> 2, n8 = n2 >> 3;
        const float4 *tw = reinterpret_cast<const float4 *>(pl.tw) + lvl[16 + k];
        for (int op = lvl[k] + lane; op < lvl[k + 1]; op += 32) {
            const unsigned o = pl.ops[op];
            const int i = (int)(o & 0xffffu), j = (int)(o >> 16);
            if (j == 0) {
                int i1 = i, i2 = i1 + n4, i3 = i2 + n4, i4 = i3 + n4;
                float t1 = x[i4] + x[i3];
                x[i4] = x[i4] - x[i3];
                x[i3] = x[i1] - t1;
                x[i1] = x[i1] + t1;
                if (n4 != 1) {
                    i1 += n8; i2 += n8; i3 += n8; i4 += n8;
                    t1 = (float)((double)(x[i3] + x[i4]) / 1.41421356237309504880);
                    const float t2 = (float)((double)(x[i3] - x[i4]) / 1.41421356237309504880);
                    x[i4] = x[i2] - t1;
                    x[i3] = -x[i2] - t1;
                    x[i2] = x[i1] - t2;
                    x[i1] = x[i1] + t2;
                }
            } else {
                const float4 w = tw[j - 1];                         // cc1, ss1, cc3, ss3
                const int i1 = i + j, i2 = i1 + n4, i3 = i2 + n4, i4 = i3 + n4;
                const int i5 = i + n4 - j, i6 = i5 + n4, i7 = i6 + n4, i8 = i7 + n4;
                float t1 = x[i3] * w.x + x[i7] * w.y;
                float t2 = x[i7] * w.x - x[i3] * w.y;
                float t3 = x[i4] * w.z + x[i8] * w.w;
                float t4 = x[i8] * w.z - x[i4] * w.w;
                const float t5 = t1 + t3, t6 = t2 + t4;
                t3 = t1 - t3;
                t4 = t2 - t4;
                t2 = x[i6] + t6;
                x[i3] = t6 - x[i6];
                x[i8] = t2;
                t2 = x[i2] - t3;
                x[i7] = -x[i2] - t3;
                x[i4] = t2;
                t1 = x[i1] + t5;
                x[i6] = x[i1] - t5;
                x[i1] = t1;
                t1 = x[i5] + t4;
                x[i5] = x[i5] - t4;
                x[i2] = t1;
            }
        }
        __syncwarp();
    }
}

// DoSigWindowing (NoiseSup.c:218-231) -> rfft -> FFTtoPSD (:249-270): psd[r] = bin lane + 32 r.
template <int PROF, int RB = Prof<PROF>::RB>
__device__ __forceinline__ void frame_psd(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const FftTw &, int lane,
                                          const float *ring, int wpos, float (&psd)[Prof<PROF>::NB])
{
    using P = Prof<PROF>;
    constexpr int N = (PROF == 8) ? 256 : 512, BITS = (PROF == 8) ? 8 : 9;
    float *x = wm.xf;
    for (int i = lane; i < N; i += 32) {
        float v = 0.0f;
        if (i < P::FL) v = ring[wrap<RB>(wpos + i)] * bm.win[i];
        x[__brev((unsigned)i) >> (32 - BITS)] = v;
    }
    __syncwarp();
    rfft_exact(x, bm.rf, lane);
    auto pw = [&](int k) -> float {
        if (k == 0) return x[0] * x[0];
        if (k == N / 2) return x[N / 2] * x[N / 2];
        return x[k] * x[k] + x[N - k] * x[N - k];
    };
#pragma unroll
    for (int r = 0; r < P::NB; r++) {
        const int b = lane + 32 * r;
        psd[r] = 0.0f;
        if (b < P::SPEC - 1) psd[r] = (float)((double)(pw(2 * b) + pw(2 * b + 1)) / 2.0);
        else if (b == P::SPEC - 1) psd[r] = pw(2 * b);
    }
    __syncwarp();
}

// One Wiener-filter stage, DoNoiseSup's fstage loop body (NoiseSup.c:1144-1367; plugin NoiseSup.cpp:1218-1374).
template <int PROF>
__device__ __forceinline__ void ns_stage(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const FftTw &tw, const IdctRegs &,
                                         const BankRegs &, int lane, Scal &s, float (&noise)[Prof<PROF>::NB], float (&den)[Prof<PROF>::NB],
                                         float eps, const int fstage, float (&o)[Prof<PROF>::CH], StageFlags &fl,
                                         float *gains_row, float *spec_trace, float *psd_keep = nullptr)
{
    using P = Prof<PROF>;
    float *ring = fstage == 0 ? wm.ring1 + 8 : wm.ring2 + 8;
    const int cnt = fstage == 0 ? s.nb_in1 : s.nb_in2;
    const int slot_old = cnt & 3;
    const int cur_base = ((cnt + 1) & 3) * P::S;

    float psd[P::NB];
    frame_psd<PROF>(wm, bm, tw, lane, ring, slot_old * P::S + P::ANA, psd);
    if (psd_keep) { psd_keep[0] = psd[1]; psd_keep[1] = psd[2]; }

    // VAD (NoiseSup.c:359-430): frameEn = 64 + sum x^2 over the newest shift, accumulated sample by sample
    float frame_en = 64.0f;
    if (fstage == 0)
        for (int i = 0; i < P::S; i++) { const float x = ring[cur_base + i]; frame_en += x * x; }
    vad_ns(s, fstage, frame_en);
    const int nbf = (int)(short)(fstage == 0 ? s.nb_frame0 : s.nb_frame1);

    // PSDMean (:289-303) + FilterCalc (:449-563) per owned bin; W / den / noise also to shared memory for the sums
    float *prev = wm.prev + fstage * (P::NB * 32);
#pragma unroll
    for (int r = 0; r < P::NB; r++) {
        const int b = lane + 32 * r;
        if (b < P::SPEC) {
            const float pm = (prev[r * 32 + lane] + psd[r]) / 2;
            prev[r * 32 + lane] = psd[r];
            const float W = filter_calc_bin(fstage, psd[r], pm, noise[r], den[r], nbf, s.flag_vad, eps);
            wm.W[b] = W;
            wm.sden[b] = den[r];
            wm.snoise[b] = noise[r];
            if (spec_trace) {
                spec_trace[(3 * fstage + 0) * P::SPEC + b] = (float)sqrt((double)psd[r]);
                spec_trace[(3 * fstage + 1) * P::SPEC + b] = noise[r];
                spec_trace[(3 * fstage + 2) * P::SPEC + b] = den[r];
                spec_trace[(6 + fstage) * P::SPEC + b] = (float)sqrt((double)pm);
            }
        }
    }
    __syncwarp();
    if (fstage == 0) {
        // SpeechQVar (:800-839): mean and variance of the first NS_FFT_LENGTH/4 gains, summed in order
        float mean = 0.0f, var = 0.0f;
        for (int i = 0; i < P::QVAR; i++) { const float w = wm.W[i]; mean += w; var += w * w; }
        SE_COLD(cold_fd_load(s, wm.cold));
        fl.var = q_var(s, mean, var, P::QVAR);
        SE_COLD(cold_fd_store(wm.cold, s));
    }

    // DoMelFB (MelProc.c:82-104) / DoGamma (aurora_etsi/MelProc.cpp:119-141): lane j = band j, bins in ascending order
    float mw = 0.0f;
    if (lane < kMel) {
        const int len = bm.bank_len[lane];
        if constexpr (PROF == 8) {
            const int st = bm.bank_start[lane];
            for (int i = 0; i < len; i++) mw += wm.W[st + i] * bm.bankT[i * 32 + lane];
        } else {
            for (int i = 0; i < len; i++) mw += wm.W[i] * bm.bankT[(lane * 32 + (i & 31)) * 4 + (i >> 5)];
        }
    }
    wm.melw[lane] = mw;
    __syncwarp();

    if (fstage == 0) {
        float te = 0.0f;
        for (int i = 0; i < kMel; i++) te += wm.melw[i];
        SE_COLD(cold_fd_load(s, wm.cold));
        SE_COLD(cold_gf_load(s, wm.cold));
        s.fd_spec_val = (float)((double)(te * te) - 3.0);
        fl.spec = q_spec(s);
        s.fd_mel_val1 = (float)((double)((wm.melw[1] + wm.melw[2]) + wm.melw[3]) / 3.0);
        fl.mel = q_mel(s);
        s.den_en0 = s.den_en1;                                   // DoGainFact stage 1, NoiseSup.c:593-599
        s.den_en1 = s.den_en2;
        float de = 0.0f;
        for (int i = 0; i < P::SPEC; i++) de += wm.sden[i];
        s.den_en2 = de;
        SE_COLD(cold_fd_store(wm.cold, s));
        SE_COLD(cold_gf_store(wm.cold, s));
    } else {
        float ne = 0.0f;
        for (int i = 0; i < P::SPEC; i++) ne += wm.snoise[i];
        SE_COLD(cold_gf_load(s, wm.cold));
        gain_fact_stage2<PROF>(s, ne);
        mw = (float)((double)(s.alfa_gf * mw) + (1.0 - (double)s.alfa_gf) * 1.0);
        SE_COLD(cold_gf_store(wm.cold, s));
        if (gains_row && lane < kMel) gains_row[lane] = mw;
        __syncwarp();
        wm.melw[lane] = mw;
        __syncwarp();
    }

    // DoMelIDCT (MelProc.c:357-378), taps 0..8 of 25 (the rest never reaches DoFilterWindowing, NoiseSup.c:660-669)
    float ht = 0.0f;
    if (lane < 9) {
        for (int f = 0; f < kMel; f++) ht += wm.melw[f] * bm.idctT[f * 16 + lane];
        ht = ht * bm.firwin[lane];
    }
    float h[9];
#pragma unroll
    for (int t = 0; t < 9; t++) h[t] = __shfl_sync(0xffffffffu, ht, t);

    // ApplyWF (:324-340): sum over j = -8 .. 8 in ascending order
    const int i0 = lane * P::CH;
    const float *xin = ring + cur_base + i0 - 8;
#pragma unroll
    for (int t = 0; t < P::CH; t++) {
        float acc = 0.0f;
        if (i0 + t < P::S) {
#pragma unroll
            for (int j = -8; j <= 8; j++) acc += h[j < 0 ? -j : j] * xin[t + 8 - j];
        }
        o[t] = acc;
    }

    if (fstage == 0) {
        const int slot = s.nb_in2 & 3;
#pragma unroll
        for (int t = 0; t < P::CH; t++)
            if (i0 + t < P::S) ring_store<PROF>(wm.ring2 + 8, slot, i0 + t, o[t]);
        s.nb_in2++;
    } else {
        s.nb_out2++;
    }
    __syncwarp();
}

// DCOffsetFil (NoiseSup.c:182-198): y[n] = x[n] - x[n-1] + 0.9990234375 * y[n-1] with the product and the second sum in
// double, sample after sample (every lane runs the same recurrence; the values pass through shared memory).
template <int PROF> __device__ __forceinline__ void dc_offset(float (&o)[Prof<PROF>::CH], Scal &s, int lane, float *scratch)
{
    using P = Prof<PROF>;
    const int i0 = lane * P::CH;
#pragma unroll
    for (int t = 0; t < P::CH; t++)
        if (i0 + t < P::S) scratch[i0 + t] = o[t];
    __syncwarp();
    float dc_x = s.dc_x, dc_y = s.dc_y;
    for (int i = 0; i < P::S; i++) {
        const float aux = scratch[i];
        const float y = (float)((double)(aux - dc_x) + 0.9990234375 * (double)dc_y);
        dc_x = aux;
        dc_y = y;
        if (lane == 0) scratch[P::S + i] = y;
    }
    s.dc_x = dc_x;
    s.dc_y = dc_y;
    __syncwarp();
#pragma unroll
    for (int t = 0; t < P::CH; t++)
        if (i0 + t < P::S) o[t] = scratch[P::S + i0 + t];
    __syncwarp();
}

// WI8CompCeps (etsi/cpp/CompCeps.c:368-549).  ld(i) = frame sample i, -1 <= i < 200.  Returns this lane's coefficient
// (lanes 0..11 = c1..c12, lane 12 = c0, lane 13 = logE).  HP16: the 16 kHz hooks (:464-479, 488-505, 518-531;
// DecodeBands16k / MergeSSandCoded / CorrectEnergy, 16kHzProcessing.c:106-158, 318-366).
template <bool HP16 = false, class LD>
__device__ __forceinline__ float comp_ceps_core(WarpMem<8> &wm, const BlockMem<8> &bm, const FftTw &, const BankRegs &, int lane, LD ld,
                                                const float *dct26T = nullptr, float code9 = 0.0f, float hp3 = 0.0f)
{
    constexpr int NBAND = HP16 ? 26 : 23;
    constexpr unsigned FULL = 0xffffffffu;
    const float floor_fb = 4.5399929762484854e-05f;      // (float)exp(-10.0)
    const float floor_e = 1.9287498479639178e-22f;       // (float)exp(-50.0)
    const float corr = 0.6418538689613342f;              // (float)log(1.0 + 0.90): rough pre-emphasis correction (16 kHz mode)
    float *x = wm.xf;
    // log energy (:415-423): sum of squares in order
    float log_e = 0.0f;
    for (int i = 0; i < 200; i++) { const float d = ld(i); log_e += d * d; }
    // pre-emphasis 0.9 (double product, :427-429), Hamming (:434), zero padding
    for (int i = lane; i < 256; i += 32) {
        float v = 0.0f;
        if (i < 200) {
            v = (float)((double)ld(i) - 0.90 * (double)ld(i - 1));
            v = v * bm.ham[i];
        }
        x[__brev((unsigned)i) >> 24] = v;
    }
    __syncwarp();
    rfft_exact(x, bm.rf, lane);
    // power spectrum in double (:451-459) -> wm.W[0..128]
    for (int k = lane; k <= 128; k += 32) {
        float pwr;
        if (k == 0 || k == 128) pwr = (float)((double)x[k] * (double)x[k]);
        else pwr = (float)((double)x[k] * (double)x[k] + (double)x[256 - k] * (double)x[256 - k]);
        wm.W[k] = pwr;
    }
    __syncwarp();
    float coded = 0.0f;
    if constexpr (HP16) {
        // GetBandsForDecoding16k (16kHzProcessing.c:318-336) + DecodeBands16k (:357-366) with the oldest buffered codes
        float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
        for (int i = 66; i < 77; i++) a0 += wm.W[i];
        for (int i = 78; i < 97; i++) a1 += wm.W[i];
        for (int i = 98; i < 129; i++) a2 += wm.W[i];
        a0 = (float)((double)a0 / 2.0); a1 = (float)((double)a1 / 2.0); a2 = (float)((double)a2 / 2.0);
        a0 = (a0 > floor_fb) ? (float)log((double)a0) : -10.0f;
        a1 = (a1 > floor_fb) ? (float)log((double)a1) : -10.0f;
        a2 = (a2 > floor_fb) ? (float)log((double)a2) : -10.0f;
        const int i3 = 3 * (lane - 23);
        const float c0_ = __shfl_sync(FULL, code9, i3 & 31), c1_ = __shfl_sync(FULL, code9, (i3 + 1) & 31),
                    c2_ = __shfl_sync(FULL, code9, (i3 + 2) & 31);
        coded = 0.0f;
        coded += 0.1f * (a0 - c0_);
        coded += 0.2f * (a1 - c1_);
        coded += 0.7f * (a2 - c2_);
    }
    // 23 mel triangles (DoMelFB, MelProc.c:82-104): lane j = band j, bins in ascending order
    float fb = 0.0f;
    if (lane < 23) {
        const int st = bm.tri_start[lane], len = bm.tri_len[lane];
        for (int i = 0; i < len; i++) fb += wm.W[st + i] * bm.triT[i * 32 + lane];
    }
    if constexpr (HP16) {
        const float hp_e = __shfl_sync(FULL, hp3, (lane - 23) & 31);
        if (lane >= 23 && lane < 26) fb = (float)((1.0 + 0.90) * (double)hp_e);       // CompCeps.c:488-505
        coded = coded + corr;
    }
    if (lane < NBAND) fb = (fb < floor_fb) ? -10.0f : (float)log((double)fb);       // :509-513
    else fb = 0.0f;
    if constexpr (HP16) {
        // MergeSSandCoded (16kHzProcessing.c:106-125)
        const float perc = 0.7f;
        if (lane >= 23 && lane < 26) fb = (float)((double)(perc * coded) + (1.0 - (double)perc) * (double)fb);
        const float b22 = __shfl_sync(FULL, fb, 22), b23 = __shfl_sync(FULL, fb, 23);
        const float avg = (float)(0.5 * (double)b22 + 0.5 * (double)b23);
        if (lane == 22 || lane == 23) fb = (float)(0.6 * (double)fb + 0.4 * (double)avg);
        // CorrectEnergy (:145-158): energyHP (float) += exp(double), band after band
        const float ex23 = __shfl_sync(FULL, fb, 23), ex24 = __shfl_sync(FULL, fb, 24), ex25 = __shfl_sync(FULL, fb, 25);
        float e_hp = 0.0f;
        e_hp = (float)((double)e_hp + exp((double)(ex23 - corr)));
        e_hp = (float)((double)e_hp + exp((double)(ex24 - corr)));
        e_hp = (float)((double)e_hp + exp((double)(ex25 - corr)));
        log_e = log_e + e_hp;
    }
    wm.melw[lane] = fb;
    __syncwarp();
    float out = 0.0f;
    if (lane < 12) {                                     // DCT (:203-227)
        const float *dt = HP16 ? dct26T : bm.dctT;
        for (int j = 0; j < NBAND; j++) out += wm.melw[j] * dt[j * 16 + lane];
    } else if (lane == 12) {
        for (int j = 0; j < NBAND; j++) out += wm.melw[j];
    } else if (lane == 13) {
        out = (log_e < floor_e) ? -50.0f : (float)log((double)log_e);
    }
    __syncwarp();
    return out;
}

__device__ __forceinline__ void comp_ceps(WarpMem<8> &wm, const BlockMem<8> &bm, const FftTw &tw, const BankRegs &br, int lane, int base,
                                          float *row)
{
    using P = Prof<8>;
    const float *d = wm.ring3;
    const float out = comp_ceps_core(wm, bm, tw, br, lane, [&](int i) { return d[wrap<P::RB>(base + i)]; });
    if (lane < 14) row[lane] = out;
}

}  // namespace SE_NSPACE
