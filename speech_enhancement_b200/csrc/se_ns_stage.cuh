// One Wiener-filter stage of one frame, executed by one warp.  Included by se_ns.cu.
//
// Follows the body of the fstage loop of DoNoiseSup (etsi/cpp/NoiseSup.c:1144-1367; plugin twin
// aurora_etsi/NoiseSup.cpp:1218-1374): DoSigWindowing -> rfft -> FFTtoPSD -> PSDMean -> VAD ->
// FilterCalc -> SpeechQVar -> DoMelFB|DoGamma -> SpeechQSpec/Mel -> DoGainFact -> DoMelIDCT ->
// DoFilterWindowing -> ApplyWF.
#pragma once

namespace se {

template <int RB> __device__ __forceinline__ int wrap(int pos)
{
    if (pos >= RB) pos -= RB;
    if (pos < 0) pos += RB;
    return pos;
}

// Windowed frame -> PSD bins owned by this lane (psd[r] for bin lane+32r).
template <int PROF>
__device__ __forceinline__ void frame_psd(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const FftTw &tw, int lane,
                                          const float *ring, int wpos, float (&psd)[Prof<PROF>::NB])
{
    using P = Prof<PROF>;
    if constexpr (PROF == 8) {
        float2 v[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int i2 = 2 * (lane + 32 * j);
            v[j] = make_float2(0.0f, 0.0f);
            if (i2 < P::FL) {
                const int pos = wrap<P::RB>(wpos + i2);
                const float2 x = *reinterpret_cast<const float2 *>(ring + pos);
                const float2 w = *reinterpret_cast<const float2 *>(bm.win + i2);
                v[j] = make_float2(x.x * w.x, x.y * w.y);
            }
        }
        fft128_warp(v, tw, lane);
        fft128_store(v, wm.Z, lane);
        __syncwarp();
        // FFTtoPSD, etsi/cpp/NoiseSup.c:249-270: |X|^2 on 129 bins, pair-averaged to 65
#pragma unroll
        for (int r = 0; r < P::NB; r++) {
            const int b = lane + 32 * r;
            psd[r] = 0.0f;
            if (b < 64) {
                const float2 x0 = rsplit256(wm.Z, 2 * b, bm.twh[2 * b]);
                const float2 x1 = rsplit256(wm.Z, 2 * b + 1, bm.twh[2 * b + 1]);
                psd[r] = ((x0.x * x0.x + x0.y * x0.y) + (x1.x * x1.x + x1.y * x1.y)) * 0.5f;
            } else if (b == 64) {
                const float2 x0 = rsplit256(wm.Z, 128, bm.twh[128]);
                psd[r] = x0.x * x0.x;
            }
        }
    } else {
        // The plugin calls rfft(x, 512, 8): one butterfly level short.  What that
        // leaves in the array is A = rfft256(x[0::2]) in [0,256), B = rfft128(x[1::4]) in [256,384) and
        // C = rfft128(x[3::4]) in [384,512), each in the Re0..Re(n/2),Im(n/2-1)..Im1 layout; FFTtoPSD
        // then pairs element i with element 512-i.  Reproduced exactly, not "fixed".
        float2 va[4], vw[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int i4 = 4 * (lane + 32 * j);
            va[j] = make_float2(0.0f, 0.0f);
            vw[j] = make_float2(0.0f, 0.0f);
            if (i4 < P::FL) {
                const int pos = wrap<P::RB>(wpos + i4);
                const float4 x = *reinterpret_cast<const float4 *>(ring + pos);
                const float4 w = *reinterpret_cast<const float4 *>(bm.win + i4);
                va[j] = make_float2(x.x * w.x, x.z * w.z);
                vw[j] = make_float2(x.y * w.y, x.w * w.w);
            }
        }
        fft128_warp(va, tw, lane);
        fft128_warp(vw, tw, lane);
        fft128_store(va, wm.Z, lane);
        fft128_store(vw, wm.Z2, lane);
        __syncwarp();
        float *raw = wm.raw;
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int k = lane + 32 * r;
            const float2 a = rsplit256(wm.Z, k, bm.twh[k]);
            raw[k] = a.x * a.x;
            if (k > 0) raw[256 - k] = a.y * a.y;
        }
        if (lane == 0) {
            const float2 a = rsplit256(wm.Z, 128, bm.twh[128]);
            raw[128] = a.x * a.x;
        }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int k = lane + 32 * r;
            float2 B, C;
            rsplit_pair(wm.Z2, k, B, C);
            if (k == 0) {
                raw[256] = B.x * B.x;
                raw[128] += C.x * C.x;
            } else {
                raw[k] += C.y * C.y;
                raw[128 - k] += C.x * C.x;
                raw[128 + k] += B.y * B.y;
                raw[256 - k] += B.x * B.x;
            }
        }
        if (lane == 0) {
            float2 B, C;
            rsplit_pair(wm.Z2, 64, B, C);
            raw[64] += C.x * C.x;
            raw[192] += B.x * B.x;
        }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < P::NB; r++) {
            const int b = lane + 32 * r;
            psd[r] = 0.0f;
            if (b < 128) psd[r] = (raw[2 * b] + raw[2 * b + 1]) * 0.5f;
            else if (b == 128) psd[r] = raw[256];
        }
    }
}

struct StageFlags { int var, spec, mel; };

template <int PROF, int FSTAGE>
__device__ __forceinline__ void ns_stage(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const FftTw &tw, int lane, Scal &s,
                                         float (&noise)[Prof<PROF>::NB], float (&den)[Prof<PROF>::NB],
                                         float (&prev)[Prof<PROF>::NB], float eps, float (&o)[Prof<PROF>::CH],
                                         StageFlags &fl, float *gains_row, float *spec_trace)
{
    using P = Prof<PROF>;
    float *ring = FSTAGE == 0 ? wm.ring1 : wm.ring2;
    const int cnt = FSTAGE == 0 ? s.nb_in1 : s.nb_in2;   // blocks written to this ring so far
    const int slot_old = cnt & 3;                        // buffer[0..S)      = prvFrame
    const int cur_base = ((cnt + 1) & 3) * P::S;         // buffer[S..2S)     = curFrame

    float psd[P::NB];
    frame_psd<PROF>(wm, bm, tw, lane, ring, slot_old * P::S + P::ANA, psd);

    // VAD (stage 0 only looks at the signal)
    float frame_sq = 0.0f;
    if (FSTAGE == 0) {
#pragma unroll
        for (int r = 0; r < P::NR; r++) {
            const int i = lane + 32 * r;
            if (i < P::S) { const float x = ring[cur_base + i]; frame_sq += x * x; }
        }
        frame_sq = warp_sum(frame_sq);
    }
    vad_ns(s, FSTAGE, frame_sq);
    const int nbf = (int)(short)(FSTAGE == 0 ? s.nb_frame0 : s.nb_frame1);   // Q2: read through int16

    // PSDMean (NoiseSup.c:289-303) + FilterCalc per owned bin
    float wsum = 0.0f, wsq = 0.0f, den_sum = 0.0f, noise_sum = 0.0f;
#pragma unroll
    for (int r = 0; r < P::NB; r++) {
        const int b = lane + 32 * r;
        if (b < P::SPEC) {
            const float pm = (prev[r] + psd[r]) * 0.5f;
            prev[r] = psd[r];
            const float W = filter_calc_bin<FSTAGE>(psd[r], pm, noise[r], den[r], nbf, s.flag_vad, eps);
            wm.W[b] = W;
            if (r < P::NB - 1) { wsum += W; wsq += W * W; }
            den_sum += den[r];
            noise_sum += noise[r];
            if (spec_trace) {
                spec_trace[(3 * FSTAGE + 0) * P::SPEC + b] = sqrtf(psd[r]);
                spec_trace[(3 * FSTAGE + 1) * P::SPEC + b] = noise[r];
                spec_trace[(3 * FSTAGE + 2) * P::SPEC + b] = den[r];
            }
        }
    }
    if (FSTAGE == 0) {
        wsum = warp_sum(wsum);
        wsq = warp_sum(wsq);
        fl.var = q_var(s, wsum, wsq, P::QVAR);
    }
    __syncwarp();

    // DoMelFB (etsi/cpp/MelProc.c:82-104) / DoGamma (aurora_etsi/MelProc.cpp:119-141, ignores StartingPoint)
    float mw = 0.0f;
    if constexpr (PROF == 8) {
        const int st = bm.bank_start[lane], len = bm.bank_len[lane];
#pragma unroll 4
        for (int i = 0; i < P::BANKROWS; i++)
            if (i < len) mw += wm.W[st + i] * bm.bankT[i * 32 + lane];
    } else {
#pragma unroll 4
        for (int i = 0; i < 128; i += 4) {
            const float4 w4 = *reinterpret_cast<const float4 *>(wm.W + i);
            mw += w4.x * bm.bankT[(i + 0) * 32 + lane];
            mw += w4.y * bm.bankT[(i + 1) * 32 + lane];
            mw += w4.z * bm.bankT[(i + 2) * 32 + lane];
            mw += w4.w * bm.bankT[(i + 3) * 32 + lane];
        }
    }
    if (lane >= kMel) mw = 0.0f;

    if (FSTAGE == 0) {
        // frame-dropping VAD statistics, NoiseSup.c:1268-1281
        const float te = warp_sum(mw);
        s.fd_spec_val = (float)((double)(te * te) - 3.0);
        fl.spec = q_spec(s);
        const float w1 = __shfl_sync(0xffffffffu, mw, 1), w2 = __shfl_sync(0xffffffffu, mw, 2),
                    w3 = __shfl_sync(0xffffffffu, mw, 3);
        s.fd_mel_val1 = (float)((double)__fadd_rn(__fadd_rn(w1, w2), w3) / 3.0);
        fl.mel = q_mel(s);
        // DoGainFact stage 0, NoiseSup.c:593-599
        s.den_en0 = s.den_en1;
        s.den_en1 = s.den_en2;
        s.den_en2 = warp_sum(den_sum);
    } else {
        gain_fact_stage2<PROF>(s, warp_sum(noise_sum));
        mw = (float)((double)__fmul_rn(s.alfa_gf, mw) + (1.0 - (double)s.alfa_gf) * 1.0);
        if (gains_row && lane < kMel) gains_row[lane] = mw;
    }

    // DoMelIDCT (MelProc.c:357-378): only taps 0..8 survive DoFilterWindowing (NoiseSup.c:660-669)
    wm.melw[lane] = mw;
    __syncwarp();
    if (lane < 9) {
        float acc = 0.0f;
#pragma unroll
        for (int f = 0; f < kMel; f++) acc += wm.melw[f] * bm.idctT[f * 16 + lane];
        wm.fir[lane] = acc * bm.firwin[lane];
    }
    __syncwarp();
    float h[9];
#pragma unroll
    for (int t = 0; t < 9; t++) h[t] = wm.fir[t];

    // ApplyWF, NoiseSup.c:324-340: 17-tap symmetric FIR over curFrame with prv/next context
    float x[P::CH + 16];
    const int i0 = lane * P::CH;
#pragma unroll
    for (int d = 0; d < P::CH + 16; d++) x[d] = ring[wrap<P::RB>(cur_base + i0 + d - 8)];
#pragma unroll
    for (int t = 0; t < P::CH; t++) {
        float acc = 0.0f;
#pragma unroll
        for (int j = -8; j <= 8; j++) acc += h[j < 0 ? -j : j] * x[t + 8 - j];
        o[t] = acc;
    }

    if (FSTAGE == 0) {
        const int nb = (s.nb_in2 & 3) * P::S;
#pragma unroll
        for (int t = 0; t < P::CH; t++)
            if (i0 + t < P::S) wm.ring2[nb + i0 + t] = o[t];
        s.nb_in2++;
    } else {
        s.nb_out2++;
    }
    __syncwarp();
}

// DCOffsetFil (NoiseSup.c:182-198) over one shift, chunk-per-lane layout, as a warp scan of the
// first-order recurrence y[n] = a y[n-1] + (x[n] - x[n-1]), a = 0.9990234375.
template <int PROF> __device__ __forceinline__ void dc_offset(float (&o)[Prof<PROF>::CH], Scal &s, int lane)
{
    using P = Prof<PROF>;
    constexpr float a = 0.9990234375f;
    float apow[P::CH + 1];
    apow[0] = 1.0f;
#pragma unroll
    for (int t = 1; t <= P::CH; t++) apow[t] = apow[t - 1] * a;
    float xprev = __shfl_up_sync(0xffffffffu, o[P::CH - 1], 1);
    if (lane == 0) xprev = s.dc_x;
    float y[P::CH];
    float acc = 0.0f, xp = xprev;
#pragma unroll
    for (int t = 0; t < P::CH; t++) {
        const bool valid = lane * P::CH + t < P::S;
        const float xt = valid ? o[t] : xp;
        acc = a * acc + (xt - xp);
        xp = xt;
        y[t] = acc;
    }
    // inclusive scan of lane-end values: E(l) = e_l + A E(l-1), A = a^CH, E(-1) = dc_y
    float v = acc;
    if (lane == 0) v += apow[P::CH] * s.dc_y;
    float Ad = apow[P::CH];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float up = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += Ad * up;
        Ad *= Ad;
    }
    float carry = __shfl_up_sync(0xffffffffu, v, 1);
    if (lane == 0) carry = s.dc_y;
#pragma unroll
    for (int t = 0; t < P::CH; t++) y[t] += apow[t + 1] * carry;
    // new filter memory: sample S-1
    constexpr int last_lane = (P::S - 1) / P::CH, last_t = (P::S - 1) % P::CH;
    s.dc_x = __shfl_sync(0xffffffffu, o[last_t], last_lane);
    s.dc_y = __shfl_sync(0xffffffffu, y[last_t], last_lane);
#pragma unroll
    for (int t = 0; t < P::CH; t++) o[t] = y[t];
}

// WI8CompCeps (etsi/cpp/CompCeps.c:368-549) on the 200-sample frame that starts `base` (ring3
// position of frame sample 0); ring3 holds the float denoised stream.
__device__ __forceinline__ void comp_ceps(WarpMem<8> &wm, const BlockMem<8> &bm, const FftTw &tw, int lane, int base,
                                          float *row)
{
    using P = Prof<8>;
    const float *d = wm.ring3;
    float e = 0.0f;
    float2 v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i2 = 2 * (lane + 32 * j);
        v[j] = make_float2(0.0f, 0.0f);
        if (i2 < 200) {
            const float dm1 = d[wrap<P::RB>(base + i2 - 1)];
            const float d0 = d[wrap<P::RB>(base + i2)];
            const float d1 = d[wrap<P::RB>(base + i2 + 1)];
            e += d0 * d0 + d1 * d1;
            // pre-emphasis 0.9 (:427-429) and Hamming (:434)
            v[j] = make_float2((d0 - 0.9f * dm1) * bm.ham[i2], (d1 - 0.9f * d0) * bm.ham[i2 + 1]);
        }
    }
    e = warp_sum(e);
    fft128_warp(v, tw, lane);
    fft128_store(v, wm.Z, lane);
    __syncwarp();
    // power spectrum (:451-459), 129 bins, into wm.W
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int k = lane + 32 * r;
        const float2 xk = rsplit256(wm.Z, k, bm.twh[k]);
        wm.W[k] = (k == 0) ? xk.x * xk.x : (xk.x * xk.x + xk.y * xk.y);
    }
    if (lane == 0) {
        const float2 xk = rsplit256(wm.Z, 128, bm.twh[128]);
        wm.W[128] = xk.x * xk.x;
    }
    __syncwarp();
    // 23 mel triangles (DoMelFB, MelProc.c:82-104), ln with floor (:509-513)
    float fb = 0.0f;
    {
        const int st = bm.tri_start[lane], len = bm.tri_len[lane];
        for (int i = 0; i < kTriRows; i++)
            if (i < len) fb += wm.W[st + i] * bm.triT[i * 32 + lane];
    }
    const float floor_fb = 4.5399929762484854e-05f;   // (float)exp(-10.0)
    fb = (fb < floor_fb) ? -10.0f : logf(fb);
    if (lane >= 23) fb = 0.0f;
    const float c0 = warp_sum(fb);
    wm.melw[lane] = fb;
    __syncwarp();
    float out = 0.0f;
    if (lane < 12) {
#pragma unroll
        for (int j = 0; j < 23; j++) out += wm.melw[j] * bm.dctT[j * 16 + lane];
    } else if (lane == 12) {
        out = c0;
    } else if (lane == 13) {
        const float floor_e = 1.9287498479639178e-22f;   // (float)exp(-50.0)
        out = (e < floor_e) ? -50.0f : logf(e);
    }
    if (lane < 14) row[lane] = out;
    __syncwarp();
}

}  // namespace se
