// ETSI 16 kHz mode behind the P8 stream kernel (FEAT == 2): QMF split of the 16 kHz input, spectral subtraction and
// coding of three 4-8 kHz mel bands next to NoiseSup, decoding + merging inside CompCeps (26 log bands).
//   Do16kProcessing   etsi/cpp/16kHzProcessing.c:711-787   (qmf_split_warp)
//   DoNoiseSup hooks  etsi/cpp/NoiseSup.c:1119-1120, 1166-1172, 1216-1235 (hp16_lowband_log),
//                     :1307-1327 (hp16_stage1: GetBandsForCoding16k, CodeBands16k, DoSpecSub16k 16kHzProcessing.c:388-480),
//                     :1402-1435 (hp16_push)
//   WI8CompCeps hooks etsi/cpp/CompCeps.c:464-479, 488-505, 518-531 (comp_ceps16_core; DecodeBands16k,
//                     MergeSSandCoded, CorrectEnergy 16kHzProcessing.c:106-158, 357-366)
// One warp per utterance, included by se_ns.cu.
#pragma once
#include "se_qmf_taps.h"

namespace SE_NSPACE {

__constant__ float c_qmfz[126] = SE_QMF_PADDED;
#undef Q

constexpr int kHpRing = 480;     // bufferData16k: 6 shifts of the flipped 4-8 kHz half
struct __align__(16) Hp16Mem {
    float line[280];             // [0,117) = QMF delay line, [117,277) = this call's 160 samples
    float hpring[kHpRing];
};
struct __align__(16) Hp16Block {
    float hpfb[3 * 64];          // mel bands 1..3 of the 5-band HP filter bank as dense rows over PSD bins 0..63
    float dct26T[26 * 16];       // dct26T[j*16 + t] = DCT row t (c_{t+1}) over 26 log bands
};

// Lane-distributed state: lane i holds element i of each small delay line.
struct Hp16State {
    float lpl;        // BandsForCoding16k[9]: 3 frames x 3 low-band log energies (oldest first)
    float codeb;      // bufferCodeForBands16k[27]
    float hpb;        // hpBands[9]
    float noise;      // noiseSE16k[3]
    float mean_en;    // warp-uniform from here on
    int nb_speech, hang_over;
};

__device__ __forceinline__ void hp16_state_init(Hp16State &h, Hp16Mem &hm, int lane)
{
    h.lpl = h.codeb = h.hpb = h.noise = h.mean_en = 0.0f;
    h.nb_speech = h.hang_over = 0;
    for (int i = lane; i < 280; i += 32) hm.line[i] = 0.0f;
    for (int i = lane; i < kHpRing; i += 32) hm.hpring[i] = 0.0f;
}

// Two-channel analysis bank, decimated by 2.  Both filters share the symmetric prototype p: the high-pass is
// h[j] = -(-1)^j p[j], so with E = sum over even taps and O = sum over odd taps the low band is E + O and the high
// band O - E, flipped to base band by the alternating output sign.  Lane l < 27 produces outputs 3l..3l+2 from
// one pass over x[6l .. 6l+121] (61 LDS.64); taps come from the constant bank as FFMA operands.
// The reference accumulates j = 0..117 in one float chain; this order differs by float rounding only.
__device__ __forceinline__ void qmf_split_warp(Hp16Mem &hm, float *ring1, int slot, int hslot, int lane)
{
#ifdef SE_NS_EXACT
    // exact mode: both 118-tap sums of every output in the reference's order j = 0..117 (16kHzProcessing.c:735-760);
    // the high-pass taps are the low-pass prototype with alternating signs (+ for odd j), the output sign alternates
#pragma unroll 1
    for (int k = lane; k < 80; k += 32) {
        float a = 0.0f, b = 0.0f;
        const float *x = hm.line + 2 * k;
        for (int j = 0; j < 118; j++) {
            const float tap = c_qmfz[4 + j];
            a += x[j] * tap;
            b += x[j] * ((j & 1) ? tap : -tap);
        }
        ring_store<8>(ring1, slot, k, a);
        hm.hpring[hslot * 80 + k] = (k & 1) ? -b : b;
    }
#else
    if (lane < 27) {
        float e0 = 0.0f, o0 = 0.0f, e1 = 0.0f, o1 = 0.0f, e2 = 0.0f, o2 = 0.0f;
        const float2 *xp = reinterpret_cast<const float2 *>(hm.line + 6 * lane);
#pragma unroll
        for (int p = 0; p < 61; p++) {
            const float2 x = xp[p];
            e0 = fmaf(x.x, c_qmfz[4 + 2 * p], e0);
            e1 = fmaf(x.x, c_qmfz[2 + 2 * p], e1);
            e2 = fmaf(x.x, c_qmfz[2 * p], e2);
            o0 = fmaf(x.y, c_qmfz[5 + 2 * p], o0);
            o1 = fmaf(x.y, c_qmfz[3 + 2 * p], o1);
            o2 = fmaf(x.y, c_qmfz[1 + 2 * p], o2);
        }
        const float lp[3] = {e0 + o0, e1 + o1, e2 + o2};
        const float hp[3] = {o0 - e0, o1 - e1, o2 - e2};
#pragma unroll
        for (int t = 0; t < 3; t++) {
            const int k = 3 * lane + t;
            if (k < 80) {
                ring_store<8>(ring1, slot, k, lp[t]);
                hm.hpring[hslot * 80 + k] = (k & 1) ? -hp[t] : hp[t];
            }
        }
    }
#endif
    __syncwarp();
    // keep the last 117 samples as the next call's delay line
    float keep[4];
#pragma unroll
    for (int r = 0; r < 4; r++) keep[r] = hm.line[160 + min(lane + 32 * r, 116)];
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 4; r++)
        if (lane + 32 * r < 117) hm.line[lane + 32 * r] = keep[r];
    __syncwarp();
}

// Stage-0 hook: log energies of PSD bins 33..38, 39..48, 49..64 pushed into the 3-frame history.
// psd1 / psd2 = this lane's bins lane+32 and lane+64.
__device__ __forceinline__ void hp16_lowband_log(Hp16State &h, float psd1, float psd2, int lane)
{
    const float floor_v = 4.5399929762484854e-05f;   // NS_SPEC_FLOOR = (float)exp(-10.0)
#ifdef SE_NS_EXACT
    // exact mode: GetBandsForCoding16k's three sums bin after bin (16kHzProcessing.c:497-511), double log
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f;
    for (int i = 33; i < 39; i++) a0 += __shfl_sync(0xffffffffu, psd1, i - 32);
    for (int i = 39; i < 49; i++) a1 += __shfl_sync(0xffffffffu, psd1, i - 32);
    for (int i = 49; i < 64; i++) a2 += __shfl_sync(0xffffffffu, psd1, i - 32);
    a2 += __shfl_sync(0xffffffffu, psd2, 0);
    const float sel = (lane == 6) ? a0 : ((lane == 7) ? a1 : a2);
    const float lg = (sel > floor_v) ? (float)log((double)sel) : -10.0f;
#else
    float a0 = (lane >= 1 && lane <= 6) ? psd1 : 0.0f;
    float a1 = (lane >= 7 && lane <= 16) ? psd1 : 0.0f;
    float a2 = (lane >= 17) ? psd1 : ((lane == 0) ? psd2 : 0.0f);
    a0 = warp_sum(a0);
    a1 = warp_sum(a1);
    a2 = warp_sum(a2);
    const float sel = (lane == 6) ? a0 : ((lane == 7) ? a1 : a2);
    const float lg = (sel > floor_v) ? logf(sel) : -10.0f;
#endif
    const float down = __shfl_down_sync(0xffffffffu, h.lpl, 3);
    h.lpl = (lane < 6) ? down : ((lane < 9) ? lg : 0.0f);
}

// Stage-1 hook.  hslot = ring slot written by this call; the frame that lines up with the cepstral frame starts 4
// shifts earlier (bufferData16k[80..279]).  frame_no = vadNS.nbFrame[1].  Returns, lane-distributed:
//   code (lanes 0..8)  = oldest low-band log energies - log HP band   (CodeForBands16k)
//   fbss (lanes 0..2)  = spectrally-subtracted HP band energies        (fb16k after DoSpecSub16k)
__device__ __forceinline__ void hp16_stage1(WarpMem<8> &wm, const BlockMem<8> &bm, const Hp16Block &hb, const Hp16Mem &hm,
                                            const FftTw &tw, int lane, Hp16State &h, int hslot, int frame_no,
                                            float &code, float &fbss)
{
    constexpr unsigned FULL = 0xffffffffu;
    float psd[3];
    int wslot = hslot + 2;
    if (wslot >= 6) wslot -= 6;
    frame_psd<8, kHpRing>(wm, bm, tw, lane, hm.hpring, wslot * 80, psd);
#ifdef SE_NS_EXACT
    // exact mode: DoMelFB over the HP bank bin after bin (MelProc.c:82-104; zero weights outside a triangle add
    // exact zeros), double logs below
#define SE_HP_LOG(v) ((float)log((double)(v)))
    wm.W[lane] = psd[0];
    wm.W[32 + lane] = psd[1];
    __syncwarp();
    float f0 = 0.0f, f1 = 0.0f, f2 = 0.0f;
    for (int i = 0; i < 64; i++) {
        const float pv = wm.W[i];
        f0 += pv * hb.hpfb[i];
        f1 += pv * hb.hpfb[64 + i];
        f2 += pv * hb.hpfb[128 + i];
    }
    __syncwarp();
#else
#define SE_HP_LOG(v) logf(v)
    float f0 = psd[0] * hb.hpfb[lane] + psd[1] * hb.hpfb[32 + lane];
    float f1 = psd[0] * hb.hpfb[64 + lane] + psd[1] * hb.hpfb[96 + lane];
    float f2 = psd[0] * hb.hpfb[128 + lane] + psd[1] * hb.hpfb[160 + lane];
    f0 = warp_sum(f0);
    f1 = warp_sum(f1);
    f2 = warp_sum(f2);
#endif
    const float floor_v = 4.5399929762484854e-05f;
    // CodeBands16k: code[3 i + j] = lp_log[j] - aux[i]
    {
        const int i = lane / 3, j = lane - 3 * i;
        const float fi = (i == 0) ? f0 : ((i == 1) ? f1 : f2);
        const float aux = (fi > floor_v) ? SE_HP_LOG(fi) : -10.0f;
        code = __shfl_sync(FULL, h.lpl, j) - aux;
    }
    // DoSpecSub16k (warp-uniform scalars; lane i < 3 owns band i's noise estimate)
    const int nbf = frame_no;
    const float lambda = (nbf < 100) ? (float)(1.0 - 1.0 / (double)(float)nbf) : 0.99f;
    float frame_en = __fadd_rn(__fadd_rn(f0, f1), f2);
    frame_en = ((double)frame_en > 0.001) ? SE_HP_LOG(frame_en) : -6.9077552789821368f;      // (float)log(0.001)
    float mean_en = h.mean_en;
    if (((double)__fsub_rn(frame_en, mean_en) < 1.2) || (nbf < 10)) {
        const float d = __fsub_rn(frame_en, mean_en);
        if (nbf < 10) mean_en = __fadd_rn(mean_en, __fmul_rn(1.0f - lambda, d));
        else if (frame_en < mean_en) mean_en = (float)((double)mean_en + (1 - 0.98) * (double)d);
        else mean_en = (float)((double)mean_en + (1 - 0.995) * (double)d);
    }
    int flag;
    if ((double)__fsub_rn(frame_en, mean_en) > 2.2) {
        flag = 1;
        h.nb_speech++;
    } else {
        if (h.nb_speech > 4) h.hang_over = 15;
        h.nb_speech = 0;
        if (h.hang_over != 0) { h.hang_over--; flag = 1; }
        else flag = 0;
    }
    h.mean_en = mean_en;
    const float fb = (lane == 0) ? f0 : ((lane == 1) ? f1 : f2);
    if (flag == 0) {
        float nz = h.noise;
        if (nbf < 10 || fb < nz) nz = __fadd_rn(__fmul_rn(lambda, nz), __fmul_rn(1.0f - lambda, fb));
        else nz = (float)(0.995 * (double)nz + (1 - 0.995) * (double)fb);
        if (nz < floor_v) nz = floor_v;            // NS_EPS_16K
        h.noise = nz;
    }
    const float fl = (float)(0.1 * (double)fb);
    const float diff = (float)((double)fb - 1.5 * (double)h.noise);
    fbss = (diff > fl) ? diff : fl;
}

#undef SE_HP_LOG

// After a denoised shift is out: advance the three delay lines.
__device__ __forceinline__ void hp16_push(Hp16State &h, float code, float fbss, int lane)
{
    constexpr unsigned FULL = 0xffffffffu;
    const float hb_down = __shfl_down_sync(FULL, h.hpb, 3);
    const float hb_new = __shfl_sync(FULL, fbss, (lane - 6) & 31);
    h.hpb = (lane < 6) ? hb_down : ((lane < 9) ? hb_new : 0.0f);
    const float cb_down = __shfl_down_sync(FULL, h.codeb, 9);
    const float cb_new = __shfl_sync(FULL, code, (lane - 18) & 31);
    h.codeb = (lane < 18) ? cb_down : ((lane < 27) ? cb_new : 0.0f);
}

}  // namespace SE_NSPACE
