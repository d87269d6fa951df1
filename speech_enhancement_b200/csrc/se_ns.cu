// Fused two-stage Wiener-filter noise suppression (+ CompCeps), one warp per utterance stream.
//
// Replaces, for a whole ragged batch in ONE launch:
//   P8  : etsi_denoise -> DoAdvProcess -> DoNoiseSup (+ DoCompCeps)
//         (etsi/cpp/AdvFrontEnd.c:125, ParmInterface.c:208, NoiseSup.c:1061, CompCeps.c:368)
//   P16 : etsi_denoise_mapping_func_Wiener (function/.../aurora_etsi/NoiseSup.cpp:1140)
//
// Mapping.  The per-frame recursion (noise PSD, decision-directed gain, VAD, gain factorisation) is
// strictly serial in time, so parallelism comes from utterances: a persistent grid of warps pulls
// utterance indices from an atomic queue.  Inside a warp: lane l owns PSD bins {l, l+32, ...} in
// registers (noise / denoised / previous-PSD state for both stages), the FFT-128 runs as three
// in-register radix-4 passes with two shared-memory transposes (se_warp_fft.cuh), the 4-block sample
// history of both stages lives in per-warp shared memory rings, filter banks / IDCT / windows are
// CTA-shared tables in shared memory, int16 input arrives through bulk-async copies (se_bulk.cuh).
// HBM traffic is exactly the algorithmic minimum: each input sample is read once, each output written
// once.  Batches too small for that mapping get one CTA per utterance instead (se_ns_pipe.cuh).
#include <type_traits>
// Per-lane loop-invariant constants (FFT twiddles, IDCT taps, mel / CompCeps triangle taps) live in shared memory
// ([field][lane], conflict-free) and are reloaded at their use sites instead of holding ~25-45 registers through the
// frame loop.  -DSE_NS_CONST_REGS restores the register version.
#ifndef SE_NS_CONST_REGS
#define SE_NS_CONST_SMEM 1
#endif
#include "se_internal.h"
// This file is compiled twice: as is (namespace se, the default fast arithmetic) and with -DSE_NS_EXACT=1
// -DSE_NSPACE=se_exact -fmad=false (the reference's arithmetic, se_ns_exact.cuh); the two sets of kernels live in
// different namespaces so both can be linked into libse_b200.so.
#ifndef SE_NSPACE
#define SE_NSPACE se
#endif
#ifdef SE_NS_EXACT
#define SE_NS_FAST_MATH 0
namespace se_exact { using namespace se; }
#endif
#include "se_warp_fft.cuh"
#include "se_tmem.cuh"

namespace SE_NSPACE {

// ------------------------------------------------------------------------------------------------
// profiles
// ------------------------------------------------------------------------------------------------
template <int PROF> struct Prof;
template <> struct Prof<8> {      // etsi/cpp/NoiseSup.h:28-79
    static constexpr int S = 80, FL = 200, SPEC = 65, ANA = 60, RB = 320;
    static constexpr int NB = 3;        // PSD bins per lane (ceil(65/32))
    static constexpr int CH = 3;        // contiguous output samples per lane (ceil(80/32))
    static constexpr int NR = 3;        // strided input samples per lane
    static constexpr int BANKROWS = 16; // max mel triangle length (checked on the host)
    static constexpr int QVAR = 64;     // NS_FFT_LENGTH/4
    static constexpr bool CEPS = true;
};
template <> struct Prof<16> {     // function/.../aurora_etsi/NoiseSup.h:22-72
    static constexpr int S = 160, FL = 480, SPEC = 129, ANA = 80, RB = 640;
    static constexpr int NB = 5, CH = 5, NR = 5;
    static constexpr int BANKROWS = 128;
    static constexpr int QVAR = 128;
    static constexpr bool CEPS = false;
};

constexpr int kTriRows = 32;     // max CompCeps triangle length (checked on the host)

// Bulk-async input prefetch (se_ns_kernel.cuh): groups of int16 blocks in flight per warp and the bytes one of them
// occupies (the group plus 16: the copy starts at the 16-byte boundary below the group's first sample).
constexpr int kInSlots = 2;
constexpr int kInGroupBytes = 640;                      // input bytes per copy: 4 blocks of 80 samples, 2 of 160
constexpr int kInSlotBytes = kInGroupBytes + 16;

template <int PROF> struct __align__(16) WarpMem {
    using P = Prof<PROF>;
#ifdef SE_NS_EXACT
    static constexpr bool kSpecScratch = true;       // W / melw: in-order sums of the exact mode, both profiles
#else
    static constexpr bool kSpecScratch = PROF == 8;  // the 16 kHz plugin profile keeps W and the mel gains in registers
#endif
    float ring1[P::RB + 16];          // 8 guard samples either side, see ring_store()
    float ring2[P::RB + 16];
    float ring3[P::CEPS ? P::RB : 4];
    float prev[2 * P::NB * 32];       // previous-frame PSD of both stages, [stage][r][lane]
    float2 Z[kZ];                     // FFT-128 spectrum (zpad layout, 132 used) + its transposes (152)
    float2 Z2[PROF == 16 ? kZ : 2];
    float raw[PROF == 16 ? 260 : 4];
    float W[kSpecScratch ? 132 : 4];
    float melw[kSpecScratch ? 32 : 4];
    float cold[PROF == 8 ? 16 : 4];   // P8: Scal fields touched once per block, parked here between their uses
#ifndef SE_NS_NO_BULK
    __align__(16) unsigned char in_stage[kInSlots][kInSlotBytes];   // cp.async.bulk destinations (int16 input blocks)
    __align__(8) unsigned long long in_bar[kInSlots];               // their mbarriers
#endif
#ifdef SE_NS_EXACT
    float xf[PROF == 8 ? 256 : 512];  // the reference's in-place rfft array
    float sden[132], snoise[132];     // denSigSE / noiseSE of the current stage, for the in-order energy sums
#endif
};

#ifdef SE_NS_EXACT
struct RfPlanDev { const unsigned *ops; const int *lvl; const float *tw; int m; };
#endif

template <int PROF> struct __align__(16) BlockMem {
    using P = Prof<PROF>;
    float win[P::FL];
    float2 twh[132];
    // the 8 kHz profile's sparse banks are read once, from global memory, into per-lane taps (lane_br): only the exact
    // build walks them in shared memory
#ifdef SE_NS_EXACT
    float bankT[PROF == 16 ? 25 * 32 * 4 : P::BANKROWS * 32];
#elif defined(SE_NS_BANK_TMEM)
    float bankT[4];                                              // P16: the Gamma bank lives in tensor memory (se_tmem.cuh)
#else
    float bankT[PROF == 16 ? 25 * 32 * 4 : 4];
#endif
    uint32_t tmem_base;               // P16 with SE_NS_BANK_TMEM: this CTA's tensor-memory allocation (kBankTmemCols columns)
    uint32_t tmem_pad[3];
    int bank_start[32], bank_len[32];
#ifdef SE_NS_EXACT
    float idctT[25 * 16];             // the fast build reads the IDCT / window taps once, from global memory (lane_ic)
    float firwin[12];
#endif
    // CompCeps tables (P8 only)
    float ham[P::CEPS ? 200 : 4];
#ifdef SE_NS_EXACT
    float triT[P::CEPS ? kTriRows * 32 : 4];
#else
    float triT[4];
#endif
    int tri_start[32], tri_len[32];
    float dctT[P::CEPS ? 23 * 16 : 4];
#ifdef SE_NS_CONST_SMEM
    float lane_tw[kTwFloats * 32];   // FftTw of every lane, [field][lane]
    float lane_ic[10 * 32];      // IdctRegs of every lane
    float lane_br[PROF == 8 ? 22 * 32 : 4];   // BankRegs taps of every lane: 10 mel + 12 CompCeps triangle taps
#endif
#ifdef SE_NS_EXACT
    RfPlanDev rf;
#endif
};

// Scalar (warp-uniform) part of the stream state: the GPU image of VAD_DATA_NS, VAD_DATA_FD,
// GAIN_FACT, DC_FILTER and the BUFFERS counters (etsi/cpp/NoiseSup.c:49-100).
struct Scal {
    int nb_in1, nb_in2, nb_out2, nb_frame0, nb_frame1;
    int flag_vad, hang_over, nb_speech, onset, n_ceps, n_rows;
    float mean_en, dc_x, dc_y, den_en0, den_en1, den_en2, low_snr, alfa_gf;
    float fd_mel_mean, fd_var_mean, fd_acc_test, fd_spec_mean, fd_mel_val0, fd_mel_val1, fd_spec_val, fd_in_vadq;
};

// 8 kHz profile: the 13 scalars that are read and written once per block live in per-warp shared memory between their
// uses, so they do not hold registers through the FFT / filter-bank code (every lane writes and reads the same
// values).  Measured: feature chain 36.3 -> 34.5 ms, P8 24.24 -> 24.13 ms; the 16 kHz plugin profile loses 2.5 % with it
// (28.6 vs 27.9 ms), so SE_COLD is a no-op there.
__device__ __forceinline__ void cold_fd_load(Scal &s, const float *c)
{
    s.fd_mel_mean = c[0]; s.fd_var_mean = c[1]; s.fd_acc_test = c[2]; s.fd_spec_mean = c[3];
    s.fd_mel_val0 = c[4]; s.fd_mel_val1 = c[5]; s.fd_spec_val = c[6]; s.fd_in_vadq = c[7];
}
__device__ __forceinline__ void cold_fd_store(float *c, const Scal &s)
{
    c[0] = s.fd_mel_mean; c[1] = s.fd_var_mean; c[2] = s.fd_acc_test; c[3] = s.fd_spec_mean;
    c[4] = s.fd_mel_val0; c[5] = s.fd_mel_val1; c[6] = s.fd_spec_val; c[7] = s.fd_in_vadq;
}
__device__ __forceinline__ void cold_gf_load(Scal &s, const float *c)
{
    s.den_en0 = c[8]; s.den_en1 = c[9]; s.den_en2 = c[10]; s.low_snr = c[11]; s.alfa_gf = c[12];
}
__device__ __forceinline__ void cold_gf_store(float *c, const Scal &s)
{
    c[8] = s.den_en0; c[9] = s.den_en1; c[10] = s.den_en2; c[11] = s.low_snr; c[12] = s.alfa_gf;
}
#define SE_COLD(x) do { if constexpr (PROF == 8) { x; } } while (0)

__device__ __forceinline__ void scal_init(Scal &s)
{
    // DoNoiseSupInit, etsi/cpp/NoiseSup.c:895-943
    s.nb_in1 = s.nb_in2 = s.nb_out2 = s.nb_frame0 = s.nb_frame1 = 0;
    s.flag_vad = s.hang_over = s.nb_speech = s.onset = s.n_ceps = s.n_rows = 0;
    s.mean_en = s.dc_x = s.dc_y = s.den_en0 = s.den_en1 = s.den_en2 = s.low_snr = 0.0f;
    s.alfa_gf = 0.8f;
    s.fd_mel_mean = s.fd_var_mean = s.fd_acc_test = s.fd_spec_mean = 0.0f;
    s.fd_mel_val0 = s.fd_mel_val1 = s.fd_spec_val = s.fd_in_vadq = 0.0f;
}

// ------------------------------------------------------------------------------------------------
// warp-uniform scalar recursions (every lane evaluates the same values; double where the reference
// promotes to double)
// ------------------------------------------------------------------------------------------------

// VAD for noise suppression, etsi/cpp/NoiseSup.c:359-430.  frame_sq = sum of the newest-shift squares.
__device__ __forceinline__ void vad_ns(Scal &s, int fstage, float frame_sq)
{
    int nbf = fstage ? s.nb_frame1 : s.nb_frame0;
    if (nbf < 2147483647) nbf++;
    if (fstage) { s.nb_frame1 = nbf; return; }
    s.nb_frame0 = nbf;
    float lambda = (nbf < 10) ? (1.0f - 1.0f / (float)nbf) : 0.97f;
#ifdef SE_NS_EXACT
    // exact mode: frame_sq already is 64 + the in-order sum; NoiseSup.c:391 in double, log(2.0) as gcc folds it
    float frame_en = (float)(0.5 + (log((double)frame_sq / 64.0) / 0.69314718055994530942) * 16.0);
#else
    float frame_en = 64.0f + frame_sq;
    // NoiseSup.c:391: 0.5 + log(frame_en/64)/log(2) * 16 (double in the reference; float log2 is within
    // an ulp of the float result, the same order as the float sum of squares feeding it)
    frame_en = 0.5f + log2f(frame_en * (1.0f / 64.0f)) * 16.0f;
#endif
    float mean_en = s.mean_en;
    if (((frame_en - mean_en) < 20.0f) || (nbf < 10)) {
        // separately rounded mul/add (no FMA contraction): this recursion feeds the VAD decision
        if ((frame_en < mean_en) || (nbf < 10))
            mean_en = __fadd_rn(mean_en, __fmul_rn(1.0f - lambda, __fsub_rn(frame_en, mean_en)));
        else
            mean_en = __fadd_rn(mean_en, __fmul_rn(1.0f - 0.99f, __fsub_rn(frame_en, mean_en)));
        if (mean_en < 80.0f) mean_en = 80.0f;
    }
    if (nbf > 4) {
        if ((frame_en - mean_en) > 15.0f) {
            s.flag_vad = 1;
            s.nb_speech++;
        } else {
            if (s.nb_speech > 4) s.hang_over = 15;
            s.nb_speech = 0;
            if (s.hang_over != 0) { s.hang_over--; s.flag_vad = 1; }
            else s.flag_vad = 0;
        }
    }
    s.mean_en = mean_en;
}

// SpeechQVar / SpeechQSpec / SpeechQMel, etsi/cpp/NoiseSup.c:686-839
__device__ __forceinline__ int q_var(Scal &s, float mean, float var, int ssize)
{
    const short fc = (short)s.nb_frame0;
    const float sv = __fsub_rn(var / (float)ssize, __fmul_rn(mean, mean) / (float)(ssize * ssize));
    if (fc < 15) s.fd_var_mean = (s.fd_var_mean > sv) ? s.fd_var_mean : sv;
    if ((double)sv < (double)s.fd_var_mean * 1.5 && (double)sv > (double)s.fd_var_mean * 0.85)
        s.fd_var_mean = (float)((double)s.fd_var_mean * 0.8 + (double)sv * 0.2);
    if ((double)sv <= (double)s.fd_var_mean * 0.25)
        s.fd_var_mean = (float)((double)s.fd_var_mean * 0.97 + (double)sv * 0.03);
    return ((double)sv > (double)s.fd_var_mean * 1.65) ? 1 : 0;
}

__device__ __forceinline__ int q_spec(Scal &s)
{
    const short fc = (short)s.nb_frame0;
    const float v = s.fd_spec_val;
    if (fc == 1) s.fd_spec_mean = v;
    if (fc < 15) {
        s.fd_acc_test = (float)(1.1 * (double)__fadd_rn(__fmul_rn(s.fd_acc_test, (float)(fc - 1)), v) / (double)(float)(fc));
        const float acc = v / s.fd_acc_test;
        if ((double)acc > 2.5) s.fd_in_vadq = 1.0f;
        if (s.fd_in_vadq == 0.0f) s.fd_spec_mean = (s.fd_spec_mean > v) ? s.fd_spec_mean : v;
    }
    if ((double)v < (double)s.fd_spec_mean * 1.5 && (double)v > (double)s.fd_spec_mean * 0.75)
        s.fd_spec_mean = (float)((double)s.fd_spec_mean * 0.8 + (double)v * 0.2);
    if ((double)v <= (double)s.fd_spec_mean * 0.5)
        s.fd_spec_mean = (float)((double)s.fd_spec_mean * 0.97 + (double)v * 0.03);
    return ((double)v > (double)s.fd_spec_mean * 1.65) ? 1 : 0;
}

__device__ __forceinline__ int q_mel(Scal &s)
{
    const short fc = (short)s.nb_frame0;
    const float sm = (float)(0.75 * (double)s.fd_mel_val1 + 0.25 * (double)s.fd_mel_val0);
    s.fd_mel_val0 = s.fd_mel_val1;
    if (fc < 15) s.fd_mel_mean = (s.fd_mel_mean > sm) ? s.fd_mel_mean : sm;
    if ((double)sm < (double)s.fd_mel_mean * 1.5 && (double)sm > (double)s.fd_mel_mean * 0.75)
        s.fd_mel_mean = (float)((double)s.fd_mel_mean * 0.8 + (double)sm * 0.2);
    if ((double)sm <= (double)s.fd_mel_mean * 0.5)
        s.fd_mel_mean = (float)((double)s.fd_mel_mean * 0.97 + (double)sm * 0.03);
    return ((double)sm > (double)s.fd_mel_mean * 3.25) ? 1 : 0;
}

// Stage-2 part of DoGainFact, etsi/cpp/NoiseSup.c:600-641 (DoGainFact_IBM, NoiseSup.cpp:653-697).
// The plugin is C++, where log10(float) is the float overload; the ETSI file is C (double log10).
template <int PROF> __device__ __forceinline__ void gain_fact_stage2(Scal &s, float noise_en)
{
    float aver = (s.den_en0 * s.den_en1 * s.den_en2) / (noise_en * noise_en * noise_en);
    if ((double)aver > 0.00001) {
#ifdef SE_NS_EXACT
        // the ETSI file is C: (20 * log10((double)aver)) / 3.0; the plugin is C++: log10 of a float is the float overload
        // (evaluated here as the correctly rounded float of the double log10), times 20 in float, / 3.0 in double
        if (PROF == 8) aver = (float)((20 * log10((double)aver)) / 3.0);
        else aver = (float)((double)(20 * (float)log10((double)aver)) / 3.0);
#else
        aver = (20.0f * log10f(aver)) / 3.0f;
#endif
    } else {
        aver = (float)(-100.0 / 3.0);
    }
    if ((((double)aver - (double)s.low_snr) < 10.0) || (s.nb_frame1 < 10)) {
        float lambda;
        if (s.nb_frame1 < 10) lambda = (float)(1.0 - 1.0 / (double)(float)s.nb_frame1);
        else if (aver < s.low_snr) lambda = (float)0.95;
        else lambda = (float)0.99;
        s.low_snr = (float)((double)s.low_snr + (1.0 - (double)lambda) * (double)(aver - s.low_snr));
    }
    if (s.den_en2 > 100.0f) {
        if ((double)aver < ((double)s.low_snr + 3.5)) {
            s.alfa_gf = (float)((double)s.alfa_gf + 0.15);
            if ((double)s.alfa_gf > 0.8) s.alfa_gf = (float)0.8;
        } else {
            s.alfa_gf = (float)((double)s.alfa_gf - 0.3);
            if ((double)s.alfa_gf < 0.1) s.alfa_gf = (float)0.1;
        }
    }
}

// Approximate-rounding division / square root (MUFU.RCP / MUFU.SQRT, <= 2 ulp) for the per-bin Wiener
// gain.  The gain recursion is smooth and contractive (0.98 decision-directed feedback), so ulp-level
// differences do not grow; parity bars (gains rel 1e-4, waveform >= 60 dB SNR) are checked by the tests.
#ifndef SE_NS_FAST_MATH
#define SE_NS_FAST_MATH 1
#endif
__device__ __forceinline__ float fdiv(float a, float b)
{
#if SE_NS_FAST_MATH
    float r;                       // MUFU.RCP + FMUL; every divisor here is >= 2e-9 (eps^2) and finite
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
    return a * r;
#else
    return a / b;
#endif
}
__device__ __forceinline__ float fsqrt(float a)
{
#if SE_NS_FAST_MATH
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
#else
    return sqrtf(a);
#endif
}

// Decision-directed Wiener gain of one bin, etsi/cpp/NoiseSup.c:449-563 (FilterCalc).
// psd: this frame's PSD bin, psd_mean: 2-frame mean.  Returns W; updates noise/den.
#ifdef SE_NS_EXACT
__device__ __forceinline__ float filter_calc_bin(const int fstage, float psd, float psd_mean, float &noise, float &den,
                                                 int nbf, int flag_vad, float eps)
{
    const float beta = 0.98f, rsb_min = 0.079432823f;
    if (fstage == 1) {
        float n2 = noise * noise;                                    // noise estimation in energy, NoiseSup.c:486-517
        if (nbf < 11) {
            const float lambda = 1 - 1 / (float)nbf;
            n2 = lambda * n2 + (1 - lambda) * psd_mean;
        } else {
            const float up = (float)(0.9 + 0.1 * (double)(psd_mean / (psd_mean + n2)) *
                                               (1.0 + 1.0 / (1.0 + 0.1 * (double)(psd_mean / n2))));
            n2 = n2 * up;
        }
        noise = (float)sqrt((double)n2);
        if (noise < eps) noise = eps;
    }
    const float nsig = (float)sqrt((double)psd);                     // :522-526
    const float pm = (float)sqrt((double)psd_mean);
    if (fstage == 0) {
        const float lambda = (nbf < 100) ? (1 - 1 / (float)nbf) : 0.99f;
        if (flag_vad == 0) {
            noise = lambda * noise + (1 - lambda) * pm;
            if (noise < eps) noise = eps;
        }
    }
    const float post = (pm / noise) - 1;                             // :551-560
    float prio = beta * (den / noise) + (1 - beta) * ((0 > post) ? 0 : post);
    float W = prio / (1 + prio);
    prio = W * pm / noise;
    prio = (prio > rsb_min) ? prio : rsb_min;
    W = prio / (1 + prio);
    den = W * nsig;
    return W;
}
#else
__device__ __forceinline__ float filter_calc_bin(const int fstage, float psd, float psd_mean, float &noise, float &den,
                                                 int nbf, int flag_vad, float eps)
{
    if (fstage == 1) {
        float n2 = noise * noise;
        if (nbf < 11) {
            const float lambda = 1.0f - 1.0f / (float)nbf;
            n2 = lambda * n2 + (1.0f - lambda) * psd_mean;
        } else {
            // NoiseSup.c:509-512 (double in the reference; the factor lies in [0.9, 1.1], float is ample)
            const float up = 0.9f + 0.1f * fdiv(psd_mean, psd_mean + n2) * (1.0f + fdiv(1.0f, 1.0f + 0.1f * fdiv(psd_mean, n2)));
            n2 *= up;
        }
        noise = fsqrt(n2);
        if (noise < eps) noise = eps;
    }
    const float nsig = fsqrt(psd);
    const float pm = fsqrt(psd_mean);
    if (fstage == 0) {
        const float lambda = (nbf < 100) ? (1.0f - 1.0f / (float)nbf) : 0.99f;
        if (flag_vad == 0) {
            noise = lambda * noise + (1.0f - lambda) * pm;
            if (noise < eps) noise = eps;
        }
    }
    const float inv_noise = fdiv(1.0f, noise);
    const float post = pm * inv_noise - 1.0f;
    float prio = 0.98f * (den * inv_noise) + (1.0f - 0.98f) * ((0.0f > post) ? 0.0f : post);
    float W = fdiv(prio, 1.0f + prio);
    prio = W * pm * inv_noise;
    prio = (prio > 0.079432823f) ? prio : 0.079432823f;
    W = fdiv(prio, 1.0f + prio);
    den = W * nsig;
    return W;
}
#endif

}  // namespace SE_NSPACE

#include "se_ns_stage.cuh"
#include "se_ns_feat.cuh"
#include "se_ns_hp16k.cuh"
#include "se_ns_kernel.cuh"
