// 64-channel gammatone path: sub-band analysis, fused IRM/IBM, zero-phase masked resynthesis.
//
// Mapping: one CTA of 64 threads per utterance, thread c = channel c, serial in time (the IIR is a
// per-sample recurrence; all parallelism is utterance x channel).  The utterance's int16 samples are
// staged once per 64-sample tile through shared memory and broadcast to the 64 channels; per-channel
// outputs go out coalesced (channel-minor scratch for the resynthesis intermediate, shared-memory
// transposes for the channel-major sub-band "wav" rows).
#include "se_internal.h"
#include "se_g64_dev.cuh"

namespace se {

__constant__ GtDevTables c_gt;
__constant__ float c_ibm_win[320];     // Hamming-320 of our SpecInfo restatement (se_oracle.c)
__constant__ float c_w_fall[160], c_w_rise[160], c_w_both[160];   // resynth() raised-cosine halves
__constant__ double c_w_fall_d[160], c_w_rise_d[160];              // the same halves as the reference's double expressions

constexpr int kTile = 64;

__device__ __forceinline__ GtCoef load_coef(int c)
{
    GtCoef k;
    k.f1 = c_gt.f1[c]; k.f2 = c_gt.f2[c]; k.gain = c_gt.gain[c]; k.mid = c_gt.mid[c];
    return k;
}

__device__ __forceinline__ GtRot load_rot(int c)
{
    GtRot k;
    k.rho = c_gt.rho[c]; k.rho2 = 2.0f * k.rho; k.rc = c_gt.rot_c[c]; k.rs = c_gt.rot_s[c]; k.turns = c_gt.turns[c];
    return k;
}

// ------------------------------------------------------------------------------------------------
// subbband(): gammatone -> hair cell -> (short), 64 channel-major rows per utterance
// (enhancement_extract_subband_linux/cpp/extractwav.cpp:36-97)
// ------------------------------------------------------------------------------------------------
template <bool EXACT>
__global__ void __launch_bounds__(64) subband_kernel(const int16_t *samples, const int64_t *offsets, int utt_begin,
                                                     int16_t *out)
{
    __shared__ float s_in[kTile];
    __shared__ short s_out[64][kTile + 2];
    const int u = utt_begin + blockIdx.x, c = threadIdx.x, lane = c & 31, warp = c >> 5;
    const int64_t off = offsets[u], n = offsets[u + 1] - off;
    const int16_t *x = samples + off;
    int16_t *o = out + 64 * off;
    const GtCoef k = load_coef(c);
    GtState g; gt_reset(g);
    HcState h; hc_reset(h, c_gt.hc);
    for (int64_t t0 = 0; t0 < n; t0 += kTile) {
        const int len = (int)((n - t0 < kTile) ? (n - t0) : kTile);
        __syncthreads();
        if (c < len) s_in[c] = (float)x[t0 + c];
        __syncthreads();
        for (int t = 0; t < len; t++) {
            const float y = gt_step<EXACT>(g, k, s_in[t]);
            s_out[c][t] = cast_i16(hc_step(h, c_gt.hc, y));
        }
        __syncthreads();
        for (int r = warp * 32; r < warp * 32 + 32; r++) {
            int16_t *row = o + (int64_t)r * n + t0;
            if (lane < len) row[lane] = s_out[r][lane];
            if (lane + 32 < len) row[lane + 32] = s_out[r][lane + 32];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// make_single_IBM fused with the two subbband() passes that feed it
// (enhancement_extract_subband_linux/cpp/show_IBM.cpp:92-213).  Per (frame f, channel c):
//   S_x = sum_{k<256} |FFT512(hamming320 * x_c[160 f .. 160 f + 320))|^2,  x in {pure, noise}
//   IRM = S_p/(S_p+S_n),  IBM = (S_n < S_p)
// S is evaluated through Parseval:  S = (512*sum(v^2) + (sum v)^2 - (sum (-1)^k v)^2)/2, the three
// sums accumulated sequentially in k exactly like oracle/se_oracle.c:so_single_ibm does.
// ------------------------------------------------------------------------------------------------
struct FrameAcc { float sq, s0, s1; };
__device__ __forceinline__ void acc_reset(FrameAcc &a) { a.sq = a.s0 = a.s1 = 0.0f; }
__device__ __forceinline__ void acc_add(FrameAcc &a, float v, int k)
{
    a.sq = __fmaf_rn(v, v, a.sq);
    a.s0 = __fadd_rn(a.s0, v);
    a.s1 = (k & 1) ? __fsub_rn(a.s1, v) : __fadd_rn(a.s1, v);
}
__device__ __forceinline__ float acc_power(const FrameAcc &a)
{
    float t = __fmul_rn(512.0f, a.sq);
    t = __fadd_rn(t, __fmul_rn(a.s0, a.s0));
    t = __fsub_rn(t, __fmul_rn(a.s1, a.s1));
    return __fmul_rn(0.5f, t);
}

// 128 threads per utterance: thread = (signal, channel), signal 0 = pure (warps 0-1), 1 = noise (warps 2-3), so
// the two subbband() passes of one utterance run side by side (twice the resident warps per utterance to hide the
// latency of the serial float/double recurrences).  One tile = one 160-sample hop, so a frame's second half
// completes exactly at a tile end; the two signals' frame powers meet in shared memory there.
constexpr int kIbmHop = 160;
__global__ void __launch_bounds__(128) ibm_kernel(const int16_t *pure, const int16_t *noise, const int64_t *offsets,
                                                  int utt_begin, float *irm, uint32_t *bits,
                                                  const int64_t *frame_offsets)
{
    __shared__ float s_x[2][kIbmHop];
    __shared__ float s_pow[2][64];
    const int u = utt_begin + blockIdx.x, sig = threadIdx.x >> 6, c = threadIdx.x & 63, warp = c >> 5;
    const int64_t off = offsets[u], n = offsets[u + 1] - off;
    const int64_t f0 = frame_offsets[u];
    const int64_t frames = (n >= 320) ? (n - 320) / 160 + 1 : 0;
    const int16_t *x = (sig ? noise : pure) + off;
    const GtCoef k = load_coef(c);
    GtState g; gt_reset(g);
    HcState hs; hc_reset(hs, c_gt.hc);
    FrameAcc a_cur, a_old;        // cur: frame h (window idx r); old: frame h-1 (idx 160+r)
    acc_reset(a_cur); acc_reset(a_old);
    int64_t h = 0;
    for (int64_t t0 = 0; t0 < n; t0 += kIbmHop, h++) {
        const int len = (int)((n - t0 < kIbmHop) ? (n - t0) : kIbmHop);
        __syncthreads();
        for (int i = c; i < len; i += 64) s_x[sig][i] = (float)x[t0 + i];
        __syncthreads();
        const float *xin = s_x[sig];
        for (int r = 0; r < len; r++) {
            const float v = (float)cast_i16(hc_step(hs, c_gt.hc, gt_step<true>(g, k, xin[r])));
            acc_add(a_cur, __fmul_rn(v, c_ibm_win[r]), r);
            acc_add(a_old, __fmul_rn(v, c_ibm_win[160 + r]), 160 + r);
        }
        if (len < kIbmHop) break;                         // ragged tail: no further frame completes
        const bool emit = h >= 1 && h - 1 < frames;       // uniform over the CTA
        if (emit) s_pow[sig][c] = acc_power(a_old);
        a_old = a_cur;
        acc_reset(a_cur);
        if (emit) {
            __syncthreads();
            if (sig == 0) {
                const float sp = s_pow[0][c], sn = s_pow[1][c];
                const int64_t f = f0 + h - 1;
                if (irm) irm[f * 64 + c] = sp / __fadd_rn(sp, sn);
                const unsigned b = __ballot_sync(0xffffffffu, sn < sp);
                if (bits && (c & 31) == 0) bits[f * 2 + warp] = b;
            }
        }
    }
}

// make_single_IBM on MATERIALISED sub-band rows (what the reference reads back from the 2 x 64 wav files that
// subbband() wrote, show_IBM.cpp:118-144): rows are laid out like se_subband_batch's output.  One CTA per frame of
// the batch, thread = (signal, channel); the three Parseval sums run over the frame's 320 samples in the same order
// as in ibm_kernel, so both entry points give identical values on identical sub-band samples.
__global__ void __launch_bounds__(128) ibm_rows_kernel(const int16_t *pure64, const int16_t *noise64, const int64_t *offsets,
                                                       int n_utts, float *irm, uint32_t *bits, const int64_t *frame_offsets)
{
    __shared__ float s_pow[2][64];
    const int64_t F = blockIdx.x;
    int lo = 0, hi = n_utts;                       // utterance u with frame_offsets[u] <= F < frame_offsets[u + 1]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (frame_offsets[mid] <= F) lo = mid; else hi = mid;
    }
    const int u = lo, sig = threadIdx.x >> 6, c = threadIdx.x & 63, warp = c >> 5;
    const int64_t off = offsets[u], n = offsets[u + 1] - off, f = F - frame_offsets[u];
    const int64_t frames = (n >= 320) ? (n - 320) / 160 + 1 : 0;
    if (f >= frames) return;                       // spare capacity rows of this utterance (uniform over the CTA)
    const int16_t *x = (sig ? noise64 : pure64) + 64 * off + (int64_t)c * n + 160 * f;
    FrameAcc a;
    acc_reset(a);
    for (int k = 0; k < 320; k++) acc_add(a, __fmul_rn((float)x[k], c_ibm_win[k]), k);
    s_pow[sig][c] = acc_power(a);
    __syncthreads();
    if (sig == 0) {
        const float sp = s_pow[0][c], sn = s_pow[1][c];
        if (irm) irm[F * 64 + c] = sp / __fadd_rn(sp, sn);
        const unsigned b = __ballot_sync(0xffffffffu, sn < sp);
        if (bits && (c & 31) == 0) bits[F * 2 + warp] = b;
    }
}

}  // namespace se

#include "se_g64_resynth.cuh"
#include "se_g64_noalloc.cuh"
