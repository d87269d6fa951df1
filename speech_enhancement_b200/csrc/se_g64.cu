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

__global__ void __launch_bounds__(64) ibm_kernel(const int16_t *pure, const int16_t *noise, const int64_t *offsets,
                                                 int utt_begin, float *irm, uint32_t *bits,
                                                 const int64_t *frame_offsets)
{
    __shared__ float s_p[kTile], s_n[kTile];
    const int u = utt_begin + blockIdx.x, c = threadIdx.x, warp = c >> 5;
    const int64_t off = offsets[u], n = offsets[u + 1] - off;
    const int64_t f0 = frame_offsets[u];
    const int64_t frames = (n >= 320) ? (n - 320) / 160 + 1 : 0;
    const GtCoef k = load_coef(c);
    GtState gp, gn; gt_reset(gp); gt_reset(gn);
    HcState hp, hn; hc_reset(hp, c_gt.hc); hc_reset(hn, c_gt.hc);
    FrameAcc ap_cur, an_cur, ap_old, an_old;   // cur: frame h (window idx r); old: frame h-1 (idx 160+r)
    acc_reset(ap_cur); acc_reset(an_cur); acc_reset(ap_old); acc_reset(an_old);
    int r = 0;
    int64_t h = 0;
    for (int64_t t0 = 0; t0 < n; t0 += kTile) {
        const int len = (int)((n - t0 < kTile) ? (n - t0) : kTile);
        __syncthreads();
        if (c < len) { s_p[c] = (float)pure[off + t0 + c]; s_n[c] = (float)noise[off + t0 + c]; }
        __syncthreads();
        for (int t = 0; t < len; t++) {
            const float vp = (float)cast_i16(hc_step(hp, c_gt.hc, gt_step<true>(gp, k, s_p[t])));
            const float vn = (float)cast_i16(hc_step(hn, c_gt.hc, gt_step<true>(gn, k, s_n[t])));
            const float w0 = c_ibm_win[r], w1 = c_ibm_win[160 + r];
            acc_add(ap_cur, __fmul_rn(vp, w0), r);
            acc_add(an_cur, __fmul_rn(vn, w0), r);
            acc_add(ap_old, __fmul_rn(vp, w1), 160 + r);
            acc_add(an_old, __fmul_rn(vn, w1), 160 + r);
            if (++r == 160) {
                r = 0;
                if (h >= 1 && h - 1 < frames) {
                    const float sp = acc_power(ap_old), sn = acc_power(an_old);
                    const int64_t f = f0 + h - 1;
                    if (irm) irm[f * 64 + c] = sp / __fadd_rn(sp, sn);
                    const unsigned b = __ballot_sync(0xffffffffu, sn < sp);
                    if (bits && (c & 31) == 0) bits[f * 2 + warp] = b;
                }
                ap_old = ap_cur; an_old = an_cur;
                acc_reset(ap_cur); acc_reset(an_cur);
                h++;
            }
        }
    }
}

}  // namespace se

#include "se_g64_resynth.cuh"
