// resynth(): zero-phase (forward + time-reversed) gammatone, mask weights, 64-channel sum.
// resyth_64sub_IBM/cpp/extractwav.cpp:9-132 (+ _ori / 1dnn variants).  Included by se_g64.cu.
#pragma once

namespace se {

// Work item of the two resynthesis passes: output samples [a, b) of utterance u.  The forward pass runs
// over [fs, fe) from rest and the reversed pass from fe-1 down to a, also from rest.
//   EXACT mode : one item per utterance, fs = 0, fe = n  (the reference's arithmetic, bit for bit).
//   default    : the utterance is cut into time chunks with kGtWarm samples of run-in on either side
//                (fs = max(0, a - W), fe = min(n, b + W)).  The 4th-order gammatone's state decays like
//                k^3 z^k; for the slowest channel (cf 50 Hz, z = 0.988) that is below 1e-7 of its peak after
//                2200 samples, so with W = 2560 the run-in error is under float rounding while the batch
//                exposes ~10x more independent work to the SMs.
struct GtItem {
    int64_t sample_off;      // first sample of utterance u in the batch
    int64_t y_off;           // first scratch row (one row = 64 floats) of this item
    int64_t frame_off;       // first mask row of utterance u
    int32_t frames;          // mask rows of utterance u
    int32_t fs, fe, a, b;
};
constexpr int kGtWarm = 2560;
constexpr int kGtChunk = 16000;      // smallest time chunk; the host doubles it while enough work items remain
constexpr int kGtChunkLatency = 2048; // SE_FLAG_LATENCY: chunk of a lone utterance (a multiple of the 64-sample tile)

// Time-chunk length of the default mode.  Every chunk pays 2 x kGtWarm samples of run-in, so long utterances get longer
// chunks -- a function of the utterance's own length (and the caller's flags) only, so that its output does not depend
// on what else is in the batch (at least 4 chunks per utterance are kept for parallelism, at most 8 x kGtChunk each).
// SE_FLAG_LATENCY trades throughput for latency: 2 048-sample chunks turn one 10 s utterance into 79 work items of
// 7 168 + 4 608 filter steps instead of 10 items of 21 120 + 18 560, for 2.9x the arithmetic.
static inline int64_t gt_chunk_len(int64_t n, int flags)
{
    if (flags & SE_FLAG_LATENCY) return kGtChunkLatency;
    int64_t chunk = kGtChunk;
    while (chunk < 8 * kGtChunk && n >= 8 * chunk) chunk *= 2;
    return chunk;
}

// ---- pass 1: y1[n][c] = gammatone_c(x)[n] / midEar_c, channel-minor scratch (coalesced) ---------
template <bool EXACT>
__global__ void __launch_bounds__(64) gt_fwd_kernel(const int16_t *samples, const GtItem *items, float *Y)
{
    __shared__ float s_in[kTile];
    const GtItem it = items[blockIdx.x];
    const int c = threadIdx.x;
    const int16_t *x = samples + it.sample_off;
    float *y = Y + it.y_off * 64 + c;
    const GtCoef k = load_coef(c);
    const float inv_mid = 1.0f / k.mid;
    GtState g; gt_reset(g);
    const int n = it.fe - it.fs;
    for (int t0 = 0; t0 < n; t0 += kTile) {
        const int len = (n - t0 < kTile) ? (n - t0) : kTile;
        __syncthreads();
        if (c < len) s_in[c] = (float)x[it.fs + t0 + c];
        __syncthreads();
#pragma unroll 4
        for (int t = 0; t < len; t++) {
            const float v = gt_step<EXACT>(g, k, s_in[t]);
            y[(int64_t)(t0 + t) * 64] = EXACT ? __fdiv_rn(v, k.mid) : v * inv_mid;
        }
    }
}

// ---- pass 2: time-reversed gammatone over y1, /midEar, x mask weight, sum over channels ---------
// weight of sample n (hop h=n/160, r=n%160): fall[r]*m[h] (frame h's falling half, added first)
//                                          + rise[r]*m[h+1] (frame h+1's rising half), gated per mode.
template <bool EXACT, int MODE>
__global__ void __launch_bounds__(64) gt_bwd_kernel(const float *Y, const GtItem *items, const float *mask, int16_t *out, float *out_f)
{
    constexpr int T = 32;
    __shared__ float tile[2][T][65];
    const GtItem it = items[blockIdx.x];
    const int c = threadIdx.x;
    const float *y = Y + it.y_off * 64 + c;
    const int64_t frames = it.frames;
    const float *mk = mask + it.frame_off * 64 + c;
    int16_t *o = out + it.sample_off;
    const GtCoef k = load_coef(c);
    const float inv_mid = 1.0f / k.mid;
    GtState g; gt_reset(g);

    auto gate = [&](int64_t f) -> float {   // effective mask value of frame f for this channel (0 = inactive)
        if (f < 0 || f >= frames) return 0.0f;
        const float m = mk[f * 64];
        if (MODE == SE_RESYNTH_IBM) return (m > 0.5f) ? 1.0f : 0.0f;
        return (m > 0.0f) ? m : 0.0f;
    };
    // run-in: samples [b, fe) only feed the filter state (default mode; empty in EXACT mode)
    int pos = it.fe - 1;
    for (; pos >= it.b; pos--) (void)gt_step<EXACT>(g, k, y[(int64_t)(pos - it.a) * 64]);

    const int n = it.b - it.a;           // outputs to produce, from time b-1 down to a
    int h = (n > 0) ? pos / 160 : 0;
    int r = (n > 0) ? pos % 160 : 0;
    float m_h = gate(h), m_h1 = gate(h + 1), m_prev = gate(h - 1);   // m_prev: prefetched for the next hop
    int buf = 0;
    for (int m0 = 0; m0 < n; m0 += T, buf ^= 1) {
        const int len = (n - m0 < T) ? (n - m0) : T;
        float in[T];
#pragma unroll
        for (int t = 0; t < T; t++) in[t] = (t < len) ? y[(int64_t)(it.b - 1 - (m0 + t) - it.a) * 64] : 0.0f;
#pragma unroll
        for (int t = 0; t < T; t++) {
            if (t < len) {
                float v = gt_step<EXACT>(g, k, in[t]);
                v = EXACT ? __fdiv_rn(v, k.mid) : v * inv_mid;
                float w;
                if (MODE == SE_RESYNTH_IBM) {
                    w = (m_h != 0.0f) ? ((m_h1 != 0.0f) ? c_w_both[r] : c_w_fall[r]) : ((m_h1 != 0.0f) ? c_w_rise[r] : 0.0f);
                } else {
                    // weight_e[] (float) += 0.5*(1+cos(..)) * v_erm : a double product added to a float
                    // (resyth_64sub_ori/cpp/extractwav.cpp:101-104); falling half of frame h first
                    w = 0.0f;
                    if (EXACT) {
                        if (m_h != 0.0f) w = (float)(c_w_fall_d[r] * (double)m_h);
                        // separately rounded double product and sum (no FMA contraction: the float output must match
                        // the reference bit for bit, SE_FLAG_OUT_F32)
                        if (m_h1 != 0.0f) w = (float)__dadd_rn((double)w, __dmul_rn(c_w_rise_d[r], (double)m_h1));
                    } else {
                        if (m_h != 0.0f) w = c_w_fall[r] * m_h;
                        if (m_h1 != 0.0f) w += c_w_rise[r] * m_h1;
                    }
                }
                tile[buf][t][c] = __fmul_rn(w, v);
                if (--r < 0) { r = 159; h--; m_h1 = m_h; m_h = m_prev; m_prev = gate(h - 1); }
            }
        }
        __syncthreads();
        if (c < len) {      // reference order: resynth_e[n] += w*rev for chan = 0..63 (extractwav.cpp:109-113)
            float acc = 0.0f;
#pragma unroll 8
            for (int ch = 0; ch < 64; ch++) acc = __fadd_rn(acc, tile[buf][c][ch]);
            // resynth_e[] (float) and its (short) cast, extractwav.cpp:119-122; SE_FLAG_OUT_F32 returns the float
            if (out_f) out_f[it.sample_off + it.b - 1 - (m0 + c)] = acc;
            else o[it.b - 1 - (m0 + c)] = cast_i16(acc);
        }
    }
}

// ---- default-mode kernels (same work items; FMA-contracted arithmetic, gain/midEar folded) --------
__global__ void __launch_bounds__(64) gt_fwd_fast_kernel(const int16_t *samples, const GtItem *items, float *Y)
{
    __shared__ float s_in[2][kTile];
    const GtItem it = items[blockIdx.x];
    const int c = threadIdx.x;
    const int16_t *x = samples + it.sample_off + it.fs;
    float *y = Y + it.y_off * 64 + c;
    const GtRot k = load_rot(c);
    const float scale = c_gt.gain[c] / c_gt.mid[c];
    GtStateR g; gtr_reset(g);
    const int n = it.fe - it.fs;
    const int warm = it.a - it.fs;           // run-in: filtered for the state only, not stored (a whole number of tiles)
    int buf = 0;
    if (c < n) s_in[0][c] = (float)x[c];
    for (int t0 = 0; t0 < n; t0 += kTile, buf ^= 1) {
        const int len = (n - t0 < kTile) ? (n - t0) : kTile;
        __syncthreads();
        if (t0 + kTile + c < n) s_in[buf ^ 1][c] = (float)x[t0 + kTile + c];     // next tile, other buffer
        float *yp = y + (int64_t)(t0 - warm) * 64;                                // rows are relative to a
        gt_rot_seed(g, k, t0);
        if (t0 < warm) {
#pragma unroll 16
            for (int t = 0; t < kTile; t++) (void)gt_step_rot(g, k, s_in[buf][t]);
        } else if (len == kTile) {
#pragma unroll 16
            for (int t = 0; t < kTile; t++) yp[t * 64] = gt_step_rot(g, k, s_in[buf][t]) * scale;
        } else {
            for (int t = 0; t < len; t++) yp[t * 64] = gt_step_rot(g, k, s_in[buf][t]) * scale;
        }
    }
}

// Reversed pass.  Tiles of 32 samples are aligned so that every tile lies inside one 160-sample hop
// (160 = 5 x 32): the two mask values that weight a hop are fetched once per tile, and the raised-cosine
// weight is  w = m[h+1] + (m[h] - m[h+1]) * fall[r]   (fall[r] + rise[r] = 1).
template <int MODE>
__global__ void __launch_bounds__(64) gt_bwd_fast_kernel(const float *Y, const GtItem *items, const float *mask, int16_t *out, float *out_f)
{
    constexpr int T = 32;
    __shared__ float tile[2][T][65];
    const GtItem it = items[blockIdx.x];
    const int c = threadIdx.x;
    const float *y = Y + it.y_off * 64 + c;
    const int frames = it.frames;
    const float *mk = mask + it.frame_off * 64 + c;
    int16_t *o = out + it.sample_off;
    const GtRot k = load_rot(c);
    const float scale = c_gt.gain[c] / c_gt.mid[c];    // this pass's own gain and midEar division
    GtStateR g; gtr_reset(g);
    int step = 0;                                      // samples filtered since the start from rest

    auto gate = [&](int f) -> float {
        if (f < 0 || f >= frames) return 0.0f;
        const float m = mk[(int64_t)f * 64];
        if (MODE == SE_RESYNTH_IBM) return (m > 0.5f) ? 1.0f : 0.0f;
        return (m > 0.0f) ? m : 0.0f;
    };
    // run-in over [b, fe): state only
    int pos = it.fe - 1;
    while (pos >= it.b) {
        const int len = (pos - it.b + 1 < 16) ? (pos - it.b + 1) : 16;
        const float *yp = y + (int64_t)(pos - it.a) * 64;
        float in[16];
#pragma unroll
        for (int t = 0; t < 16; t++) in[t] = (t < len) ? yp[-t * 64] : 0.0f;
        gt_rot_seed(g, k, step);
#pragma unroll
        for (int t = 0; t < 16; t++)
            if (t < len) (void)gt_step_rot(g, k, in[t]);
        pos -= len;
        step += len;
    }
    int buf = 0;
    int len = (pos >= it.a) ? (pos & 31) + 1 : 0;      // first tile ends on a multiple of 32
    while (pos >= it.a) {
        if (len > pos - it.a + 1) len = pos - it.a + 1;
        const int h = pos / 160, r_top = pos - h * 160;
        const float m_h = gate(h), m_h1 = gate(h + 1);
        const float wa = (m_h - m_h1) * scale, wb = m_h1 * scale;
        const float *yp = y + (int64_t)(pos - it.a) * 64;
        float in[T];
        gt_rot_seed(g, k, step);
        if (len == T) {
#pragma unroll
            for (int t = 0; t < T; t++) in[t] = yp[-t * 64];
#pragma unroll
            for (int t = 0; t < T; t++) {
                const float v = gt_step_rot(g, k, in[t]);
                tile[buf][t][c] = fmaf(wa, c_w_fall[r_top - t], wb) * v;
            }
        } else {
            for (int t = 0; t < len; t++) {
                const float v = gt_step_rot(g, k, yp[-t * 64]);
                tile[buf][t][c] = fmaf(wa, c_w_fall[r_top - t], wb) * v;
            }
        }
        __syncthreads();
        if (c < len) {
            float acc = 0.0f;
#pragma unroll 16
            for (int ch = 0; ch < 64; ch++) acc += tile[buf][c][ch];
            if (out_f) out_f[it.sample_off + pos - c] = acc;
            else o[pos - c] = cast_i16(acc);
        }
        pos -= len;
        step += len;
        len = T;
        buf ^= 1;
    }
}

// ---- known-answer kernel: gammaToneFilter on float signals, one thread per signal ---------------
template <bool EXACT>
__global__ void gammatone_kernel(const float *in, float *out, const int32_t *channel, int64_t n, int count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const GtCoef k = load_coef(channel[i]);
    GtState g; gt_reset(g);
    for (int64_t t = 0; t < n; t++) out[i * n + t] = gt_step<EXACT>(g, k, in[i * n + t]);
}

// gammaToneFilter with caller-supplied coefficients (any cf / bw / midEarCoeff, HuWang.h:49): coef[i] = {f1, f2, gain}
template <bool EXACT>
__global__ void gammatone_coef_kernel(const float *in, float *out, const float *coef, int64_t n, int count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    GtCoef k;
    k.f1 = coef[3 * i]; k.f2 = coef[3 * i + 1]; k.gain = coef[3 * i + 2]; k.mid = 1.0f;
    GtState g; gt_reset(g);
    for (int64_t t = 0; t < n; t++) out[i * n + t] = gt_step<EXACT>(g, k, in[i * n + t]);
}

// ---- known-answer kernel: hairCell (extractwav.cpp:213-258) on float signals, one thread per signal ---------
__global__ void haircell_kernel(const float *in, float *out, int64_t n, int count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    HcState h; hc_reset(h, c_gt.hc);
    for (int64_t t = 0; t < n; t++) out[i * n + t] = hc_step(h, c_gt.hc, in[i * n + t]);
}

// ---- addnoise (enhancement_extract_subband_linux/cpp/extractwav.cpp:6-35) -----------------------
// Energy sums accumulate in float with a double intermediate, sequentially (one thread / utterance).
__global__ void addnoise_gain_kernel(const int16_t *pure, const int16_t *noise, const int64_t *offsets, int n_utts,
                                     float db_lin, float *gain)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= n_utts) return;
    float ps = 0.0f, ns = 0.0f;
    for (int64_t i = offsets[u]; i < offsets[u + 1]; i++) {
        ps = (float)((double)ps + (double)((int)pure[i] * (int)pure[i]) * 1.0);
        ns = (float)((double)ns + (double)((int)noise[i] * (int)noise[i]) * 1.0);
    }
    gain[u] = sqrtf(ps / ns / db_lin);
}

__global__ void addnoise_mix_kernel(const int16_t *pure, int16_t *noise, int16_t *noisy, const int64_t *offsets,
                                    const float *gain)
{
    // 64 CTAs per utterance, utterance index on grid.x (grid.y is capped at 65 535, a corpus of short files is not)
    const int u = blockIdx.x >> 6, part = blockIdx.x & 63;
    const float g = gain[u];
    for (int64_t i = offsets[u] + part * blockDim.x + threadIdx.x; i < offsets[u + 1]; i += (int64_t)64 * blockDim.x) {
        const short z = cast_i16(__fmul_rn((float)noise[i], g));   // Q4: cast before the multiply
        noise[i] = z;
        noisy[i] = (short)((int)pure[i] + (int)z);
    }
}

}  // namespace se

#include "se_g64_host.cuh"
