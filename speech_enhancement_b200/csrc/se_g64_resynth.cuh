// resynth(): zero-phase (forward + time-reversed) gammatone, mask weights, 64-channel sum.
// resyth_64sub_IBM/cpp/extractwav.cpp:9-132 (+ _ori / 1dnn variants).  Included by se_g64.cu.
#pragma once

namespace se {

// ---- pass 1: y1[n][c] = gammatone_c(x)[n] / midEar_c, channel-minor scratch (coalesced) ---------
template <bool EXACT>
__global__ void __launch_bounds__(64) gt_fwd_kernel(const int16_t *samples, const int64_t *offsets, int utt_begin,
                                                    int64_t base_off, float *Y)
{
    __shared__ float s_in[kTile];
    const int u = utt_begin + blockIdx.x, c = threadIdx.x;
    const int64_t off = offsets[u], n = offsets[u + 1] - off;
    float *y = Y + (off - base_off) * 64 + c;
    const GtCoef k = load_coef(c);
    const float inv_mid = 1.0f / k.mid;
    GtState g; gt_reset(g);
    for (int64_t t0 = 0; t0 < n; t0 += kTile) {
        const int len = (int)((n - t0 < kTile) ? (n - t0) : kTile);
        __syncthreads();
        if (c < len) s_in[c] = (float)samples[off + t0 + c];
        __syncthreads();
#pragma unroll 4
        for (int t = 0; t < len; t++) {
            const float v = gt_step<EXACT>(g, k, s_in[t]);
            y[(t0 + t) * 64] = EXACT ? __fdiv_rn(v, k.mid) : v * inv_mid;
        }
    }
}

// ---- pass 2: time-reversed gammatone over y1, /midEar, x mask weight, sum over channels ---------
// weight of sample n (hop h=n/160, r=n%160): fall[r]*m[h] (frame h's falling half, added first)
//                                          + rise[r]*m[h+1] (frame h+1's rising half), gated per mode.
template <bool EXACT, int MODE>
__global__ void __launch_bounds__(64) gt_bwd_kernel(const float *Y, const int64_t *offsets, int utt_begin,
                                                    int64_t base_off, const float *mask, const int64_t *frame_offsets,
                                                    int16_t *out)
{
    constexpr int T = 32;
    __shared__ float tile[2][T][65];
    const int u = utt_begin + blockIdx.x, c = threadIdx.x;
    const int64_t off = offsets[u], n = offsets[u + 1] - off;
    const float *y = Y + (off - base_off) * 64 + c;
    const int64_t fo = frame_offsets[u];
    const int64_t frames = frame_offsets[u + 1] - fo;
    const float *mk = mask + fo * 64 + c;
    const GtCoef k = load_coef(c);
    const float inv_mid = 1.0f / k.mid;
    GtState g; gt_reset(g);

    auto gate = [&](int64_t f) -> float {   // effective mask value of frame f for this channel (0 = inactive)
        if (f < 0 || f >= frames) return 0.0f;
        const float m = mk[f * 64];
        if (MODE == SE_RESYNTH_IBM) return (m > 0.5f) ? 1.0f : 0.0f;
        return (m > 0.0f) ? m : 0.0f;
    };
    int64_t pos = n - 1;                 // time index of the sample being produced
    int64_t h = (n > 0) ? pos / 160 : 0;
    int r = (n > 0) ? (int)(pos % 160) : 0;
    float m_h = gate(h), m_h1 = gate(h + 1), m_prev = gate(h - 1);   // m_prev: prefetched for the next hop
    int buf = 0;
    for (int64_t m0 = 0; m0 < n; m0 += T, buf ^= 1) {
        const int len = (int)((n - m0 < T) ? (n - m0) : T);
        float in[T];
#pragma unroll
        for (int t = 0; t < T; t++) in[t] = (t < len) ? y[(n - 1 - (m0 + t)) * 64] : 0.0f;
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
                        if (m_h1 != 0.0f) w = (float)((double)w + c_w_rise_d[r] * (double)m_h1);
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
            out[off + (n - 1 - (m0 + c))] = cast_i16(acc);
        }
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
    const int u = blockIdx.y;
    const float g = gain[u];
    for (int64_t i = offsets[u] + blockIdx.x * blockDim.x + threadIdx.x; i < offsets[u + 1];
         i += (int64_t)gridDim.x * blockDim.x) {
        const short z = cast_i16(__fmul_rn((float)noise[i], g));   // Q4: cast before the multiply
        noise[i] = z;
        noisy[i] = (short)((int)pure[i] + (int)z);
    }
}

}  // namespace se

#include "se_g64_host.cuh"
