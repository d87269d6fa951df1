// In-warp 128-point complex FFT held in registers (4 points per lane) with shuffle butterflies, and
// the real-FFT split helpers built on it.  Replaces the reference's in-place split-radix rfft
// (etsi/cpp/rfft.c:45-180) for the sizes the hot path uses:
//   * 256-point real FFT  (P8 NoiseSup, CompCeps)          = one complex FFT-128 + split
//   * the plugin's truncated rfft(x,512,8)   = rfft256(x[0::2]) ++ rfft128(x[1::4]) ++
//     rfft128(x[3::4])                                      = two complex FFT-128 + splits
#pragma once
#include <cuda_runtime.h>

namespace SE_NSPACE {

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 shfl_xor2(float2 v, int m)
{
    return make_float2(__shfl_xor_sync(0xffffffffu, v.x, m), __shfl_xor_sync(0xffffffffu, v.y, m));
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
    return v;
}

// Per-lane twiddles, computed once per kernel.  The inter-lane stage twiddles are pre-selected per
// lane: the lower lane of a butterfly pair keeps (a + b) and gets twiddle 1, the upper lane keeps
// (b - a) * w, so every lane executes the same  t = b + sgn * a ;  a = t * w'  (no select, no branch).
struct FftTw {
    float2 q1, q2, q3;        // exp(-2 pi i * lane*q / 128)
    float2 s16, s8, s4, s2;   // upper lane of the span: exp(-2 pi i * (lane & (s-1)) / (2 s)); lower lane: 1
    unsigned sign;            // bit k set: lane is the upper one at span 2^k (k = 0..4)
};

__device__ __forceinline__ float2 expmi(float frac_of_pi)   // exp(-i * pi * frac)
{
    float s, c;
    sincospif(frac_of_pi, &s, &c);
    return make_float2(c, -s);
}

__device__ __forceinline__ void fft_tw_init(FftTw &t, int lane)
{
    const float2 one = make_float2(1.0f, 0.0f);
    t.q1 = expmi((float)lane * (1.0f / 64.0f));
    t.q2 = expmi((float)(2 * lane) * (1.0f / 64.0f));
    t.q3 = expmi((float)(3 * lane) * (1.0f / 64.0f));
    t.s16 = (lane & 16) ? expmi((float)(lane & 15) * (1.0f / 16.0f)) : one;
    t.s8 = (lane & 8) ? expmi((float)(lane & 7) * (1.0f / 8.0f)) : one;
    t.s4 = (lane & 4) ? expmi((float)(lane & 3) * (1.0f / 4.0f)) : one;
    t.s2 = (lane & 2) ? expmi((float)(lane & 1) * (1.0f / 2.0f)) : one;
    t.sign = (unsigned)lane;
}

template <int SPAN>
__device__ __forceinline__ void bfly(float2 &a, const float2 w, const float sgn)
{
    const float2 b = shfl_xor2(a, SPAN);
    // lower lane: a + b ; upper lane: (partner - own) = x[i] - x[i+span]
    const float2 t = make_float2(fmaf(sgn, a.x, b.x), fmaf(sgn, a.y, b.y));
    a = (SPAN == 1) ? t : cmul(t, w);
}

// In: v[j] = z[lane + 32 j].  Out: v[q] = Z[q + 4*bitrev5(lane)].
__device__ __forceinline__ void fft128_warp(float2 (&v)[4], const FftTw &t, int lane)
{
    // radix-4 over j (in registers): y_q = sum_j v_j (-i)^(j q)
    const float2 a02 = cadd(v[0], v[2]), s02 = csub(v[0], v[2]);
    const float2 a13 = cadd(v[1], v[3]), s13 = csub(v[1], v[3]);
    const float2 y0 = cadd(a02, a13);
    const float2 y2 = csub(a02, a13);
    const float2 y1 = make_float2(s02.x + s13.y, s02.y - s13.x);   // s02 - i*s13
    const float2 y3 = make_float2(s02.x - s13.y, s02.y + s13.x);   // s02 + i*s13
    v[0] = y0;
    v[1] = cmul(y1, t.q1);
    v[2] = cmul(y2, t.q2);
    v[3] = cmul(y3, t.q3);
    // 32-point DIF across lanes
    const float g16 = (lane & 16) ? -1.0f : 1.0f, g8 = (lane & 8) ? -1.0f : 1.0f, g4 = (lane & 4) ? -1.0f : 1.0f,
                g2 = (lane & 2) ? -1.0f : 1.0f, g1 = (lane & 1) ? -1.0f : 1.0f;
#pragma unroll
    for (int q = 0; q < 4; q++) bfly<16>(v[q], t.s16, g16);
#pragma unroll
    for (int q = 0; q < 4; q++) bfly<8>(v[q], t.s8, g8);
#pragma unroll
    for (int q = 0; q < 4; q++) bfly<4>(v[q], t.s4, g4);
#pragma unroll
    for (int q = 0; q < 4; q++) bfly<2>(v[q], t.s2, g2);
#pragma unroll
    for (int q = 0; q < 4; q++) bfly<1>(v[q], t.s2, g1);
}

// Conflict-free padded index for the 128-entry spectrum scratch (132 float2).
__device__ __forceinline__ int zpad(int k) { return k + (k >> 5); }

__device__ __forceinline__ void fft128_store(const float2 (&v)[4], float2 *Z, int lane)
{
    const int r = (int)(__brev((unsigned)lane) >> 27);
#pragma unroll
    for (int q = 0; q < 4; q++) Z[zpad(q + 4 * r)] = v[q];
}

// Bin k (0..128) of the 256-point real FFT whose even/odd samples were packed as z = x[2m] + i x[2m+1];
// Z is the FFT-128 of z, tw = (cos, sin)(2 pi k / 256).
__device__ __forceinline__ float2 rsplit256(const float2 *Z, int k, const float2 tw)
{
    const float2 zk = Z[zpad(k & 127)], zn = Z[zpad((128 - k) & 127)];
    const float er = 0.5f * (zk.x + zn.x), ei = 0.5f * (zk.y - zn.y);
    const float dr = 0.5f * (zk.x - zn.x), di = 0.5f * (zk.y + zn.y);
    return make_float2(er + (tw.x * di - tw.y * dr), ei - (tw.x * dr + tw.y * di));
}

// Two 128-point real FFTs packed as w = b + i c; W is the FFT-128 of w.  k = 0..64.
__device__ __forceinline__ void rsplit_pair(const float2 *W, int k, float2 &B, float2 &C)
{
    const float2 wk = W[zpad(k & 127)], wn = W[zpad((128 - k) & 127)];
    B = make_float2(0.5f * (wk.x + wn.x), 0.5f * (wk.y - wn.y));
    const float dr = 0.5f * (wk.x - wn.x), di = 0.5f * (wk.y + wn.y);
    C = make_float2(di, -dr);
}

}  // namespace SE_NSPACE
