// In-warp 128-point complex FFT held in registers (4 points per lane): radix-4 passes in registers with
// lane <-> point transposes through shared memory (default) or shuffle butterflies (-DSE_FFT_SHFL, the
// round-1 form), and the real-FFT split helpers built on it.  Replaces the reference's in-place split-radix rfft
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

__device__ __forceinline__ float2 expmi(float frac_of_pi)   // exp(-i * pi * frac)
{
    float s, c;
    sincospif(frac_of_pi, &s, &c);
    return make_float2(c, -s);
}

// Spectrum scratch of one FFT-128: 132 float2 hold the result (zpad layout); the transposing FFT below also
// stages its two lane<->point transposes in it and needs 152.
constexpr int kZ = 152;

#ifdef SE_FFT_SHFL
// Per-lane twiddles, computed once per kernel.  The inter-lane stage twiddles are pre-selected per
// lane: the lower lane of a butterfly pair keeps (a + b) and gets twiddle 1, the upper lane keeps
// (b - a) * w, so every lane executes the same  t = b + sgn * a ;  a = t * w'  (no select, no branch).
struct FftTw {
    float2 q1, q2, q3;        // exp(-2 pi i * lane*q / 128)
    float2 s16, s8, s4, s2;   // upper lane of the span: exp(-2 pi i * (lane & (s-1)) / (2 s)); lower lane: 1
    unsigned sign;            // bit k set: lane is the upper one at span 2^k (k = 0..4)
};

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

constexpr int kTwFloats = 15;
__device__ __forceinline__ void fft_tw_to_array(const FftTw &t, float (&v)[kTwFloats])
{
    const float a[kTwFloats] = {t.q1.x, t.q1.y, t.q2.x, t.q2.y, t.q3.x, t.q3.y, t.s16.x, t.s16.y, t.s8.x, t.s8.y,
                                t.s4.x, t.s4.y, t.s2.x, t.s2.y, __uint_as_float(t.sign)};
#pragma unroll
    for (int i = 0; i < kTwFloats; i++) v[i] = a[i];
}
__device__ __forceinline__ FftTw fft_tw_from_table(const float *p /* tab + lane, stride 32 */)
{
    FftTw t;
    t.q1 = make_float2(p[0], p[32]);     t.q2 = make_float2(p[64], p[96]);    t.q3 = make_float2(p[128], p[160]);
    t.s16 = make_float2(p[192], p[224]); t.s8 = make_float2(p[256], p[288]);  t.s4 = make_float2(p[320], p[352]);
    t.s2 = make_float2(p[384], p[416]);  t.sign = __float_as_uint(p[448]);
    return t;
}
#else   // ---- default: radix-4 passes in registers, lane <-> point transposes through shared memory -------------
// 128 = 4 x 4 x 4 x 2.  Three radix-4 passes run in registers on the four points a lane holds; between them the
// points change lanes through the warp's spectrum scratch (two conflict-free STS.64/LDS.64 transposes) instead of
// through five shuffle stages (2 SHFL + 6 FP32 per point and stage); the last radix-2 is one shuffle stage between
// adjacent lanes.  ~125 instead of ~180 instructions per transform, and 8 shuffles instead of 40.
//   pass 1: lane l, points l + 32 j          -> y_q[l]      = W128^(l q)  sum_j z[l + 32 j] W4^(j q)
//   pass 2: lane (q, m), points m + 8 i      -> z_(q,c)[m]  = W32^(m c)   sum_i y_q[m + 8 i] W4^(i c)
//   pass 3: lane (q, c, r), points r + 2 i   -> w_d[r]      = W8^(r d)    sum_i z_(q,c)[r + 2 i] W4^(i d)
//   radix 2 over r (lane ^ 1)                -> X[q + 4 c + 16 d + 64 e] = w_d[0] + (-1)^e w_d[1],  e = lane & 1
struct FftTw {
    float2 q1, q2, q3;        // exp(-2 pi i * lane * k / 128)
    float2 p1, p2, p3;        // exp(-2 pi i * (lane & 7) * k / 32)
    float2 r1, r2, r3;        // odd lanes: exp(-2 pi i * k / 8); even lanes: 1
    float sgn;                // odd lanes -1, even lanes +1
};
constexpr int kTwFloats = 19;

__device__ __forceinline__ void fft_tw_init(FftTw &t, int lane)
{
    const float2 one = make_float2(1.0f, 0.0f);
    t.q1 = expmi((float)lane * (1.0f / 64.0f));
    t.q2 = expmi((float)(2 * lane) * (1.0f / 64.0f));
    t.q3 = expmi((float)(3 * lane) * (1.0f / 64.0f));
    const int m = lane & 7;
    t.p1 = expmi((float)m * (1.0f / 16.0f));
    t.p2 = expmi((float)(2 * m) * (1.0f / 16.0f));
    t.p3 = expmi((float)(3 * m) * (1.0f / 16.0f));
    t.r1 = (lane & 1) ? expmi(0.25f) : one;
    t.r2 = (lane & 1) ? expmi(0.5f) : one;
    t.r3 = (lane & 1) ? expmi(0.75f) : one;
    t.sgn = (lane & 1) ? -1.0f : 1.0f;
}
__device__ __forceinline__ void fft_tw_to_array(const FftTw &t, float (&v)[kTwFloats])
{
    const float a[kTwFloats] = {t.q1.x, t.q1.y, t.q2.x, t.q2.y, t.q3.x, t.q3.y, t.p1.x, t.p1.y, t.p2.x, t.p2.y,
                                t.p3.x, t.p3.y, t.r1.x, t.r1.y, t.r2.x, t.r2.y, t.r3.x, t.r3.y, t.sgn};
#pragma unroll
    for (int i = 0; i < kTwFloats; i++) v[i] = a[i];
}
__device__ __forceinline__ FftTw fft_tw_from_table(const float *p /* tab + lane, stride 32 */)
{
    FftTw t;
#ifdef SE_FFT_TW_DERIVE
    // Measured and NOT adopted: load the two base twiddles and the sign, derive the squares / cubes and select the
    // constant last-pass twiddles -- 14 fewer shared-memory wavefronts per frame (the pipe that bounds these kernels):
    // P16 -0.6 %, P8 -1.5 %, but the 2-ulp twiddles double the FFT's rounding noise and the cepstra of speech with
    // digital-silence runs (tests/test_gpu_fuzz.py, kind 6) leave their bar.  Parity first.
    t.q1 = make_float2(p[0], p[32]);
    t.p1 = make_float2(p[192], p[224]);
    t.sgn = p[576];
    t.q2 = cmul(t.q1, t.q1);
    t.q3 = cmul(t.q2, t.q1);
    t.p2 = cmul(t.p1, t.p1);
    t.p3 = cmul(t.p2, t.p1);
    const bool odd = t.sgn < 0.0f;
    const float h = 0.70710678118654752440f;
    t.r1 = make_float2(odd ? h : 1.0f, odd ? -h : 0.0f);
    t.r2 = make_float2(odd ? 0.0f : 1.0f, odd ? -1.0f : 0.0f);
    t.r3 = make_float2(odd ? -h : 1.0f, odd ? -h : 0.0f);
#else
    t.q1 = make_float2(p[0], p[32]);     t.q2 = make_float2(p[64], p[96]);    t.q3 = make_float2(p[128], p[160]);
    t.p1 = make_float2(p[192], p[224]);  t.p2 = make_float2(p[256], p[288]);  t.p3 = make_float2(p[320], p[352]);
    t.r1 = make_float2(p[384], p[416]);  t.r2 = make_float2(p[448], p[480]);  t.r3 = make_float2(p[512], p[544]);
    t.sgn = p[576];
#endif
    return t;
}

// y_k = sum_j v_j (-i)^(j k), then y_k *= w_k (k = 1..3)
__device__ __forceinline__ void radix4_tw(float2 (&v)[4], const float2 w1, const float2 w2, const float2 w3)
{
    const float2 a02 = cadd(v[0], v[2]), s02 = csub(v[0], v[2]);
    const float2 a13 = cadd(v[1], v[3]), s13 = csub(v[1], v[3]);
    v[0] = cadd(a02, a13);
    v[1] = cmul(make_float2(s02.x + s13.y, s02.y - s13.x), w1);   // s02 - i*s13
    v[2] = cmul(csub(a02, a13), w2);
    v[3] = cmul(make_float2(s02.x - s13.y, s02.y + s13.x), w3);   // s02 + i*s13
}
#endif

// Conflict-free padded index for the 128-entry spectrum scratch (132 float2).
__device__ __forceinline__ int zpad(int k) { return k + (k >> 5); }

__device__ __forceinline__ void fft128_store(const float2 (&v)[4], float2 *Z, int lane)
{
    const int r = (int)(__brev((unsigned)lane) >> 27);
#pragma unroll
    for (int q = 0; q < 4; q++) Z[zpad(q + 4 * r)] = v[q];
}

// NF complex FFT-128 at once (their instruction streams interleave).  In: v[f][j] = z_f[lane + 32 j].  Out: Z[f] holds
// the spectrum in zpad layout; the caller synchronises the warp before reading it.  Z[f] may still be read by other
// lanes when this is called (the routine synchronises before its first store).
template <int NF>
__device__ __forceinline__ void fft128_spectrum(float2 (&v)[NF][4], float2 *const (&Z)[NF], const FftTw &t, int lane)
{
#ifdef SE_FFT_SHFL
#pragma unroll
    for (int f = 0; f < NF; f++) fft128_warp(v[f], t, lane);
    __syncwarp();
#pragma unroll
    for (int f = 0; f < NF; f++) fft128_store(v[f], Z[f], lane);
#else
#pragma unroll
    for (int f = 0; f < NF; f++) radix4_tw(v[f], t.q1, t.q2, t.q3);
    // transpose 1: y_q[l] -> row (l >> 3) (stride 40), column 8 q + (l & 7); lane L then reads column L of every row
    const int w1 = (lane >> 3) * 40 + (lane & 7);
    __syncwarp();
#pragma unroll
    for (int f = 0; f < NF; f++)
#pragma unroll
        for (int q = 0; q < 4; q++) Z[f][w1 + 8 * q] = v[f][q];
    __syncwarp();
#pragma unroll
    for (int f = 0; f < NF; f++)
#pragma unroll
        for (int i = 0; i < 4; i++) v[f][i] = Z[f][40 * i + lane];
#pragma unroll
    for (int f = 0; f < NF; f++) radix4_tw(v[f], t.p1, t.p2, t.p3);
    // transpose 2: lane (q, m) -> row (m >> 1) (stride 34), column 8 q + 2 c + (m & 1)
    const int w2 = ((lane & 7) >> 1) * 34 + (lane >> 3) * 8 + (lane & 1);
    __syncwarp();
#pragma unroll
    for (int f = 0; f < NF; f++)
#pragma unroll
        for (int c = 0; c < 4; c++) Z[f][w2 + 2 * c] = v[f][c];
    __syncwarp();
#pragma unroll
    for (int f = 0; f < NF; f++)
#pragma unroll
        for (int i = 0; i < 4; i++) v[f][i] = Z[f][34 * i + lane];
#pragma unroll
    for (int f = 0; f < NF; f++) radix4_tw(v[f], t.r1, t.r2, t.r3);
    // radix 2 between adjacent lanes, then the natural-order store: bin q + 4 c + 16 d + 64 e
    const int k0 = (lane >> 3) + 4 * ((lane >> 1) & 3) + 66 * (lane & 1);      // 64 e + zpad's (k >> 5) = 2 e
    __syncwarp();
#pragma unroll
    for (int f = 0; f < NF; f++)
#pragma unroll
        for (int d = 0; d < 4; d++) {
            const float2 b = shfl_xor2(v[f][d], 1);
            Z[f][k0 + 16 * d + (d >> 1)] = make_float2(fmaf(t.sgn, v[f][d].x, b.x), fmaf(t.sgn, v[f][d].y, b.y));
        }
#endif
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
