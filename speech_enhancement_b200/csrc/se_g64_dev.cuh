// Device building blocks of the 64-channel gammatone path: one IIR step of gammaToneFilter and one
// step of the Meddis hair cell, written so that the EXACT variants round every product and sum
// separately, in the reference's order (bit-identical to its float arithmetic), while the fast
// variants let the compiler contract to FMA.
#pragma once
#include <cuda_runtime.h>

namespace se {

struct GtCoef { float f1, f2, gain, mid; };
struct GtState { float p0, p1, p2, p3, q0, q1, q2, q3; };

__device__ __forceinline__ void gt_reset(GtState &s) { s.p0 = s.p1 = s.p2 = s.p3 = s.q0 = s.q1 = s.q2 = s.q3 = 0.0f; }

// gammaToneFilter inner loop, resyth_64sub_IBM/cpp/extractwav.cpp:191-211.
// Returns output[n] = p[3]*gain BEFORE the state update (one-sample delay).
template <bool EXACT> __device__ __forceinline__ float gt_step(GtState &s, const GtCoef &c, float in)
{
    float out, x0, x1, x2, x3, y0, y1, y2, y3;
    if (EXACT) {
        out = __fmul_rn(s.p3, c.gain);
        x0 = __fsub_rn(__fmul_rn(c.f1, s.p0), __fmul_rn(c.f2, s.q0));
        y0 = __fadd_rn(__fmul_rn(c.f2, s.p0), __fmul_rn(c.f1, s.q0));
        x1 = __fsub_rn(__fmul_rn(c.f1, s.p1), __fmul_rn(c.f2, s.q1));
        y1 = __fadd_rn(__fmul_rn(c.f2, s.p1), __fmul_rn(c.f1, s.q1));
        x2 = __fsub_rn(__fmul_rn(c.f1, s.p2), __fmul_rn(c.f2, s.q2));
        y2 = __fadd_rn(__fmul_rn(c.f2, s.p2), __fmul_rn(c.f1, s.q2));
        x3 = __fsub_rn(__fmul_rn(c.f1, s.p3), __fmul_rn(c.f2, s.q3));
        y3 = __fadd_rn(__fmul_rn(c.f2, s.p3), __fmul_rn(c.f1, s.q3));
        s.p0 = __fadd_rn(__fmul_rn(in, c.f1), x0);
        s.q0 = __fadd_rn(__fmul_rn(in, c.f2), y0);
        s.p1 = __fadd_rn(s.p0, x1);
        s.q1 = __fadd_rn(s.q0, y1);
        s.p2 = __fadd_rn(__fadd_rn(s.p1, x1), x2);
        s.q2 = __fadd_rn(__fadd_rn(s.q1, y1), y2);
        s.p3 = __fadd_rn(__fadd_rn(__fadd_rn(s.p2, x1), __fmul_rn(2.0f, x2)), x3);
        s.q3 = __fadd_rn(__fadd_rn(__fadd_rn(s.q2, y1), __fmul_rn(2.0f, y2)), y3);
    } else {
        out = s.p3 * c.gain;
        x0 = c.f1 * s.p0 - c.f2 * s.q0; y0 = c.f2 * s.p0 + c.f1 * s.q0;
        x1 = c.f1 * s.p1 - c.f2 * s.q1; y1 = c.f2 * s.p1 + c.f1 * s.q1;
        x2 = c.f1 * s.p2 - c.f2 * s.q2; y2 = c.f2 * s.p2 + c.f1 * s.q2;
        x3 = c.f1 * s.p3 - c.f2 * s.q3; y3 = c.f2 * s.p3 + c.f1 * s.q3;
        s.p0 = in * c.f1 + x0;
        s.q0 = in * c.f2 + y0;
        s.p1 = s.p0 + x1;
        s.q1 = s.q0 + y1;
        s.p2 = s.p1 + x1 + x2;
        s.q2 = s.q1 + y1 + y2;
        s.p3 = s.p2 + x1 + 2.0f * x2 + x3;
        s.q3 = s.q2 + y1 + 2.0f * y2 + y3;
    }
    return out;
}

// Default-mode step: same recurrence with FMA contraction, returns p[3] BEFORE the update without the gain
// (callers fold gain / midEar into one scale factor).
__device__ __forceinline__ float gt_step_raw(GtState &s, const GtCoef &c, float in)
{
    const float out = s.p3;
    const float x0 = c.f1 * s.p0 - c.f2 * s.q0, y0 = c.f2 * s.p0 + c.f1 * s.q0;
    const float x1 = c.f1 * s.p1 - c.f2 * s.q1, y1 = c.f2 * s.p1 + c.f1 * s.q1;
    const float x2 = c.f1 * s.p2 - c.f2 * s.q2, y2 = c.f2 * s.p2 + c.f1 * s.q2;
    const float x3 = c.f1 * s.p3 - c.f2 * s.q3, y3 = c.f2 * s.p3 + c.f1 * s.q3;
    s.p0 = fmaf(in, c.f1, x0);
    s.q0 = fmaf(in, c.f2, y0);
    s.p1 = s.p0 + x1;
    s.q1 = s.q0 + y1;
    const float ax = x1 + x2, ay = y1 + y2;
    s.p2 = s.p1 + ax;
    s.q2 = s.q1 + ay;
    s.p3 = (s.p2 + ax) + (x2 + x3);
    s.q3 = (s.q2 + ay) + (y2 + y3);
    return out;
}

// Rotating-frame form of the same recurrence for the default mode.  With c_k = p_k + i q_k and the pole
// z = rho e^{i theta}, substitute c_k[n] = e^{i theta n} d_k[n]: the four complex rotations z * c_k (16 real
// operations per sample) become four real scalings by rho that fold into the binomial sums as FMAs, at the
// price of modulating the input and demodulating the output with the phasor e^{-+ i theta n}:
//   d0' = rho (x e^{-i theta n} + d0);  d1' = d0' + rho d1;  d2' = d1' + rho d1 + rho d2;
//   d3' = d2' + rho d1 + 2 rho d2 + rho d3;  out[n] = Re(e^{i theta n} d3[n])
// 22 FP32 operations per sample instead of 31.  The phasor is advanced by one complex multiply per sample
// and re-seeded from a double-precision angle at every tile start (gt_rot_seed), so it cannot drift.
struct GtRot { float rho, rho2, rc, rs; double turns; };
struct GtStateR { float d0r, d0i, d1r, d1i, d2r, d2i, d3r, d3i, pr, pi; };

__device__ __forceinline__ void gtr_reset(GtStateR &s)
{
    s.d0r = s.d0i = s.d1r = s.d1i = s.d2r = s.d2i = s.d3r = s.d3i = 0.0f;
    s.pr = 1.0f;
    s.pi = 0.0f;
}

// phasor (cos, sin)(theta * n) for step index n since the filter was started from rest
__device__ __forceinline__ void gt_rot_seed(GtStateR &s, const GtRot &k, int n)
{
    const double x = k.turns * (double)n;
    const float frac = (float)(x - floor(x));
    sincospif(2.0f * frac, &s.pi, &s.pr);
}

// returns d3 demodulated BEFORE the update (the one-sample delay of the reference), without the gain
__device__ __forceinline__ float gt_step_rot(GtStateR &s, const GtRot &k, float in)
{
    const float out = s.pr * s.d3r - s.pi * s.d3i;
    const float n0r = k.rho * fmaf(in, s.pr, s.d0r);
    const float n0i = k.rho * fmaf(-in, s.pi, s.d0i);
    const float n1r = fmaf(k.rho, s.d1r, n0r), n1i = fmaf(k.rho, s.d1i, n0i);
    const float n2r = fmaf(k.rho, s.d2r, fmaf(k.rho, s.d1r, n1r));
    const float n2i = fmaf(k.rho, s.d2i, fmaf(k.rho, s.d1i, n1i));
    const float n3r = fmaf(k.rho, s.d3r, fmaf(k.rho2, s.d2r, fmaf(k.rho, s.d1r, n2r)));
    const float n3i = fmaf(k.rho, s.d3i, fmaf(k.rho2, s.d2i, fmaf(k.rho, s.d1i, n2i)));
    s.d0r = n0r; s.d0i = n0i; s.d1r = n1r; s.d1i = n1i; s.d2r = n2r; s.d2i = n2i; s.d3r = n3r; s.d3i = n3i;
    const float pr = s.pr * k.rc - s.pi * k.rs;
    s.pi = s.pr * k.rs + s.pi * k.rc;
    s.pr = pr;
    return out;
}

// hairCell, resyth_64sub_IBM/cpp/extractwav.cpp:213-258.  hc[] = {ymdt,xdt,ydt,lplusrdt,rdt,gdt,hdt,c0,q0,w0}.
struct HcState { float c, q, w; };
__device__ __forceinline__ void hc_reset(HcState &s, const float *hc) { s.c = hc[7]; s.q = hc[8]; s.w = hc[9]; }

__device__ __forceinline__ float hc_step(HcState &s, const float *hc, float in)
{
    // kt is evaluated in double in the reference (MED_A, MED_B are double literals)
    const double t = (double)in + 3.0;
    const float kt = (t > 0.0) ? (float)((double)hc[5] * t / (t + 300.0)) : 0.0f;
    const float replenish = (s.q < 1.0f) ? __fsub_rn(hc[0], __fmul_rn(hc[2], s.q)) : 0.0f;
    const float eject = __fmul_rn(kt, s.q);
    const float loss = __fmul_rn(hc[3], s.c);
    const float reuptake = __fmul_rn(hc[4], s.c);
    const float reprocess = __fmul_rn(hc[1], s.w);
    s.q = __fadd_rn(__fsub_rn(__fadd_rn(s.q, replenish), eject), reprocess);
    if (s.q < 0.0f) s.q = 0.0f;
    s.c = __fsub_rn(__fadd_rn(s.c, eject), loss);
    if (s.c < 0.0f) s.c = 0.0f;
    s.w = __fsub_rn(__fadd_rn(s.w, reuptake), reprocess);
    if (s.w < 0.0f) s.w = 0.0f;
    return __fmul_rn(hc[6], s.c);
}

// (short) of a float as x86 does it: cvttss2si to int32, keep the low 16 bits.
// Out of int32 range (or NaN) cvttss2si returns the "integer indefinite" 0x80000000, whose low 16 bits are 0;
// __float2int_rz would saturate positive overflow to 0x7fffffff (low 16 bits -1) instead.
__device__ __forceinline__ short cast_i16(float v)
{
    if (!(v > -2147483904.0f && v < 2147483648.0f)) return (short)0;
    return (short)__float2int_rz(v);
}

}  // namespace se
