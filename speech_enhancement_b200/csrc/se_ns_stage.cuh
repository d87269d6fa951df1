// One Wiener-filter stage of one frame, executed by one warp.  Included by se_ns.cu.
//
// Follows the body of the fstage loop of DoNoiseSup (etsi/cpp/NoiseSup.c:1144-1367; plugin twin
// aurora_etsi/NoiseSup.cpp:1218-1374): DoSigWindowing -> rfft -> FFTtoPSD -> PSDMean -> VAD ->
// FilterCalc -> SpeechQVar -> DoMelFB|DoGamma -> SpeechQSpec/Mel -> DoGainFact -> DoMelIDCT ->
// DoFilterWindowing -> ApplyWF.
#pragma once

namespace SE_NSPACE {

struct IdctRegs { float c[9]; float firwin; };

#ifdef SE_NS_CONST_SMEM
__device__ __forceinline__ void lane_consts_store(float *tw_tab, float *ic_tab, const FftTw &t, const IdctRegs &ic, int lane)
{
    float v[kTwFloats];
    fft_tw_to_array(t, v);
#pragma unroll
    for (int i = 0; i < kTwFloats; i++) tw_tab[i * 32 + lane] = v[i];
#pragma unroll
    for (int i = 0; i < 9; i++) ic_tab[i * 32 + lane] = ic.c[i];
    ic_tab[9 * 32 + lane] = ic.firwin;
}
__device__ __forceinline__ FftTw lane_tw_load(const float *tab, int lane)
{
    return fft_tw_from_table(tab + lane);
}
#define SE_TW_LOCAL(bm, tw, lane) const FftTw tw##_l = lane_tw_load((bm).lane_tw, lane)
#define SE_TW(tw) tw##_l
#else
#define SE_TW_LOCAL(bm, tw, lane)
#define SE_TW(tw) tw
#endif


template <int RB> __device__ __forceinline__ int wrap(int pos)
{
    if (pos >= RB) pos -= RB;
    if (pos < 0) pos += RB;
    return pos;
}

#ifndef SE_NS_EXACT
// Windowed frame -> PSD bins owned by this lane (psd[r] for bin lane+32r).
template <int PROF, int RB = Prof<PROF>::RB>
__device__ __forceinline__ void frame_psd(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const FftTw &tw, int lane,
                                          const float *ring, int wpos, float (&psd)[Prof<PROF>::NB])
{
    using P = Prof<PROF>;
    if constexpr (PROF == 8) {
        float2 v[1][4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int i2 = 2 * (lane + 32 * j);
            v[0][j] = make_float2(0.0f, 0.0f);
            if (i2 < P::FL) {
                const int pos = wrap<RB>(wpos + i2);
                const float2 x = *reinterpret_cast<const float2 *>(ring + pos);
                const float2 w = *reinterpret_cast<const float2 *>(bm.win + i2);
                v[0][j] = make_float2(x.x * w.x, x.y * w.y);
            }
        }
        {
            SE_TW_LOCAL(bm, tw, lane);
            float2 *const zs[1] = {wm.Z};
            fft128_spectrum<1>(v, zs, SE_TW(tw), lane);
        }
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
        float2 v[2][4];       // [0]: z = x[0::4] + i x[2::4] (-> A), [1]: w = x[1::4] + i x[3::4] (-> B, C)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int i4 = 4 * (lane + 32 * j);
            v[0][j] = make_float2(0.0f, 0.0f);
            v[1][j] = make_float2(0.0f, 0.0f);
            if (i4 < P::FL) {
                const int pos = wrap<RB>(wpos + i4);
                const float4 x = *reinterpret_cast<const float4 *>(ring + pos);
                const float4 w = *reinterpret_cast<const float4 *>(bm.win + i4);
                v[0][j] = make_float2(x.x * w.x, x.z * w.z);
                v[1][j] = make_float2(x.y * w.y, x.w * w.w);
            }
        }
        {
            SE_TW_LOCAL(bm, tw, lane);
            float2 *const zs[2] = {wm.Z, wm.Z2};
            fft128_spectrum<2>(v, zs, SE_TW(tw), lane);
        }
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
            if (b < 128) {
                const float2 e = *reinterpret_cast<const float2 *>(raw + 2 * b);    // one 8-byte load: half the wavefronts
                psd[r] = (e.x + e.y) * 0.5f;
            } else if (b == 128) psd[r] = raw[256];
        }
    }
}

#endif  // !SE_NS_EXACT

struct StageFlags { int var, spec, mel; };

// Gamma bank of the 16 kHz plugin profile in tensor memory (SE_NS_BANK_TMEM): column 4 j + r of lane l (in every lane
// quadrant) = weight of band j for bin l + 32 r -- the float4 the shared-memory version reads at bank4[j * 32 + l].
// Called by every thread of the CTA, which must have at least four warps; pairs with bank_tmem_release at kernel end.
constexpr uint32_t kBankTmemCols = 128;
template <int PROF> __device__ __forceinline__ void bank_tmem_setup(BlockMem<PROF> &bm, const float *bank_global)
{
#if defined(SE_NS_BANK_TMEM) && !defined(SE_NS_EXACT)
    if constexpr (PROF == 16) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (warp == 0) se_tmem::alloc(&bm.tmem_base, kBankTmemCols);
        se_tmem::fence_before_sync();
        __syncthreads();
        se_tmem::fence_after_sync();
        if (warp < 4) {
            const float4 *g = reinterpret_cast<const float4 *>(bank_global);
            const uint32_t taddr = se_tmem::warp_addr(bm.tmem_base, warp, 0);
            for (int j = 0; j < kMel; j++) {
                const float4 c = g[j * 32 + lane];
                se_tmem::st4(taddr + 4 * j, c.x, c.y, c.z, c.w);
            }
            se_tmem::wait_st();
        }
        se_tmem::fence_before_sync();
        __syncthreads();
        se_tmem::fence_after_sync();
    }
#endif
}
template <int PROF> __device__ __forceinline__ void bank_tmem_release(BlockMem<PROF> &bm)
{
#if defined(SE_NS_BANK_TMEM) && !defined(SE_NS_EXACT)
    if constexpr (PROF == 16) {
        se_tmem::fence_before_sync();
        __syncthreads();
        se_tmem::fence_after_sync();
        if ((threadIdx.x >> 5) == 0) se_tmem::dealloc(bm.tmem_base, kBankTmemCols);
    }
#endif
}

// Sample rings: 4 shifts of history per stage, stored with 8 guard samples on either side (a copy of the
// other end) so the 17-tap filter reads its S+16 window with plain contiguous addresses.
// ring points at the core: ring[-8..0) mirrors ring[RB-8..RB), ring[RB..RB+8) mirrors ring[0..8).
template <int PROF> __device__ __forceinline__ void ring_store(float *ring, int slot, int i, float v)
{
    using P = Prof<PROF>;
    ring[slot * P::S + i] = v;
    if (slot == 3 && i >= P::S - 8) ring[i - P::S] = v;
    if (slot == 0 && i < 8) ring[P::RB + i] = v;
}

// Per-lane registers of the 25x9 IDCT (DoMelIDCT restricted to the 9 taps DoFilterWindowing keeps):
// lane L = t + 9 g (t < 9 taps, g < 3 groups) owns bands f = 9 g + i, i < 9.

__device__ __forceinline__ void idct_regs_init(IdctRegs &ic, const float *idctT /*[25][16]*/, const float *firwin9, int lane)
{
    const int t = lane % 9, g = lane / 9;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const int f = 9 * g + i;
        ic.c[i] = (g < 3 && f < kMel) ? idctT[f * 16 + t] : 0.0f;
    }
    ic.firwin = firwin9[t];
}

// Sparse filter banks of the 8 kHz profile held as per-lane tap registers (weights zero-padded, so the
// loops need no predicate): the 25 NoiseSup mel triangles (<= 10 taps, lane = band) and the 23 CompCeps
// triangles (<= 22 taps: lane j < 23 takes taps 0..11 of band j, lane j + 9 (23..31) takes taps 12..23 of
// bands 14..22; one shuffle folds the halves).
constexpr int kMelTaps = 10, kTriTaps = 12;
struct BankRegs {
    int mel_st, tri_st;
    float mel[kMelTaps], tri[kTriTaps];
};

__device__ __forceinline__ void bank_regs_init(BankRegs &br, const NsDevTables &tb, const CcDevTables &cc, int lane)
{
    // once per kernel, straight from the global tables (tb.bankT: [tap][32 bands], cc.triT likewise)
    br.mel_st = tb.bank_start[lane];
    const int ml = tb.bank_len[lane];
#pragma unroll
    for (int i = 0; i < kMelTaps; i++) br.mel[i] = (i < ml) ? tb.bankT[i * 32 + lane] : 0.0f;
    const int band = lane < 23 ? lane : lane - 9, t0 = lane < 23 ? 0 : kTriTaps;
    br.tri_st = cc.tri_start[band] + t0;
    const int tl = cc.tri_len[band];
#pragma unroll
    for (int i = 0; i < kTriTaps; i++) br.tri[i] = (t0 + i < tl) ? cc.triT[(t0 + i) * 32 + band] : 0.0f;
}
#ifdef SE_NS_CONST_SMEM
__device__ __forceinline__ void bank_regs_store(float *tab /*[22][32]*/, const BankRegs &br, int lane)
{
#pragma unroll
    for (int i = 0; i < kMelTaps; i++) tab[i * 32 + lane] = br.mel[i];
#pragma unroll
    for (int i = 0; i < kTriTaps; i++) tab[(kMelTaps + i) * 32 + lane] = br.tri[i];
}
#define SE_BR_MEL(bm, br, i, lane) (bm).lane_br[(i) * 32 + (lane)]
#define SE_BR_TRI(bm, br, i, lane) (bm).lane_br[(kMelTaps + (i)) * 32 + (lane)]
#else
#define SE_BR_MEL(bm, br, i, lane) (br).mel[i]
#define SE_BR_TRI(bm, br, i, lane) (br).tri[i]
#endif

#ifndef SE_NS_EXACT
// One Wiener-filter stage after its PSD is known (fstage is warp-uniform at run time in the batch kernel: both stages
// share one copy of the code, which halves the kernel's instruction footprint).  psd[r] = PSD bin lane + 32 r of the
// frame (frame_psd).  Split from the PSD so that the pipelined single-utterance kernel (se_ns_pipe.cuh) can compute
// the two on different warps.
// SPLIT (the pipelined kernel's two "recursion" warps): the frame energy comes in through `sp`, the frame-dropping
// statistics (fd_stats) and the FIR (apply_fir) run on other warps -- `sp` takes what they need: the lane partials of
// the gain sums, the mel gains and the 9 filter taps.  Same expressions, same order, as the whole-stage form.
struct StageSplit { float frame_sq, wsum_part, wsq_part, mw, h[9]; };

// sum of squares of curFrame (the VAD's frame energy, stage 0)
template <int PROF> __device__ __forceinline__ float frame_energy(const float *ring, int cur_base, int lane)
{
    using P = Prof<PROF>;
    float frame_sq = 0.0f;
#pragma unroll
    for (int r = 0; r < P::NR; r++) {
        const int i = lane + 32 * r;
        if (i < P::S) { const float x = ring[cur_base + i]; frame_sq += x * x; }
    }
    return warp_sum(frame_sq);
}

// ApplyWF, NoiseSup.c:324-340: 17-tap symmetric FIR over curFrame with prv/next context
template <int PROF>
__device__ __forceinline__ void apply_fir(const float *ring, int cur_base, const float (&h)[9], int lane, float (&o)[Prof<PROF>::CH])
{
    using P = Prof<PROF>;
    float x[P::CH + 16];
    const int i0 = lane * P::CH;
    const float *xin = ring + cur_base + i0 - 8;        // guard samples make this window contiguous
#pragma unroll
    for (int d = 0; d < P::CH + 16; d++) x[d] = xin[d];
#pragma unroll
    for (int t = 0; t < P::CH; t++) {
        float acc = 0.0f;
#pragma unroll
        for (int j = -8; j <= 8; j++) acc += h[j < 0 ? -j : j] * x[t + 8 - j];
        o[t] = acc;
    }
}

// frame-dropping VAD statistics of stage 0 (NoiseSup.c:1268-1281, SpeechQVar / SpeechQSpec / SpeechQMel) from the lane
// partials of the gain sums and the mel gains (0 on lanes >= 25); s.nb_frame0 is the stage-0 frame counter
template <int PROF>
__device__ __forceinline__ void fd_stats(WarpMem<PROF> &wm, Scal &s, float wsum, float wsq, float mw, StageFlags &fl)
{
    using P = Prof<PROF>;
    wsum = warp_sum(wsum);
    wsq = warp_sum(wsq);
    SE_COLD(cold_fd_load(s, wm.cold));
    fl.var = q_var(s, wsum, wsq, P::QVAR);
    const float te = warp_sum(mw);
    s.fd_spec_val = (float)((double)(te * te) - 3.0);
    fl.spec = q_spec(s);
    const float w1 = __shfl_sync(0xffffffffu, mw, 1), w2 = __shfl_sync(0xffffffffu, mw, 2),
                w3 = __shfl_sync(0xffffffffu, mw, 3);
    s.fd_mel_val1 = (float)((double)__fadd_rn(__fadd_rn(w1, w2), w3) / 3.0);
    fl.mel = q_mel(s);
    SE_COLD(cold_fd_store(wm.cold, s));
}

template <int PROF, bool SPLIT = false>
__device__ __forceinline__ void ns_stage_rest(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const IdctRegs &ic,
                                              const BankRegs &br, int lane, Scal &s, float (&noise)[Prof<PROF>::NB], float (&den)[Prof<PROF>::NB],
                                              float eps, const int fstage, const float (&psd)[Prof<PROF>::NB], float (&o)[Prof<PROF>::CH],
                                              StageFlags &fl, float *gains_row, float *spec_trace, StageSplit *sp = nullptr)
{
    using P = Prof<PROF>;
    float *ring = fstage == 0 ? wm.ring1 + 8 : wm.ring2 + 8;
    const int cnt = fstage == 0 ? s.nb_in1 : s.nb_in2;   // blocks written to this ring so far
    const int cur_base = ((cnt + 1) & 3) * P::S;         // buffer[S..2S)     = curFrame

    // VAD (stage 0 only looks at the signal)
    float frame_sq = 0.0f;
    if (fstage == 0) {
        if constexpr (SPLIT) frame_sq = sp->frame_sq;
        else frame_sq = frame_energy<PROF>(ring, cur_base, lane);
    }
    vad_ns(s, fstage, frame_sq);
    const int nbf = (int)(short)(fstage == 0 ? s.nb_frame0 : s.nb_frame1);   // Q2: read through int16

    // PSDMean (NoiseSup.c:289-303) + FilterCalc per owned bin
    float W[P::NB];
    float *prev = wm.prev + fstage * (P::NB * 32);
    float wsum = 0.0f, wsq = 0.0f, den_sum = 0.0f, noise_sum = 0.0f;
#pragma unroll
    for (int r = 0; r < P::NB; r++) {
        const int b = lane + 32 * r;
        W[r] = 0.0f;
        if (b < P::SPEC) {
            const float pm = (prev[r * 32 + lane] + psd[r]) * 0.5f;
            prev[r * 32 + lane] = psd[r];
            W[r] = filter_calc_bin(fstage, psd[r], pm, noise[r], den[r], nbf, s.flag_vad, eps);
            if (PROF == 8) wm.W[b] = W[r];
            if (r < P::NB - 1) { wsum += W[r]; wsq += W[r] * W[r]; }
            den_sum += den[r];
            noise_sum += noise[r];
            if (spec_trace) {
                // the SPECTRUM state as FilterCalc leaves it (NoiseSup.c:519-560): nSigSE, noiseSE, denSigSE, and
                // sqrt(PSDMeaned) (a local of DoNoiseSup in the reference) -- se_ns_spectra_batch
                spec_trace[(3 * fstage + 0) * P::SPEC + b] = fsqrt(psd[r]);
                spec_trace[(3 * fstage + 1) * P::SPEC + b] = noise[r];
                spec_trace[(3 * fstage + 2) * P::SPEC + b] = den[r];
                spec_trace[(6 + fstage) * P::SPEC + b] = fsqrt(pm);
            }
        }
    }
    if (fstage == 0) {
        if constexpr (SPLIT) {
            sp->wsum_part = wsum;
            sp->wsq_part = wsq;
        } else {
            wsum = warp_sum(wsum);
            wsq = warp_sum(wsq);
            SE_COLD(cold_fd_load(s, wm.cold));
            fl.var = q_var(s, wsum, wsq, P::QVAR);
            SE_COLD(cold_fd_store(wm.cold, s));
        }
    }

    // DoMelFB (etsi/cpp/MelProc.c:82-104) / DoGamma (aurora_etsi/MelProc.cpp:119-141, ignores StartingPoint)
    float mw = 0.0f;
    if constexpr (PROF == 8) {
        __syncwarp();
#pragma unroll
        for (int i = 0; i < kMelTaps; i++) mw = fmaf(wm.W[br.mel_st + i], SE_BR_MEL(bm, br, i, lane), mw);
    } else {
        // dense 25 x 128: every lane multiplies its own 4 bins (registers) into 25 partial sums, then a
        // butterfly reduce-scatter leaves band j's total on lane j (31 shuffles instead of 25 x 5)
        float p[32];
#if defined(SE_NS_BANK_TMEM) && !defined(SE_NS_EXACT)
        // the lane's 100 weights come from its row of tensor memory (bank_tmem_setup), two bands per load: the
        // shared-memory pipe, which bounds this kernel, no longer carries them
        const uint32_t taddr = se_tmem::warp_addr(bm.tmem_base, (int)(threadIdx.x >> 5), 0);
#pragma unroll
        for (int j = 0; j + 1 < kMel; j += 2) {
            uint32_t c[8];
            se_tmem::ld8(taddr + 4 * j, c);
            se_tmem::wait_ld8(c);
            p[j] = W[0] * __uint_as_float(c[0]) + W[1] * __uint_as_float(c[1]) + W[2] * __uint_as_float(c[2]) + W[3] * __uint_as_float(c[3]);
            p[j + 1] = W[0] * __uint_as_float(c[4]) + W[1] * __uint_as_float(c[5]) + W[2] * __uint_as_float(c[6]) + W[3] * __uint_as_float(c[7]);
        }
        {
            uint32_t c[4];
            se_tmem::ld4(taddr + 4 * (kMel - 1), c);
            se_tmem::wait_ld4(c);
            p[kMel - 1] = W[0] * __uint_as_float(c[0]) + W[1] * __uint_as_float(c[1]) + W[2] * __uint_as_float(c[2]) + W[3] * __uint_as_float(c[3]);
        }
#else
        const float4 *bank4 = reinterpret_cast<const float4 *>(bm.bankT);
#pragma unroll
        for (int j = 0; j < kMel; j++) {
            const float4 c = bank4[j * 32 + lane];
            p[j] = W[0] * c.x + W[1] * c.y + W[2] * c.z + W[3] * c.w;
        }
#endif
#pragma unroll
        for (int j = kMel; j < 32; j++) p[j] = 0.0f;
#pragma unroll
        for (int sp = 16; sp >= 1; sp >>= 1) {
            const bool up = (lane & sp) != 0;
#pragma unroll
            for (int i = 0; i < sp; i++) {
                const float send = up ? p[i] : p[i + sp];
                const float keep = up ? p[i + sp] : p[i];
                p[i] = keep + __shfl_xor_sync(0xffffffffu, send, sp);
            }
        }
        mw = p[0];
    }
    if (lane >= kMel) mw = 0.0f;

    if (SPLIT && fstage == 0) {
        sp->mw = mw;
        // DoGainFact stage 0, NoiseSup.c:593-599
        SE_COLD(cold_gf_load(s, wm.cold));
        s.den_en0 = s.den_en1;
        s.den_en1 = s.den_en2;
        s.den_en2 = warp_sum(den_sum);
        SE_COLD(cold_gf_store(wm.cold, s));
    } else if (fstage == 0) {
        // frame-dropping VAD statistics, NoiseSup.c:1268-1281
        const float te = warp_sum(mw);
        SE_COLD(cold_fd_load(s, wm.cold));
        SE_COLD(cold_gf_load(s, wm.cold));
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
        SE_COLD(cold_fd_store(wm.cold, s));
        SE_COLD(cold_gf_store(wm.cold, s));
    } else {
        SE_COLD(cold_gf_load(s, wm.cold));
        gain_fact_stage2<PROF>(s, warp_sum(noise_sum));
        mw = (float)((double)__fmul_rn(s.alfa_gf, mw) + (1.0 - (double)s.alfa_gf) * 1.0);
        SE_COLD(cold_gf_store(wm.cold, s));
        if (gains_row && lane < kMel) gains_row[lane] = mw;
    }

    // DoMelIDCT (MelProc.c:357-378): only taps 0..8 survive DoFilterWindowing (NoiseSup.c:660-669).
    // lane t + 9 g accumulates bands 9 g .. 9 g + 8 of tap t; three partials are folded with two shuffles.
    float h[9];
    {
        const int g9 = (lane / 9) * 9;
        float acc = 0.0f;
#pragma unroll
#ifdef SE_NS_CONST_SMEM
        for (int i = 0; i < 9; i++) acc += __shfl_sync(0xffffffffu, mw, (g9 + i) & 31) * bm.lane_ic[i * 32 + lane];
        acc += __shfl_down_sync(0xffffffffu, acc, 9) + __shfl_down_sync(0xffffffffu, acc, 18);
        acc *= bm.lane_ic[9 * 32 + lane];
#else
        for (int i = 0; i < 9; i++) acc += __shfl_sync(0xffffffffu, mw, (g9 + i) & 31) * ic.c[i];
        acc += __shfl_down_sync(0xffffffffu, acc, 9) + __shfl_down_sync(0xffffffffu, acc, 18);
        acc *= ic.firwin;
#endif
#pragma unroll
        for (int t = 0; t < 9; t++) h[t] = __shfl_sync(0xffffffffu, acc, t);
    }

    if constexpr (SPLIT) {
#pragma unroll
        for (int t = 0; t < 9; t++) sp->h[t] = h[t];
        if (fstage == 0) s.nb_in2++;
        else s.nb_out2++;
        __syncwarp();
        return;
    }
    apply_fir<PROF>(ring, cur_base, h, lane, o);
    const int i0 = lane * P::CH;

    if (fstage == 0) {
        const int slot = s.nb_in2 & 3;
#pragma unroll
        for (int t = 0; t < P::CH; t++)
            if (i0 + t < P::S) ring_store<PROF>(wm.ring2 + 8, slot, i0 + t, o[t]);
        s.nb_in2++;
    } else {
        s.nb_out2++;
    }
    __syncwarp();
}

// One whole stage on one warp: windowed frame -> PSD -> the rest.
template <int PROF>
__device__ __forceinline__ void ns_stage(WarpMem<PROF> &wm, const BlockMem<PROF> &bm, const FftTw &tw, const IdctRegs &ic,
                                         const BankRegs &br, int lane, Scal &s, float (&noise)[Prof<PROF>::NB], float (&den)[Prof<PROF>::NB],
                                         float eps, const int fstage, float (&o)[Prof<PROF>::CH], StageFlags &fl,
                                         float *gains_row, float *spec_trace, float *psd_keep = nullptr)
{
    using P = Prof<PROF>;
    const float *ring = fstage == 0 ? wm.ring1 + 8 : wm.ring2 + 8;
    const int cnt = fstage == 0 ? s.nb_in1 : s.nb_in2;
    float psd[P::NB];
    frame_psd<PROF>(wm, bm, tw, lane, ring, (cnt & 3) * P::S + P::ANA, psd);   // buffer[0..S) = prvFrame = slot cnt & 3
    if (psd_keep) { psd_keep[0] = psd[1]; psd_keep[1] = psd[2]; }      // 16 kHz mode: bins 32..64 of this frame
    ns_stage_rest<PROF>(wm, bm, ic, br, lane, s, noise, den, eps, fstage, psd, o, fl, gains_row, spec_trace);
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
// `ld(i)` returns frame sample i, -1 <= i < 200.  Returns this lane's coefficient (lanes 0..11 = c1..c12,
// lane 12 = c0, lane 13 = logE, 0 elsewhere).
// HP16 (16 kHz mode, CompCeps.c:464-479, 488-505, 518-531): three more log bands (70 % decoded from the low band's
// spectrum shape + 30 % spectrally-subtracted HP energies, both delayed 3 frames) after the 23 mel bands, DCT over
// 26 bands, HP energy added to logE.  dct_t: [bands][16] transposed DCT; code9 / hp3: lane-distributed oldest
// entries of the code and HP-band delay lines (lanes 0..8 / 0..2).
// First half (the pipelined kernel runs the halves on two warps): pre-emphasis, Hamming, FFT, power spectrum into wm.W[0..128];
// `e` = frame energy, d0..d2 = the 16 kHz mode's decoding-band sums (lane partials).
template <bool HP16 = false, class LD>
__device__ __forceinline__ void comp_ceps_front(WarpMem<8> &wm, const BlockMem<8> &bm, const FftTw &tw, int lane, LD ld,
                                                float &e, float &d0, float &d1, float &d2)
{
    e = 0.0f;
    float2 v[1][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i2 = 2 * (lane + 32 * j);
        v[0][j] = make_float2(0.0f, 0.0f);
        if (i2 < 200) {
            const float sm1 = ld(i2 - 1);
            const float s0 = ld(i2);
            const float s1 = ld(i2 + 1);
            e += s0 * s0 + s1 * s1;
            // pre-emphasis 0.9 (:427-429) and Hamming (:434)
            v[0][j] = make_float2((s0 - 0.9f * sm1) * bm.ham[i2], (s1 - 0.9f * s0) * bm.ham[i2 + 1]);
        }
    }
    e = warp_sum(e);
    {
        SE_TW_LOCAL(bm, tw, lane);
        float2 *const zs[1] = {wm.Z};
        fft128_spectrum<1>(v, zs, SE_TW(tw), lane);
    }
    __syncwarp();
    // power spectrum (:451-459), 129 bins, into wm.W
    d0 = d1 = d2 = 0.0f;                       // HP16: GetBandsForDecoding16k, bins 66..76 | 78..96 | 98..128
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int k = lane + 32 * r;
        const float2 xk = rsplit256(wm.Z, k, bm.twh[k]);
        const float pw = (k == 0) ? xk.x * xk.x : (xk.x * xk.x + xk.y * xk.y);
        wm.W[k] = pw;
        if constexpr (HP16) {
            if (k >= 66 && k < 77) d0 += pw;
            if (k >= 78 && k < 97) d1 += pw;
            if (k >= 98) d2 += pw;
        }
    }
    if (lane == 0) {
        const float2 xk = rsplit256(wm.Z, 128, bm.twh[128]);
        wm.W[128] = xk.x * xk.x;
        if constexpr (HP16) d2 += xk.x * xk.x;
    }
    __syncwarp();
}

// Second half: mel triangles over wm.W, log, DCT; returns this lane's coefficient.
template <bool HP16 = false>
__device__ __forceinline__ float comp_ceps_back(WarpMem<8> &wm, const BlockMem<8> &bm, const BankRegs &br, int lane, float e,
                                               float d0, float d1, float d2, const float *dct26T = nullptr, float code9 = 0.0f,
                                               float hp3 = 0.0f)
{
    constexpr int NBAND = HP16 ? 26 : 23;
    // 23 mel triangles (DoMelFB, MelProc.c:82-104), ln with floor (:509-513)
    float fb = 0.0f;
#pragma unroll
    for (int i = 0; i < kTriTaps; i++) fb = fmaf(wm.W[br.tri_st + i], SE_BR_TRI(bm, br, i, lane), fb);
    {
        const float hi = __shfl_down_sync(0xffffffffu, fb, 9);      // taps 12.. of bands 14..22 live on lanes 23..31
        if (lane >= 14 && lane < 23) fb += hi;
    }
    const float floor_fb = 4.5399929762484854e-05f;   // (float)exp(-10.0)
    if constexpr (HP16) {
        constexpr unsigned FULL = 0xffffffffu;
        d0 = warp_sum(d0) * 0.5f;
        d1 = warp_sum(d1) * 0.5f;
        d2 = warp_sum(d2) * 0.5f;
        d0 = (d0 > floor_fb) ? logf(d0) : -10.0f;
        d1 = (d1 > floor_fb) ? logf(d1) : -10.0f;
        d2 = (d2 > floor_fb) ? logf(d2) : -10.0f;
        // lanes 23..25 = HP band i: decoded estimate (weights 0.1 / 0.2 / 0.7) and the subtracted energy x 1.9
        const int i3 = 3 * (lane - 23);
        const float c0_ = __shfl_sync(FULL, code9, i3 & 31), c1_ = __shfl_sync(FULL, code9, (i3 + 1) & 31),
                    c2_ = __shfl_sync(FULL, code9, (i3 + 2) & 31);
        const float hp_e = __shfl_sync(FULL, hp3, (lane - 23) & 31);
        const float corr = 0.6418538689613342f;           // (float)log(1.0 + 0.90) == (float)log(1.0 + (double)0.9f)
        float coded = __fmul_rn(0.1f, d0 - c0_);
        coded = __fadd_rn(coded, __fmul_rn(0.2f, d1 - c1_));
        coded = __fadd_rn(coded, __fmul_rn(0.7f, d2 - c2_));
        coded += corr;
        if (lane >= 23 && lane < 26) fb = 1.9f * hp_e;
        fb = (fb < floor_fb) ? -10.0f : logf(fb);
        if (lane >= 23 && lane < 26) fb = 0.7f * coded + 0.3f * fb;     // MergeSSandCoded
        if (lane >= 26) fb = 0.0f;
        const float b22 = __shfl_sync(FULL, fb, 22), b23 = __shfl_sync(FULL, fb, 23);
        const float avg = 0.5f * b22 + 0.5f * b23;
        if (lane == 22 || lane == 23) fb = 0.6f * fb + 0.4f * avg;
        const float ex = expf(fb - corr);                 // CorrectEnergy
        e += __fadd_rn(__fadd_rn(__shfl_sync(FULL, ex, 23), __shfl_sync(FULL, ex, 24)), __shfl_sync(FULL, ex, 25));
    } else {
        fb = (fb < floor_fb) ? -10.0f : logf(fb);
        if (lane >= 23) fb = 0.0f;
    }
    const float c0 = warp_sum(fb);
    wm.melw[lane] = fb;
    __syncwarp();
    float out = 0.0f;
    if (lane < 12) {
        const float *dt = HP16 ? dct26T : bm.dctT;
#pragma unroll
        for (int j = 0; j < NBAND; j++) out += wm.melw[j] * dt[j * 16 + lane];
    } else if (lane == 12) {
        out = c0;
    } else if (lane == 13) {
        const float floor_e = 1.9287498479639178e-22f;   // (float)exp(-50.0)
        out = (e < floor_e) ? -50.0f : logf(e);
    }
    __syncwarp();
    return out;
}

template <bool HP16 = false, class LD>
__device__ __forceinline__ float comp_ceps_core(WarpMem<8> &wm, const BlockMem<8> &bm, const FftTw &tw, const BankRegs &br, int lane, LD ld,
                                                const float *dct26T = nullptr, float code9 = 0.0f, float hp3 = 0.0f)
{
    float e, d0, d1, d2;
    comp_ceps_front<HP16>(wm, bm, tw, lane, ld, e, d0, d1, d2);
    return comp_ceps_back<HP16>(wm, bm, br, lane, e, d0, d1, d2, dct26T, code9, hp3);
}

__device__ __forceinline__ void comp_ceps(WarpMem<8> &wm, const BlockMem<8> &bm, const FftTw &tw, const BankRegs &br, int lane, int base,
                                          float *row)
{
    using P = Prof<8>;
    const float *d = wm.ring3;
    const float out = comp_ceps_core(wm, bm, tw, br, lane, [&](int i) { return d[wrap<P::RB>(base + i)]; });
    if (lane < 14) row[lane] = out;
}

#endif  // !SE_NS_EXACT

}  // namespace SE_NSPACE

#ifdef SE_NS_EXACT
#include "se_ns_exact.cuh"
#endif
