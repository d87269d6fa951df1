// Back half of the ETSI feature chain, fused behind NoiseSup in the P8 stream kernel (FEAT variant):
// waveform processing -> CompCeps -> blind equalisation -> frame-dropping VAD, i.e. the block of
// DoAdvProcess the reference ships commented out (etsi/cpp/ParmInterface.c:274-311).
//   DoWaveProc  etsi/cpp/WaveProc.c:386-453 (TeagerEng :173, GetTeagerFilter :204, GetMaximaPositions :74)
//   DoPostProc  etsi/cpp/PostProc.c:117-150
//   DoVADProc   etsi/cpp/VAD.c:214-330
// One warp per utterance, as everywhere in this kernel; included by se_ns.cu.
#pragma once

namespace SE_NSPACE {

constexpr int kFeatCols = 15;    // c1..c12, c0, logE, VAD flag (FeatureBuffer[0..NUM_CEP_COEFF+1])

struct __align__(16) FeatMem {
    float frame[208];            // [0] = sample -1 (pre-emphasis history), [1..200] = the frame
    int q[208];                  // rounded Teager energy with 4 replicated samples either side
    int sm[200];                 // 9-point sums of q
    float vad[7 * 16];           // the VAD's 7-frame delay line (OutBuffer)
};

// State of the back half that persists across frames (warp-uniform except lms: lane i < 12 = weightLMS[i]).
struct FeatState {
    float lms;
    int focus, hang_over, h_count, v_count, speech_bits, rows;
};

__device__ __forceinline__ void feat_state_init(FeatState &f, FeatMem &fm, int lane)
{
    f.lms = 0.0f;
    f.focus = 0;
    f.hang_over = 23;            // HANGOVERV, VAD.c:113
    f.h_count = f.v_count = f.speech_bits = f.rows = 0;
    for (int i = lane; i < 7 * 16; i += 32) fm.vad[i] = 0.0f;       // calloc'ed OutBuffer, VAD.c:150
}

// DoWaveProc on fm.frame[1..200], in place.  Lane l owns the 7 consecutive samples [7l, 7l+7) (lanes 0..28),
// so Teager energy, the 9-point sums and the window all come from registers; only the rounded energies and
// their sums pass through shared memory (stride-7 words: conflict-free).  Warp-wide maxima use redux.sync.
__device__ __forceinline__ void wave_proc_warp(FeatMem &fm, int lane)
{
    constexpr int L = 200, T = 7;
    constexpr unsigned FULL = 0xffffffffu;
    float *d = fm.frame + 1;
    const int s0 = lane * T;
    float x[T + 2];                               // x[t] = d[s0 - 1 + t], frame edges replicated
#pragma unroll
    for (int t = 0; t < T + 2; t++) x[t] = d[min(max(s0 - 1 + t, 0), L - 1)];
    // low-energy check (:400-404).  The reference sums sequentially; a frame whose energy sits within
    // float rounding of 100.0 may take the other branch here.
    float e = 0.0f;
#ifdef SE_NS_EXACT
    for (int i = 0; i < L; i++) e += d[i] * d[i];        // exact mode: the reference's in-order sum (every lane)
#else
#pragma unroll
    for (int t = 0; t < T; t++)
        if (s0 + t < L) e = fmaf(x[t + 1], x[t + 1], e);
    e = warp_sum(e);
#endif
    if (!(e >= 100.0f)) return;

    // Teager energy, scaled by 0.25 and rounded to int (:173-184, :222-227)
#pragma unroll
    for (int t = 0; t < T; t++) {
        const int i = s0 + t;
        if (i < L) {
            const float tg = fabsf(__fsub_rn(__fmul_rn(x[t + 1], x[t + 1]), __fmul_rn(x[t], x[t + 2])));
            const int qi = __double2int_rd((double)tg * 0.25 + 0.5);
            fm.q[4 + i] = qi;
            if (i == 0) fm.q[0] = fm.q[1] = fm.q[2] = fm.q[3] = qi;
            if (i == L - 1) fm.q[204] = fm.q[205] = fm.q[206] = fm.q[207] = qi;
        }
    }
    __syncwarp();
    // 9-point sums as a sliding window (:228-254), wrap-around arithmetic like the reference's int32 sum
    int gv = 0, gi = 0;
    {
        unsigned qq[T + 8];
#pragma unroll
        for (int t = 0; t < T + 8; t++) qq[t] = (unsigned)fm.q[min(s0 + t, 207)];
        unsigned acc = 0;
#pragma unroll
        for (int k = 0; k < 9; k++) acc += qq[k];
#pragma unroll
        for (int t = 0; t < T; t++) {
            if (t > 0) acc += qq[t + 8] - qq[t - 1];
            const int v = (int)acc;
            if (s0 + t < L) {
                fm.sm[s0 + t] = v;
                if (v > gv) { gv = v; gi = s0 + t; }      // first strict maximum of this lane's samples
            }
        }
    }
    __syncwarp();
    // global maximum: largest value, smallest index, must be > 0 (:86-93)
    const int top = __reduce_max_sync(FULL, gv);
    int n_pos = 0, my_pos = 0;
    if (top > 0) {
        const int g = __reduce_min_sync(FULL, gv == top ? gi : L);
        // neighbours 25..79 samples away, farthest wins ties, only non-negative sums qualify (:97-135)
        int left_reg = 0, right_reg = 0, nl = 0, nr = 0;
#pragma unroll 1
        for (int dir = -1; dir <= 1; dir += 2) {
            int m = g;
            while (dir < 0 ? (m - 25 > 0) : (m + 25 < L)) {
                int bv = -1, bt = -1;
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int dd = 25 + lane + 32 * h;
                    const int p = m + dir * dd;
                    if (dd < 80 && p >= 0 && p < L) {
                        const int v = fm.sm[p];
                        if (v >= 0 && v >= bv) { bv = v; bt = dd; }
                    }
                }
                const int best = __reduce_max_sync(FULL, bv);
                if (best < 0) break;
                m += dir * __reduce_max_sync(FULL, bv == best ? bt : -1);
                if (dir < 0) { if (lane == nl) left_reg = m; nl++; }
                else { if (lane == nr) right_reg = m; nr++; }
            }
        }
        n_pos = nl + 1 + nr;
        // ascending list on lanes 0..n_pos-1 (sort_it, :44-66)
        const int from_left = __shfl_sync(FULL, left_reg, (nl - 1 - lane) & 31);
        const int from_right = __shfl_sync(FULL, right_reg, (lane - nl - 1) & 31);
        my_pos = lane < nl ? from_left : (lane == nl ? g : from_right);
    }
    // window (:259-289): 0.4 everywhere, 0.6 over the first 80 % of each inter-peak gap from 4 before the peak.
    // bit t of `hib` = window of sample s0+t is high (t = 0..7: own samples + the next one)
    const float lo = (float)((1.0f - 0.2f) / 2.0), hi = (float)((1.0f + 0.2f) / 2.0);
    const int nxt = __shfl_down_sync(FULL, my_pos, 1), prv = __shfl_up_sync(FULL, my_pos, 1);
    const int gap = (lane < n_pos - 1) ? nxt - my_pos : my_pos - prv;
    const int ia = my_pos - 4, ib = ia + (80 * gap + 99) / 100;
    unsigned hib = 0;
    if (n_pos > 1) {
        for (int k = 0; k < n_pos; k++) {
            const int A = max(__shfl_sync(FULL, ia, k) - s0, 0), B = min(__shfl_sync(FULL, ib, k) - s0, T + 1);
            if (B > A) hib |= ((1u << B) - 1u) & ~((1u << A) - 1u);
        }
    }
#pragma unroll
    for (int t = 0; t < T; t++) {
        const int i = s0 + t;
        if (i < L) {
            const int tn = (i == L - 1) ? t : t + 1;
            const float w0 = ((hib >> t) & 1u) ? hi : lo, w1 = ((hib >> tn) & 1u) ? hi : lo;
            d[i] = __fmul_rn(x[t + 1], __fadd_rn(w0, w1));                              // :430-447
        }
    }
    __syncwarp();
}

__constant__ float kLmsTarget[12] = {-6.618909f, 0.198269f, -0.740308f, 0.055132f, -0.227086f, 0.144280f,
                                     -0.112451f, -0.146940f, -0.327466f, 0.134571f, 0.027884f, -0.114905f};

// DoPostProc (Noc0 = 0: energy term = logE on lane 13).  c = this lane's coefficient.
__device__ __forceinline__ float post_proc_warp(float c, FeatState &f, int lane)
{
    const float lambda = 0.0087890625f;
    const float log_e = __shfl_sync(0xffffffffu, c, 13);
    float wp = __fdiv_rn(__fsub_rn(__fmul_rn(log_e, 64.0f), 211.0f), 64.0f);
    if (wp < 0.0f) wp = 0.0f;
    else if (wp > 1.0f) wp = lambda;
    else wp = __fmul_rn(wp, lambda);
    if (lane < 12) {
        const float dif = __fsub_rn(__fsub_rn(c, f.lms), kLmsTarget[lane]);
        c = __fsub_rn(c, f.lms);
        f.lms = __fadd_rn(f.lms, __fmul_rn(dif, wp));
    }
    return c;
}

// The release half of DoVADProc / DoVADFlush (VAD.c:256-303, :377-419): trigger logic over the 7 speech flags
// (bit 0 = the slot at Focus, scanned last), then the row one past Focus leaves the delay line.
__device__ __forceinline__ void vad_release_warp(float &c, FeatState &f, const FeatMem &fm, int lane, int frame_counter)
{
    const int x = f.speech_bits;
    const bool run3 = (x & (x >> 1) & (x >> 2)) != 0;                   // longest run of speech flags >= V_TRIGGER
    const bool run4 = (x & (x >> 1) & (x >> 2) & (x >> 3)) != 0;        // ... >= H_TRIGGER
    if (run4) {
        f.h_count = f.hang_over;
        if (frame_counter <= 35) f.hang_over = 50;                      // SANGOVER
    }
    if (f.h_count && !run3) f.h_count--;
    if (run3) f.v_count = 5;                                            // VANGOVER
    if (f.v_count && !run3) f.v_count--;
    const int oldest = f.focus == 6 ? 0 : f.focus + 1;
    c = (lane < 14) ? fm.vad[oldest * 16 + lane] : 0.0f;
    if (lane == 14) c = (f.v_count || f.h_count || run3) ? 1.0f : 0.0f;
}

// DoVADProc.  c = this lane's coefficient (lanes 0..13).  Returns true when a frame is released; then
// `c` holds this lane's column of the released row (lane 14 = VAD flag).
__device__ __forceinline__ bool vad_proc_warp(float &c, FeatState &f, FeatMem &fm, int lane, bool speech_found, int frame_counter)
{
    f.focus = f.focus == 6 ? 0 : f.focus + 1;
    if (lane < 14) fm.vad[f.focus * 16 + lane] = c;
    f.speech_bits = ((f.speech_bits << 1) | (speech_found ? 1 : 0)) & 0x7f;
    __syncwarp();
    if (frame_counter <= 10) return false;                              // BUFFER_SIZE + 3, VAD.c:256
    vad_release_warp(c, f, fm, lane, frame_counter);
    __syncwarp();
    return true;
}

// DoVADFlush (VAD.c:342-431), one step: Focus moves on without a write, so the scan order of the flags rotates.
// The row is rewritten only when the incremented FrameCounter is past 10; otherwise FeatureBuffer (fb) keeps what
// it held and the driver loop `while (FlushAdvProcess(...)) write` emits that stale row.
__device__ __forceinline__ void vad_flush_warp(float &fb, FeatState &f, const FeatMem &fm, int lane, int &frame_counter)
{
    f.focus = f.focus == 6 ? 0 : f.focus + 1;
    frame_counter++;
    if (frame_counter > 10) {
        f.speech_bits = ((f.speech_bits << 1) | (f.speech_bits >> 6)) & 0x7f;
        vad_release_warp(fb, f, fm, lane, frame_counter);
    }
}

}  // namespace SE_NSPACE
