// Internal declarations shared by the .cu translation units of libse_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/se_b200.h"
#include "se_tables.h"

namespace se {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define SE_CUDA(call)                                                              \
    do {                                                                           \
        cudaError_t e__ = (call);                                                  \
        if (e__ != cudaSuccess) return ::se::cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// Device-resident constant tables of one NoiseSup profile (global memory; each CTA stages what it
// needs into shared memory once).
// ---------------------------------------------------------------------------------------------
struct NsDevTables {
    const float *sig_window;   // [frame]
    const float *fir_window;   // [9]  : ir_window[8..16]  (DoFilterWindowing uses the upper half only)
    const float *bankT;        // P8: [BANK_MAXLEN][32] transposed, zero padded; P16: float4 [25][32] (lane's 4 bins)
    const int *bank_start;     // [32]
    const int *bank_len;       // [32]
    const float *idctT;        // [25][16] : idctT[f*16+t] = idct[t][f], t<9
    const float *tw_half;      // [129] float2 : exp(-2 pi i k / 256), k=0..128 (real-FFT split twiddles)
    float eps;
    // exact arithmetic mode (se_ctx_set_exact): the reference's rfft schedule for this profile (RfftPlan, se_tables.h)
    const unsigned *rf_ops;
    const int *rf_lvl;         // [32]: level_begin[0..15], tw_begin[0..15]
    const float *rf_tw;        // float4 per twiddle set
    int rf_m;
};

struct CcDevTables {
    const float *hamming;      // [200] full symmetric window
    const float *triT;         // [TRI_MAXLEN][32]
    const int *tri_start;      // [32]
    const int *tri_len;        // [32]
    const float *dctT;         // [23][16] : dctT[j*16+t] = dct[t][j], t<12
    // 16 kHz mode (se_ns_hp16k.cuh)
    const float *hpfb;         // [3][64]  : HP mel bands 1..3 as dense rows over PSD bins
    const float *dct26T;       // [26][16]
};

// Per-stream NoiseSup state kept in global memory between launches (persistent streams) —
// the GPU image of NoiseSupStructX.nsVar (etsi/cpp/NoiseSup.c:117-128).
constexpr int kNsMaxSpec = 129;
constexpr int kNsMaxRing = 640;
struct NsState {
    int nb_in1, nb_in2, nb_out2;      // BUFFERS counters
    int nb_frame0, nb_frame1;         // vadNS.nbFrame[2]
    int flag_vad, hang_over, nb_speech;
    int onset;                        // FEParamsX.NonZeroFrameOnset (P8)
    int n_ceps;                       // cepstral frames emitted so far (P8)
    int n_rows;                       // gain rows emitted so far (P16)
    int pad0;
    float mean_en, dc_x, dc_y;
    float den_en0, den_en1, den_en2, low_snr, alfa_gf;
    float fd_mel_mean, fd_var_mean, fd_acc_test, fd_spec_mean, fd_mel_val0, fd_mel_val1, fd_spec_val, fd_in_vadq;
    float noise1[kNsMaxSpec], den1[kNsMaxSpec], prev_psd1[kNsMaxSpec];
    float noise2[kNsMaxSpec], den2[kNsMaxSpec], prev_psd2[kNsMaxSpec];
    float ring1[kNsMaxRing], ring2[kNsMaxRing], ring3[kNsMaxRing];
};

// Arguments of one fused NoiseSup launch (one entry per utterance is derived from offsets).
struct NsLaunch {
    const void *samples;         // int16 or float
    int is_f32;
    const int64_t *offsets;      // device copy of the WHOLE batch's offsets
    int n_utts;                  // utterances in this launch: queue slots [utt_begin, utt_begin + n_utts)
    int utt_begin;
    const int32_t *order;        // optional: queue slot -> utterance index (longest first), device
    // a launch may cover a chunk of the batch whose data pointers address chunk-local staging buffers:
    // element index = global offset - base
    int64_t sample_base, block_base, ceps_base, trace_base;
    int64_t in_shift;            // added to the INPUT index only: `samples` may address the whole batch while the outputs are chunk-local
    // P8 outputs
    int16_t *den_i16;
    float *den_f32;
    float *ceps;                 // [row][14] raw cepstra, or [row][15] feature rows when `features`
    int features;                // P8: run WaveProc -> CompCeps -> PostProc -> VAD behind NoiseSup
    int flush;                   // ... and FlushAdvProcess after the last call (6 more rows)
    const int64_t *ceps_offsets; // device
    int32_t *n_ceps;
    uint8_t *vad_trace;
    const int64_t *trace_offsets;
    // P16 outputs
    float *out_f32;
    int32_t *flags;              // [5][total_blocks]
    int64_t total_blocks;
    float *gains;
    int32_t *n_rows;
    const int64_t *block_offsets;
    // persistent streams (optional): state[u] loaded at start and stored at the end
    NsState *state;
    int load_state;
    int32_t *produced;           // optional per-block bitmask: 1 = shift output, 2 = stage 0 ran, 4 = skipped
    int zero_check;              // 0 none (DoNoiseSup), 1 leading frames only (DoAdvProcess), 2 anywhere (func_Wiener)
    float *spec_trace;           // optional [block][8][spec] spectral state per block (se_ns_spectra_batch)
    // byte range of `samples` that bulk-async copies may touch ([bulk_lo, bulk_hi), both 0 = none): the kernel
    // prefetches int16 input with cp.async.bulk, which moves whole 16-byte-aligned spans that can start up to 14 bytes
    // before a block and end up to 16 bytes after it; blocks whose span leaves this range use plain loads
    const char *bulk_lo, *bulk_hi;
    // work queue
    int *work_counter;           // device int, zeroed before launch
};

struct GtDevTables {
    float f1[kGtChannels], f2[kGtChannels], gain[kGtChannels], mid[kGtChannels];
    float hc[16];
    // polar form of the pole z = f1 + i f2 (taken from the FLOAT f1, f2 the reference filters with):
    // radius, unit rotation (cos, sin) and angle in turns, for the rotating-frame default-mode kernels
    float rho[kGtChannels], rot_c[kGtChannels], rot_s[kGtChannels];
    double turns[kGtChannels];
};

}  // namespace se

// One in-order pipeline lane: a stream with its own staging buffers and work-queue counter.  Lane 0
// is the context's visible stream (se_ctx_set_stream); lanes 1.. exist so that SE_MEM_HOST calls can
// overlap H2D, kernels and D2H of successive utterance chunks.
#ifndef SE_LANES
#define SE_LANES 3
#endif
constexpr int kLanes = SE_LANES;
struct se_lane {
    cudaStream_t stream;
    int *d_work_counter;
    void *d_stage[16];
    uint64_t stage_bytes[16];
    cudaEvent_t done;
};

struct se_ctx {
    // Every public entry point that works on a context (or on a stream created from it) holds this lock for the call:
    // handles may be created on one host thread and used on another (the reference's handles are plain heap structs);
    // calls on ONE context serialise, calls on different contexts overlap.
    std::recursive_mutex mu;
    int exact;                  // se_ctx_set_exact: NoiseSup / CompCeps / feature chains in the reference's arithmetic
    int device;
    int sm_count;
    cudaStream_t own_stream;
    se_lane lanes[kLanes];
    se_lane *cur;               // lane the launch helpers currently target
    uint64_t chunk_bytes;       // SE_MEM_HOST pipelining: target I/O bytes per chunk
    int chunk_min_utts;         // ... and the fewest utterances worth a launch of their own
    int chunk_lead;             // se_ctx_set_pipeline_lead bits: 1 = small first chunk on the pipelined kernel, 2 = all chunks'
                                // inputs uploaded back to back on up_stream ahead of the kernels (default 3)
    cudaStream_t up_stream;     // upload stream of the SE_MEM_HOST pipeline
    std::vector<cudaEvent_t> up_events;   // chunk c's input has arrived
    int feature_flush;          // se_ctx_set_feature_flush: the feature-chain calls append FlushAdvProcess's 6 rows
    int ns_pipeline;            // se_ctx_set_ns_pipeline: 0 auto, 1 never, 2 always (se_ns_pipe.cuh)
    // tables
    se::NsTables h_p8, h_p16;
    se::RfftPlan h_rf8, h_rf16;
    se::CcTables h_cc;
    se::GtTables h_gt;
    se::Hp16Tables h_hp16;
    se::NsDevTables d_p8, d_p16;
    se::CcDevTables d_cc;
    std::vector<void *> table_allocs;
    // scratch
    void *d_scratch;
    uint64_t scratch_bytes, scratch_limit;
    // D2H copies into PAGEABLE caller memory go through this pinned arena and reach their destination after the
    // stream synchronisation that ends the call (se::d2h_async / se::finish_host_copies)
    void *bounce;
    uint64_t bounce_cap, bounce_used;
    struct Bounce { void *dst; const void *src; uint64_t bytes; };
    std::vector<Bounce> bounces;
    // profiling
    int profiling;
    struct Ev { cudaEvent_t a, b; int k; };
    std::vector<Ev> pending;
    double k_ms[SE_K_COUNT];
    int64_t k_launches[SE_K_COUNT];
    int64_t launches;
};

namespace se {

// host helpers (se_api.cu)
int ensure_stage(se_ctx *ctx, int slot, uint64_t bytes, void **out);
int ensure_scratch(se_ctx *ctx, uint64_t bytes);
int d2h_async(se_ctx *ctx, void *host, const void *dev, uint64_t bytes);
void finish_host_copies(se_ctx *ctx);
void prof_begin(se_ctx *ctx, int k);
void prof_end(se_ctx *ctx, int k);

// kernel launchers
int launch_ns_p8(se_ctx *ctx, const NsLaunch &a);
int launch_ns_p16(se_ctx *ctx, const NsLaunch &a);
int ns_pipe_capacity(se_ctx *ctx, int profile, const NsLaunch &a);
int launch_rfft(se_ctx *ctx, float *x, int n, int m, int count);
int launch_qmf_split(se_ctx *ctx, const int16_t *x, int64_t n, int count, float *lp, float *hp);
int launch_wave_proc(se_ctx *ctx, float *frames, int count);
int launch_haircell(se_ctx *ctx, const float *in, float *out, int64_t n, int count);
int launch_comp_ceps(se_ctx *ctx, const float *frames, float *coef, int count);

struct G64Launch {
    const int16_t *samples, *samples2;   // samples2: noise (IBM kernel)
    const int64_t *offsets;              // device
    int utt_begin, utt_count;
    int flags;
};
int launch_subband(se_ctx *ctx, const G64Launch &a, int16_t *out);
int launch_ibm(se_ctx *ctx, const G64Launch &a, float *irm, uint64_t *bits, const int64_t *frame_offsets);
// mask_ready (optional): recorded on another stream once `mask` is on the device; only the reversed pass waits for it
int launch_resynth(se_ctx *ctx, const G64Launch &a, const int64_t *h_offsets, const float *mask,
                   const int64_t *h_frame_offsets, int mode, int16_t *out, cudaEvent_t mask_ready = nullptr);
int launch_gammatone(se_ctx *ctx, const float *in, float *out, const int32_t *channel, int64_t n, int count,
                     int flags);
int launch_gammatone_coef(se_ctx *ctx, const float *in, float *out, const float *coef, int64_t n, int count, int flags);
int launch_ibm_rows(se_ctx *ctx, const int16_t *pure64, const int16_t *noise64, const int64_t *offsets, int n_utts,
                    float *irm, uint64_t *bits, const int64_t *frame_offsets, int64_t total_frames);
int launch_resynth_noalloc(se_ctx *ctx, const int16_t *d_samples, int64_t n, const float *d_mask, int mode, int flags,
                           int16_t *d_out);
int launch_addnoise(se_ctx *ctx, const int16_t *pure, int16_t *noise, int16_t *noisy, const int64_t *offsets,
                    int n_utts, int db);
int upload_g64_tables(se_ctx *ctx);

}  // namespace se

// The same fused NoiseSup kernels compiled a second time (se_ns.cu with -DSE_NS_EXACT -fmad=false) in the reference's
// own arithmetic: its rfft butterfly schedule, its summation orders, IEEE division / square root, double-promoted
// sub-expressions.  Selected per context by se_ctx_set_exact().
namespace se_exact {
int launch_ns_p8(se_ctx *ctx, const se::NsLaunch &a);
int launch_ns_p16(se_ctx *ctx, const se::NsLaunch &a);
int launch_rfft(se_ctx *ctx, float *x, int n, int m, int count);
int launch_qmf_split(se_ctx *ctx, const int16_t *x, int64_t n, int count, float *lp, float *hp);
int launch_wave_proc(se_ctx *ctx, float *frames, int count);
int launch_comp_ceps(se_ctx *ctx, const float *frames, float *coef, int count);
}  // namespace se_exact
