/* se_b200.h — C ABI of libse_b200.so, the B200 (sm_100a) implementation of the frame-level
 * speech-enhancement hot path of guokiddo1/speech_enhancement.
 *
 * Plain C: pointers and sizes only, no C++/torch types.  Every entry point returns 0 on success or a
 * negative SE_E* code and never exits the process (the reference fprintf+exit(0)s on allocation
 * failure, etsi/cpp/NoiseSup.c:983-987).  se_last_error() returns a thread-local message.
 *
 * There is NO CPU fallback: every compute entry point fails with SE_ENODEVICE if no CUDA device is
 * usable.  The reference-named drop-in entry points (etsi_denoise, rfft, etsi_denoise_mapping_*, ...)
 * are declared in se_reference_api.h and are thin adapters over the *_batch calls below.
 *
 * Batches are ragged: `samples` holds the utterances back to back and offsets[u]..offsets[u+1]
 * (n_utts+1 entries, in samples) delimits utterance u.  `mem` says where ALL data pointers of the call
 * live: SE_MEM_HOST (the library stages H2D/D2H itself) or SE_MEM_DEVICE (pointers are device
 * pointers on the context's GPU; nothing is copied, the call is asynchronous on the context stream
 * until se_ctx_sync()).  offsets / frame-offset arrays are always HOST arrays and start at 0.
 * Output rows the path does not produce (capacity beyond n_rows / n_ceps, latency blocks of gains / flags) are zeroed
 * in SE_MEM_HOST mode; in SE_MEM_DEVICE mode the caller's device arrays are written in place and such rows are left
 * as they were (nothing is cleared behind the caller's back on the device-resident path).
 */
#ifndef SE_B200_H
#define SE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SE_OK          0
#define SE_EINVAL     -1
#define SE_ENODEVICE  -2   /* no CUDA device / driver: the product has no CPU path            */
#define SE_ECUDA      -3   /* CUDA runtime error, see se_last_error()                          */
#define SE_ENOMEM     -4
#define SE_EUNSUPPORTED -5

#define SE_MEM_HOST    0
#define SE_MEM_DEVICE  1

typedef struct se_ctx se_ctx;

const char *se_last_error(void);
int  se_device_count(void);
/* One context per GPU (reference analogue: one "thread instance" per worker,
 * function/.../aurora_etsi/NoiseSupExports.h:36).  Owns a stream, constant tables and scratch.
 * Threading: every call on a context (or on a stream created from it) holds the context's internal lock, so a
 * context or stream handle may be created on one host thread and used on others -- calls on ONE context serialise,
 * like the plugin's thread instance; different contexts run concurrently, and their calls overlap on the GPU
 * (results reach pageable caller arrays through a pinned arena so that no call holds the driver while its kernel
 * runs).  The reference-named adapters of se_reference_api.h use one context per calling thread; a handle they
 * return (FEParamsX, plugin thread instance) stays bound to the context of the thread that created it. */
int  se_ctx_create(int device, se_ctx **out);
void se_ctx_destroy(se_ctx *ctx);
/* Launch on a caller-owned CUDA stream (cudaStream_t passed as void*); NULL restores the own stream. */
int  se_ctx_set_stream(se_ctx *ctx, void *cuda_stream);
int  se_ctx_sync(se_ctx *ctx);
/* SE_MEM_HOST batch calls are pipelined: the batch is cut into contiguous utterance chunks of about
 * chunk_bytes of input+output (default 512 MiB) but never fewer than min_utts utterances (default 1024,
 * enough to keep the persistent warps busy); H2D, kernels and D2H of successive chunks overlap.
 * Results do not depend on the chunking (utterances are independent). */
int  se_ctx_set_pipeline(se_ctx *ctx, uint64_t chunk_bytes, int min_utts);
/* Two refinements of that pipeline for the NoiseSup calls, as bits of `mode` (default 3):
 *   1  the call opens with a small first chunk -- as many utterances as the single-utterance kernel
 *      (se_ctx_set_ns_pipeline) runs at once -- so results start flowing back after ~2 ms instead of after the batch
 *      kernel's per-utterance latency;
 *   2  the inputs of ALL chunks are uploaded back to back on a stream of their own, ahead of the kernels, instead of
 *      chunk by chunk between the downloads: the upload engine finishes early and the download engine, which bounds
 *      the call, runs the rest of it alone (host links are slower with both directions busy). */
int  se_ctx_set_pipeline_lead(se_ctx *ctx, int mode);
/* FlushAdvProcess (etsi/cpp/ParmInterface.c:347-355 -> DoVADFlush, VAD.c:342-431): with enable != 0,
 * se_etsi_features_batch and se_etsi16k_features_batch end every utterance the way the ETSI driver loop
 * `while (FlushAdvProcess(FeatureBuffer, fe)) write` does: SE_FEATURE_FLUSH_ROWS more rows leave the VAD's 7-frame
 * delay line (for utterances too short to have started it, the reference's loop writes its stale FeatureBuffer six
 * times; so does this).  feat_offsets capacity per utterance grows by SE_FEATURE_FLUSH_ROWS.  Default off: none of
 * the reference's drivers flushes. */
#define SE_FEATURE_FLUSH_ROWS 6
int  se_ctx_set_feature_flush(se_ctx *ctx, int enable);
/* Kernel choice of se_etsi_denoise_batch / se_p16_wiener_batch (no counterpart in the reference, whose etsi_denoise /
 * func_Wiener are one serial loop per call -- etsi/cpp/AdvFrontEnd.c:125-210, aurora_etsi/NoiseSup.cpp:1140-1407):
 * batches too small to give every SM its warps run on a kernel that spreads ONE utterance over a CTA of 5-8 warps
 * (input + stage-0 PSD | stage-0 recursion | FIR + stage-1 PSD | stage-1 recursion | FIR + outputs | CompCeps halves |
 * statistics, software-pipelined over blocks) -- 3-4x lower latency for a lone utterance, bit-identical results.  mode 0 (default): chosen by batch size;
 * 1: always the one-warp-per-utterance batch kernel; 2: the pipelined kernel whenever it is able to run. */
int  se_ctx_set_ns_pipeline(se_ctx *ctx, int mode);
/* Arithmetic mode of the NoiseSup family (se_etsi_denoise_batch, se_etsi_features_batch, se_etsi16k_features_batch,
 * se_p16_wiener_batch, se_ns_spectra_batch, se_ns_stream_*, se_rfft_batch, se_comp_ceps_batch, se_wave_proc_batch,
 * se_qmf_split_batch) -- the counterpart of SE_FLAG_EXACT of the G64 calls.
 *   0 (default): fast arithmetic -- in-register shuffle FFT, FMA contraction, MUFU reciprocal / square root, float
 *      logarithms, tree-shaped sums.  Within the parity bars (waveform >= 60 dB, spectra / gains / cepstra rel 1e-4 of
 *      their scale) but not bit-identical, and a 1-ulp difference can move one of WaveProc's integer peak positions.
 *   1: the SAME kernels compiled a second time in the reference's own arithmetic: its split-radix rfft butterfly
 *      schedule (etsi/cpp/rfft.c:45-180; for the plugin the truncated rfft(x,512,8)), every sum in the reference's order,
 *      IEEE division and square root, the double-promoted sub-expressions of the C sources, no FMA contraction.  The
 *      only freedom left is libm's log / log10 / exp in double (<= 1 ulp in both libraries before rounding to float).
 *      Used by the parity tests to show that every frame of the feature chains meets rel 1e-4; ~3-5x slower. */
int  se_ctx_set_exact(se_ctx *ctx, int enable);
/* Upper bound for internal device scratch (gammatone intermediate), bytes.  Default 48 GiB. */
int  se_ctx_set_scratch_limit(se_ctx *ctx, uint64_t bytes);

/* Pinned host memory for SE_MEM_HOST callers that want PCIe-rate staging (plain malloc works too,
 * at pageable-copy speed).  Returns NULL on failure. */
void *se_host_alloc(uint64_t bytes);
void  se_host_free(void *p);

/* ---- multi-GPU sharding (host-side, no device work) ---------------------------------------------
 * Utterances are independent, so a batch shards by contiguous utterance ranges with no collective
 * (the reference's own N-worker model: one work item = one file,
 * function/.../aurora_speech_enhancement/aurora_speech_enhancement.cpp:110-120).  first_utt gets
 * n_shards+1 entries; shard k = utterances [first_utt[k], first_utt[k+1]), balanced by sample count. */
int se_shard_plan(const int64_t *offsets, int n_utts, int n_shards, int32_t *first_utt);

/* ---- kernel timing (CUDA events on the launching stream) -------------------------------------- */
#define SE_K_NS_P8     0   /* fused NoiseSup(+CompCeps) stream kernel, 8 kHz profile  */
#define SE_K_NS_P16    1   /* fused NoiseSup stream kernel, 16 kHz plugin profile     */
#define SE_K_GT_FWD    2   /* gammatone forward pass (resynth pass 1)                  */
#define SE_K_GT_BWD    3   /* gammatone reversed pass + mask weights + channel sum     */
#define SE_K_SUBBAND   4   /* gammatone + hair cell + int16 sub-band writer            */
#define SE_K_IBM       5   /* fused sub-band -> IRM / IBM                              */
#define SE_K_ADDNOISE  6
#define SE_K_COUNT     8
int se_ctx_profiling(se_ctx *ctx, int enable);          /* record event pairs around own kernels  */
int se_ctx_kernel_time(se_ctx *ctx, int kernel_id, double *total_ms, int64_t *launches); /* syncs */
int se_ctx_reset_kernel_times(se_ctx *ctx);
int64_t se_ctx_launch_count(se_ctx *ctx);               /* kernels launched by this library so far */

/* ---- P8: ETSI NoiseSup (+CompCeps), replaces etsi_denoise (etsi/cpp/AdvFrontEnd.c:125-210) ------
 * denoised : int16, same layout as samples; the first (Z+4)*80 samples of an utterance are 0 (Z =
 *            leading all-zero 80-sample frames, etsi/cpp/ParmInterface.c:244-250) and the tail
 *            len%80 is written as 0 (the reference leaves it untouched).            [required]
 * ceps     : float [ceps_offsets[u] + m][14] = c1..c12,c0,logE for cepstral frame m, produced by
 *            the CompCeps step of etsi/cpp/ParmInterface.c:274-293 (shipped commented out in the
 *            reference; etsi/cpp/CompCeps.c:368-549).  NULL = NoiseSup only.
 * ceps_offsets : host int64[n_utts+1] frame offsets (se_etsi_num_ceps_frames gives an upper bound).
 * n_ceps   : int32[n_utts] frames actually produced (same memory kind as data), may be NULL.
 * den_f32  : optional float denoised stream (before the int16 cast), may be NULL.
 * vad_trace: optional uint8 per 80-sample call, laid out at offsets[u]/80-ish: trace_offsets[u]+call;
 *            bit0 flagVAD, bit1 SpeechFoundVar, bit2 Spec, bit3 Mel, bit4 VADNS, bit6 frame
 *            produced, bit7 zero-frame skipped.  trace_offsets host int64[n_utts+1], may be NULL. */
int64_t se_etsi_num_ceps_frames(int64_t n_samples);     /* max(0, n/80 - 6)                     */
int se_etsi_denoise_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                          int16_t *denoised, float *ceps, const int64_t *ceps_offsets, int32_t *n_ceps,
                          float *den_f32, uint8_t *vad_trace, const int64_t *trace_offsets, int mem);

/* ---- P8: the whole ETSI feature chain, DoAdvProcess with the block the reference ships commented out
 *      re-enabled (etsi/cpp/ParmInterface.c:274-311): DoNoiseSup -> DoWaveProc (WaveProc.c:386) ->
 *      DoCompCeps (CompCeps.c:368) -> DoPostProc (PostProc.c:117) -> DoVADProc (VAD.c:214), plus the
 *      null vectors of leading all-zero frames (ParmInterface.c:314-327).  No flush at the end of an
 *      utterance (FlushAdvProcess is never called by the reference's drivers either).
 * features : float [feat_offsets[u] + row][15] = c1..c12, c0, logE after blind equalisation, VAD flag
 *            (FeatureBuffer[0..14]); rows come out in call order: Z null rows, then cepstral frame
 *            m at processed call m+12 (6 calls of NoiseSup/framing latency + the VAD's 7-frame delay
 *            line): max(0, len/80 - Z - 12) rows.                                      [required]
 * feat_offsets : host int64[n_utts+1]; capacity per utterance >= se_etsi_num_feature_rows(len).
 * n_rows   : int32[n_utts] rows written (same memory kind as data), may be NULL.
 * denoised : optional int16 denoised stream, as se_etsi_denoise_batch. */
int64_t se_etsi_num_feature_rows(int64_t n_samples);    /* n/80: upper bound                    */
int se_etsi_features_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                           float *features, const int64_t *feat_offsets, int32_t *n_rows, int16_t *denoised,
                           int mem);

/* ETSI 16 kHz mode: AdvProcessAlloc(16000) (etsi/cpp/ParmInterface.c:99-109) + DoAdvProcess with
 * the feature block re-enabled.  Each call consumes 160 samples: Do16kProcessing (etsi/cpp/16kHzProcessing.c:711-787)
 * splits them into a 0-4 kHz half (-> NoiseSup -> WaveProc -> CompCeps) and a 4-8 kHz half whose three mel-band
 * energies are spectrally subtracted and coded against the low band inside DoNoiseSup (NoiseSup.c:1216-1235,
 * 1307-1327, 1402-1435; DoSpecSub16k 16kHzProcessing.c:388-480) and decoded/merged inside WI8CompCeps
 * (CompCeps.c:464-531): 26 log bands -> c1..c12, c0, logE (+ HP energy) -> PostProc -> VAD.
 * samples  : int16 16 kHz, ragged batch as above; the tail len % 160 is ignored.
 * features : float[feat_offsets[u] + row][15], rows as se_etsi_features_batch with calls = len/160.   [required]
 * feat_offsets : capacity per utterance >= se_etsi16k_num_feature_rows(len).
 * denoised_lp / den_offsets : optional int16 denoised LOW-BAND stream at 8 kHz, 80 samples per call (what
 *            etsi_denoise_16k's DenoiseBuffer holds, AdvFrontEnd.c:264), indexed by its own host int64[n_utts+1]
 *            offsets with capacity >= 80 * (len/160); unused capacity is zeroed.
 * The reference's driver etsi_denoise_16k (AdvFrontEnd.c:228-313) copies 160 samples per call out of that
 * 80-sample buffer, a heap over-read; this call returns what is defined. */
int64_t se_etsi16k_num_feature_rows(int64_t n_samples); /* n/160: upper bound                   */
int se_etsi16k_features_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                              float *features, const int64_t *feat_offsets, int32_t *n_rows, int16_t *denoised_lp,
                              const int64_t *den_offsets, int mem);

/* Module-level entry: Do16kProcessing (etsi/cpp/16kHzProcessing.c:711-787) as a stream over `count` independent
 * signals of n int16 samples each (n a multiple of 160): lp / hp = float[count][n/2], the 0-4 kHz half and the
 * 4-8 kHz half flipped to base band (118-tap QMF pair, decimated by 2, delay line starting from rest). */
int se_qmf_split_batch(se_ctx *ctx, const int16_t *samples, int64_t n, int count, float *lp, float *hp, int mem);

/* Module-level entry: DoCompCeps (etsi/cpp/CompCeps.c:309-318,368-549) on n independent frames of 201 floats
 * (frame[0] = the sample before the frame, needed by the 0.9 pre-emphasis) -> 14 floats c1..c12,c0,logE each. */
int se_comp_ceps_batch(se_ctx *ctx, const float *frames201, int n_frames, float *coef14, int mem);
/* Known-answer entry: DoWaveProc (etsi/cpp/WaveProc.c:386-453) on n frames of 200 floats, in place. */
int se_wave_proc_batch(se_ctx *ctx, float *frames, int n_frames, int mem);

/* ---- P16: plugin NoiseSup, replaces etsi_denoise_mapping_func_Wiener
 *      (function/20141106_speech_enhancement/aurora_etsi/NoiseSup.cpp:1140-1407) ------------------
 * samples are int16 (is_f32=0) or float (is_f32=1, the plugin's own input type).
 * out_f32  : float denoised, same layout as samples; blocks where nothing is produced are 0.
 * flags    : int32 [5][total_blocks] = SpeechFoundVar, Spec, Mel, VADNS, FrameCounter, block b of
 *            utterance u at block_offsets[u]+b; entries of skipped/latency blocks are 0.
 * gains    : float [block_offsets[u] + row][25], one row per stage-2 frame (the ".wiener" text rows,
 *            NoiseSup.cpp:1318-1326); n_rows int32[n_utts].
 * block_offsets : host int64[n_utts+1], block_offsets[u+1]-block_offsets[u] >= len_u/160. */
int se_p16_wiener_batch(se_ctx *ctx, const void *samples, int is_f32, const int64_t *offsets, int n_utts,
                        float *out_f32, int32_t *flags, float *gains, int32_t *n_rows,
                        const int64_t *block_offsets, int mem);

/* ---- spectral state per block, for parity checks at the spectrum level -------------
 * Runs the same fused NoiseSup kernel as the two calls above (profile 8: se_etsi_denoise_batch, int16 input;
 * profile 16: se_p16_wiener_batch) and writes, for block b of utterance u,
 * spectra[block_offsets[u] + b][8][spec] (spec = 65 | 129) = the SPECTRUM state of NoiseSupStructX after the call
 * (etsi/cpp/NoiseSup.c:102-115, as FilterCalc :449-563 leaves it):
 *   [0] nSigSE1  [1] noiseSE1  [2] denSigSE1  [3] nSigSE2  [4] noiseSE2  [5] denSigSE2
 *   [6] / [7] sqrt(PSDMeaned) of stage 1 / 2 (FFTtoPSD :249-270 -> PSDMean :289-303 -> sqrt :522-526).
 * Planes of a stage are only written in blocks where that stage ran (produced[] bit 1 = stage 1 of the two-stage
 * filter ran, bit 0 = stage 2 ran and a shift was output, bit 2 = block skipped by the zero gate); others stay 0.
 * spectra is required, produced (int32 per block, same indexing) optional.  A test / diagnostic entry: 8 spectra per
 * block are 13-26x the path's normal output bytes. */
int se_ns_spectra_batch(se_ctx *ctx, int profile, const void *samples, int is_f32, const int64_t *offsets, int n_utts,
                        float *spectra, int32_t *produced, const int64_t *block_offsets, int mem);

/* ---- persistent NoiseSup streams (state carried across calls, like the plugin's thread instance
 *      reused over 64 channel files, function/.../aurora_etsi_test/aurora_etsi_test.cpp:21-141) --- */
typedef struct se_ns_stream se_ns_stream;
int  se_ns_stream_create(se_ctx *ctx, int profile /*8|16*/, se_ns_stream **out);
void se_ns_stream_destroy(se_ns_stream *s);
int  se_ns_stream_reset(se_ns_stream *s);
/* Zero-frame gate applied in front of the stream: 0 none (DoNoiseSup; default for profile 8),
 * 1 leading all-zero blocks only (DoAdvProcess, etsi/cpp/ParmInterface.c:244-250), 2 anywhere
 * (func_Wiener, aurora_etsi/NoiseSup.cpp:1159-1170; default for profile 16). */
int  se_ns_stream_set_zero_gate(se_ns_stream *s, int mode);
/* Push n_blocks*shift float samples (HOST pointers; the call is synchronous).  out receives shift
 * floats per block.  produced[b] is a bit mask: 1 = a denoised shift was output for block b
 * (DoNoiseSup's BOOLEAN, etsi/cpp/NoiseSup.c:1436-1439), 2 = stage 0 ran, 4 = block skipped by the
 * zero gate.  P16: func_Wiener semantics; flags is int32[5][n_blocks], gains float[n_blocks][25]
 * (*n_rows rows valid).  P8: ceps float[n_blocks][14], *n_ceps rows valid (the CompCeps step of
 * ParmInterface.c:274-293).  Any output pointer may be NULL. */
int  se_ns_stream_push(se_ns_stream *s, const float *in, int n_blocks, float *out, int32_t *produced,
                       int32_t *flags, float *gains, int32_t *n_rows, float *ceps, int32_t *n_ceps);

/* ---- G64: 64-channel gammatone analysis / masks / resynthesis ----------------------------------- */
#define SE_FLAG_EXACT   1   /* evaluate the IIR recurrences with separately rounded mul/add, in the
                               reference's order (bit-identical to its float arithmetic); default is
                               FMA-contracted (faster, ~-120 dB relative difference)               */
#define SE_FLAG_OUT_F32 2   /* se_resynth_batch: `out` is a float array that receives resynth_e[] before the
                               (short) cast (resyth_64sub_IBM/cpp/extractwav.cpp:109-122), for waveform-SNR checks */
#define SE_FLAG_LATENCY 4   /* se_resynth_batch, default (non-EXACT) mode: cut every utterance into 2 048-sample
                               time chunks instead of >= 1 s ones -- 8x more independent work items per utterance
                               for 2.9x the arithmetic; for calls with a handful of utterances (the drop-in's
                               resynth() with SE_B200_FAST=1 uses it).  Same accuracy bars as the default mode;
                               the output is not bit-identical to a call without the flag                  */
/* subbband (enhancement_extract_subband_linux/cpp/extractwav.cpp:36-97): int16 hair-cell output per
 * channel; out holds, for utterance u, 64 rows of len_u samples starting at 64*offsets[u]. */
int se_subband_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                     int16_t *out, int flags, int mem);
/* make_single_IBM (enhancement_extract_subband_linux/cpp/show_IBM.cpp:92-213) fused with the two
 * subbband() calls that feed it: per (frame, channel) IRM = Sp/(Sp+Sn) and IBM = (Sn < Sp).
 * frames_u = (len_u-320)/160+1.  irm: float[frame_offsets[u]+f][64]; ibm_bits: uint64 per frame
 * (bit c = channel c).  Either output may be NULL.  Always evaluated in exact mode. */
int64_t se_ibm_num_frames(int64_t n_samples);
int se_ibm_batch(se_ctx *ctx, const int16_t *pure, const int16_t *noise, const int64_t *offsets, int n_utts,
                 float *irm, uint64_t *ibm_bits, const int64_t *frame_offsets, int mem);

/* make_single_IBM exactly as the reference drives it (show_IBM.cpp:118-170): from the MATERIALISED sub-band rows
 * that subbband() wrote for the clean and the noise signal (pure64 / noise64 laid out like se_subband_batch's
 * output: 64 rows of len_u int16 starting at 64*offsets[u]).  Same outputs as se_ibm_batch; identical values when
 * the rows are what se_subband_batch produces for the same waveforms. */
int se_ibm_rows_batch(se_ctx *ctx, const int16_t *pure64, const int16_t *noise64, const int64_t *offsets, int n_utts,
                      float *irm, uint64_t *ibm_bits, const int64_t *frame_offsets, int mem);

/* extract_Wiener (function/20141106_speech_enhancement/aurora_etsi_test/aurora_etsi_test.cpp:16-141) for a batch of
 * noisy utterances, device-resident from end to end: subbband() -> the 64 hair-cell channel signals of each utterance
 * (what the driver reads back as <id>_noisy_<chan>.wav), each cut to a multiple of 160 samples (:100-101), all 64 run
 * through etsi_denoise_mapping_func_Wiener on ONE thread instance without re-init (:21-22, 90-141) -- the NoiseSup state
 * carries over from channel to channel, so an utterance is one 64 * (len - len % 160)-sample stream for one warp.
 * The 64 x L sub-band signals never leave the GPU and outData is not produced (the driver frees it unused, :133).
 * gains : float[block_offsets[u] + row][25], the rows of the 64 ".wiener" files of utterance u back to back (file c
 *         holds the rows whose stage-2 frame falls into channel c; n_rows[u] rows in total; the rest stay 0).
 * flags : int32[5][block_offsets[n_utts]] as se_p16_wiener_batch, block b of channel c at block_offsets[u] + c*(len/160) + b.
 * block_offsets : host int64[n_utts+1], EXACTLY the running sum of se_extract_wiener_num_blocks(len) = 64 * (len/160).
 * Intermediates (130 B per sample) live in the context scratch; batches beyond se_ctx_set_scratch_limit run in groups. */
int64_t se_extract_wiener_num_blocks(int64_t n_samples);
int se_extract_wiener_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                            float *gains, int32_t *flags, int32_t *n_rows, const int64_t *block_offsets, int mem);
/* resynth (resyth_64sub_IBM/cpp/extractwav.cpp:9-132 and its _ori / 1dnn variants).
 * mask: float[frame_offsets[u]+f][64]; frames_u = se_resynth_num_frames(len_u, mode). */
#define SE_RESYNTH_IBM   0
#define SE_RESYNTH_ORI   1
#define SE_RESYNTH_1DNN  2
int64_t se_resynth_num_frames(int64_t n_samples, int mode);
int se_resynth_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                     const float *mask, const int64_t *frame_offsets, int mode, int flags,
                     int16_t *out, int mem);
/* resynth() for ONE utterance without any allocation in the call.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arithmetic, same work-item plan and therefore the same output, bit
 * for bit, as se_resynth_batch on that utterance (mode / flags as there, SE_FLAG_EXACT and SE_FLAG_OUT_F32 included).
 * d_samples (int16[n_samples]), d_mask (float[se_resynth_num_frames(n_samples, mode)][64]) and d_out (int16[n_samples],
 * or float with SE_FLAG_OUT_F32) are DEVICE pointers on the context's GPU; the call is asynchronous on the context
 * stream (se_ctx_set_stream / se_ctx_sync).  It calls no cudaMalloc, cudaFree or allocating library routine and copies
 * nothing from the host: the [sample][64] float intermediate of the two gammatone passes and the work items (planned
 * by a device kernel) live in a file-scope static __device__ array of 64 MiB that EVERY GPU which loads libse_b200.so
 * holds until the process exits.  Limit: n_samples <= SE_RESYNTH_NOALLOC_MAX_SAMPLES (12.5 s at 16 kHz); longer
 * utterances are rejected with SE_EINVAL and a message (use se_resynth_batch).  All calls on one GPU -- from any
 * context, stream or thread -- share that one array: calls that OVERLAP on the GPU race and corrupt each other's
 * output; issue them on one stream or order them with events.  The array keeps the last call's contents; nothing in it
 * needs to start at zero (every element a call reads is written earlier in the same call). */
#define SE_RESYNTH_NOALLOC_MAX_SAMPLES 200000
int se_resynth_noalloc(se_ctx *ctx, const int16_t *d_samples, int64_t n_samples, const float *d_mask, int mode, int flags,
                       int16_t *d_out);
/* addnoise (enhancement_extract_subband_linux/cpp/extractwav.cpp:6-35): noise is scaled IN PLACE. */
int se_addnoise_batch(se_ctx *ctx, const int16_t *pure, int16_t *noise, int16_t *noisy,
                      const int64_t *offsets, int n_utts, int db, int mem);

/* ---- primitives exposed for known-answer tests ---------------------------------------------------
 * Batched rfft exactly as the reference lays it out (Re0..Re(n/2), Im(n/2-1)..Im1;
 * etsi/cpp/rfft.c:29-30).  (n,m) in {(256,8), (512,8) [the plugin's truncated schedule, SURVEY Q1]}. */
int se_rfft_batch(se_ctx *ctx, float *x, int n, int m, int count, int mem);
/* gammaToneFilter (extractwav.cpp:168-212) for `count` signals of length n with per-signal channel. */
int se_gammatone_batch(se_ctx *ctx, const float *in, float *out, const int32_t *channel, int64_t n,
                       int count, int flags, int mem);
/* The same filter with caller-supplied coefficients (gammaToneFilter takes its gammaTone channel BY VALUE, HuWang.h:49,
 * so any cf / bw / midEarCoeff is legal): coef is a HOST array [count][3] = f1, f2, gain as extractwav.cpp:176-183
 * derives them (the drop-in adapter gammaToneFilter() of se_reference_api.hpp does that). */
int se_gammatone_coef_batch(se_ctx *ctx, const float *in, float *out, const float *coef, int64_t n, int count, int flags,
                            int mem);
/* hairCell (extractwav.cpp:213-258, Meddis 1988) for `count` signals of length n, each from the resting state. */
int se_haircell_batch(se_ctx *ctx, const float *in, float *out, int64_t n, int count, int mem);
/* init-time tables as the device sees them (host copies), for table parity tests */
int se_get_table(se_ctx *ctx, const char *name, float *out, int capacity, int *count);

#ifdef __cplusplus
}
#endif
#endif
