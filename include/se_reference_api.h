/* se_reference_api.h — the reference's OWN function-level API for the hot path, exported by
 * libse_b200_dropin.so (thin adapters over the se_*_batch / se_ns_stream_* calls of se_b200.h).
 *
 * A maintainer of guokiddo1/speech_enhancement links this library instead of compiling
 * etsi/cpp/{AdvFrontEnd,ParmInterface,NoiseSup,CompCeps,MelProc,rfft,BufferIn}.c,
 * function/.../aurora_etsi/{NoiseSup,MelProc,rfft}.cpp and the cpp/extractwav.cpp of each tool directory.  Same names, same
 * argument meaning, same return conventions.  Every call runs on the GPU (device chosen by the
 * SE_B200_DEVICE environment variable, default 0); with no usable GPU the calls FAIL (return value
 * noted per function and a message on stderr) — there is no CPU path.
 *
 * Per-call adapters are launch-bound by construction (one utterance, or one 80-sample frame, per
 * launch).  Throughput lives in the additive *_batch calls of se_b200.h.
 */
#ifndef SE_REFERENCE_API_H
#define SE_REFERENCE_API_H
#include <stdint.h>
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- etsi/cpp/AdvFrontEnd.h:11-15 ---------------------------------------------------------------
 * Returns FALSE(0) on success and TRUE(1) on failure, like the reference (AdvFrontEnd.c:205-209).
 * p_denoised[0 .. i_frame - i_frame%80) is written (first (Z+4)*80 samples 0); the tail is left
 * untouched, as in the reference (AdvFrontEnd.c:96-97).                                         */
int etsi_denoise(short *p_data, short *p_denoised, long i_frame);
/* AdvFrontEnd.c:213-226 — including its quirks: the shifted copy only happens when etsi_denoise
 * returned TRUE (i.e. on FAILURE), and it copies i_frame-320 BYTES.  On success p_denoised is left
 * untouched, exactly as the reference does.                                                      */
int etsi_denoise_synchronization(short *p_data, short *p_denoised, long i_frame);
/* etsi_denoise_16k (AdvFrontEnd.c:228-313) copies 160 samples per call out of the 80-sample DenoiseBuffer
 * (:264,290-293): half of what it returns is a heap over-read, so there is no defined output to match.  These
 * return TRUE(1) = failure and print why on stderr; the 16 kHz mode itself (features + the defined low-band
 * denoised stream) is se_etsi16k_features_batch in se_b200.h.                                      */
int etsi_denoise_16k(short *p_data, short *p_denoised, long i_frame);
int etsi_denoise_16k_synchronization(short *p_data, short *p_denoised, long i_frame);

/* ---- function/20141106_speech_enhancement/aurora_speech_enhancement/aurora_speech_enhancement.cpp -----------
 * BufferDenoise (:25-79): denoise a long 8 kHz recording in windows of batch_length samples (the tool uses
 * 3 min = 1 440 000), each window after the first started etsi_offset_points = 320 samples early so that
 * etsi_denoise's 320-sample latency is absorbed, outputs shifted back by 320, last 320 samples zeroed.  Same
 * arithmetic and the same `work` buffer semantics as the reference (each etsi_denoise call overwrites
 * work[0 .. len - len%80) and leaves the rest, so the < 80 stale samples the last window's copy picks up are
 * reproduced); `work` needs batch_length + 320 shorts.  Here ALL windows of the recording go to the GPU as one
 * ragged batch (they are independent), instead of one after the other.  samples < 320 is rejected (the
 * tool's caller skips such files, :167-170).  Returns 0 / negative SE_E* (BufferDenoise itself: void).      */
void BufferDenoise(short *src_wave, short *des_wave, int samples, int batch_length, short *work);
int se_buffer_denoise(const short *src_wave, short *des_wave, int samples, int batch_length /* 0 = 1440000 */);
/* The RMS normalisation the same tool applies before BufferDenoise (:172-184): scales p_data in place to an
 * RMS of `target` (the tool: 3000) with its double/float promotions and (short) cast; returns the factor.  */
double se_rms_normalise(short *p_data, int samples, float target);

/* ---- etsi/cpp/rfft.h:19 — in place, output Re0..Re(n/2), Im(n/2-1)..Im1.  (n,m) must be (256,8) or
 * the plugin's (512,8); anything else prints an error and leaves x untouched.     */
void rfft(float *x, int n, int m);

/* ---- etsi/cpp/ParmInterface.h:192-223 with an OPAQUE FEParamsX ------------------------------------
 * The reference's FEParamsX is a public struct of module pointers; here it is a handle.  Semantics
 * of the four calls are the reference's: DoAdvProcess consumes FrameShift=80 int16 samples, returns
 * TRUE(1) when a denoised shift was written to OutBuf (2+2 frames latency, NoiseSup.c:1436-1439) and
 * writes 14 cepstral coefficients c1..c12,c0,logE to FeatBuf when a cepstral frame is due (the
 * CompCeps step of ParmInterface.c:274-293; *feat_ready set to 1).  A leading all-zero frame returns TRUE(1) with
 * FeatBuf[0..13] = 0 and OutBuf untouched (ParmInterface.c:314-327).
 * Threading: a handle stays bound to the GPU context of the host thread that allocated it; it may be USED from any
 * thread (calls on one context serialise on its internal lock, se_b200.h).                       */
typedef struct FEParamsX FEParamsX;
FEParamsX *AdvProcessAlloc(int SamplingFrequency);            /* only 8000 is on the hot path      */
void AdvProcessInit(FEParamsX *This);
int DoAdvProcess(short *CurrentFrame, short *OutBuf, float *FeatBuf, FEParamsX *This);
int se_adv_feature_ready(FEParamsX *This);                    /* additive: 1 if the last call wrote FeatBuf */
/* Module entry points (etsi/cpp/NoiseSupExports.h:20-26, CompCepsExports.h:17-20) on the same opaque handle;
 * their Alloc/Delete halves are part of AdvProcessAlloc/AdvProcessDelete here.
 * DoNoiseSup: 80 floats in; TRUE(1) when 80 denoised floats were written to OutData (NoiseSup.c:1436-1439),
 * no zero-frame gate.  DoCompCeps: Data = 200-sample frame with Data[-1] valid; Coef = c1..c12,c0,logE. */
typedef struct NoiseSupStructX NoiseSupStructX;
NoiseSupStructX *DoNoiseSupAlloc(void);                       /* NoiseSupExports.h:20: one NoiseSup state (GPU) */
void DoNoiseSupDelete(NoiseSupStructX *NSX);                  /* NoiseSupExports.h:26                           */
void DoNoiseSupInit(FEParamsX *This);
int DoNoiseSup(float *InData, float *OutData, FEParamsX *This);
void DoCompCepsInit(FEParamsX *This);
int DoCompCeps(float *Data, float *Coef, FEParamsX *This);
/* FlushAdvProcess (etsi/cpp/ParmInterface.h:217, ParmInterface.c:347-355 -> DoVADFlush, VAD.c:342-431).  As shipped,
 * DoAdvProcess never feeds the VAD delay line (ParmInterface.c:274-311 is commented out), so the flush walks the
 * freshly initialised 7-slot ring: TRUE six times, then FALSE; FrameCounter advances, and once it exceeds 10 an
 * all-zero row with VAD flag 0 is copied to FeatureBuffer[0..14].  Reproduced exactly.  (The flush of the
 * RE-ENABLED feature chain is se_ctx_set_feature_flush + se_etsi_features_batch in se_b200.h.)              */
int FlushAdvProcess(float *FeatBuf, FEParamsX *This);
void AdvProcessDelete(FEParamsX **This);

/* ---- function/20141106_speech_enhancement/aurora_etsi/NoiseSupExports.h:9-42 -------------------- */
#ifndef SE_NO_PLUGIN_TYPES
typedef void *INSTANCE;
typedef void **PINSTANCE;
typedef int32_t int32s;
#define SE_BUFFSIZE 160
struct esti_denoise_in { float *inData; int dataNum; };
struct esti_denoise_out {
    float *outData;
    int *pSpeechFoundVar, *pSpeechFoundSpec, *pSpeechFoundMel, *pSpeechFoundVADNS, *pFrameCounter;
};
#endif
/* init calls return 1 on success, 0 on failure (NoiseSup.cpp:921,941-944,1083) */
int32s etsi_denoise_mapping_global_init(PINSTANCE sm_glb_pins, INSTANCE sm_glb_res);
int32s etsi_denoise_mapping_thread_init(PINSTANCE sm_thd_pins, INSTANCE sm_glb_ins);
/* fp_Wiener is a FILE*: one line of 25 "%f " per stage-2 frame (NoiseSup.cpp:1318-1326); may be NULL
 * here (the reference would crash).  Returns FALSE(0) (NoiseSup.cpp:1406); on a GPU failure returns 1.
 * outData / flag arrays are only written where the reference writes them (caller zeroes).        */
int32s etsi_denoise_mapping_func_Wiener(INSTANCE sm_glb_ins, INSTANCE sm_thd_ins, INSTANCE in_ins,
                                        INSTANCE out_ins, INSTANCE fp_Wiener);
void etsi_denoise_mapping_thread_release(PINSTANCE sm_thd_pins);
void etsi_denoise_mapping_global_release(PINSTANCE sm_glb_pins);

/* ---- <tool>/cpp/extractwav.cpp, HuWang.h:50,62 — C twins of the asdk::CWave-taking C++ functions -------
 * (asdk is un-vendored; the C++ signatures are header-only adapters in se_reference_api.hpp).    */
/* subbband(): out64xn = 64 channel-major rows of n int16 (what the reference writes as 64 wavs). */
int se_subband(const short *wav, long n, short *out64xn);
/* resynth(): mode 0 = resyth_64sub_IBM, 1 = resyth_64sub_ori, 2 = 1dnn_resynth.  mask is
 * [num_frame][64]; in IBM mode entries > 0.5 are overwritten with 1.0 like the reference does
 * (resyth_64sub_IBM/cpp/extractwav.cpp:99).  Bit-exact kernels by default; with SE_B200_FAST=1 in the
 * environment the time-chunked default-mode kernels run instead (within 1 LSB, far lower latency).  */
int se_resynth(const short *wav, long n, float *mask, int num_frame, int mode, short *out);
/* make_single_IBM arithmetic fused with the two subbband() calls feeding it.                     */
int se_single_ibm(const short *pure, const short *noise, long n, float *irm /*[frames][64]*/,
                  unsigned char *ibm /*[frames][64]*/);
/* addnoise(): noise is scaled in place; noisy = pure + noise.                                    */
int se_addnoise(const short *pure, short *noise, short *noisy, long n, int db);
/* gammaToneFilter(): channel index instead of the by-value gammaTone struct (HuWang.h:50).      */
int se_gammatone_filter(const float *input, float *output, int channel, long sigLength);
/* hairCell(float *input, float *output, long sigLength) (HuWang.h:62, extractwav.cpp:213-258).  */
int se_hair_cell(const float *input, float *output, long sigLength);
/* gammaToneFilter() with the by-value channel parameters of HuWang.h:49 (any cf / bw / midEarCoeff); the C++
 * signature gammaToneFilter(float*, float*, gammaTone, long) is a header-only adapter in se_reference_api.hpp. */
int se_gammatone_filter_params(const float *input, float *output, float cf, float bw, float midEarCoeff,
                               long sigLength);
/* make_single_IBM() arithmetic on the 2 x 64 sub-band signals it reads back from the wavs subbband() wrote
 * (show_IBM.cpp:118-170): pure64xn / noise64xn are 64 channel-major rows of n int16; irm is [frames][64].
 * The C++ signature make_single_IBM(char*, FILE*, FILESNAME) is a header-only adapter in se_reference_api.hpp. */
int se_single_ibm_rows(const short *pure64xn, const short *noise64xn, long n, float *irm);

#ifdef __cplusplus
}
#endif
#endif
