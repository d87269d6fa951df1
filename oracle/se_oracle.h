/* TEST INFRASTRUCTURE — the parity oracle.  NOT part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * Plain-C CPU restatement of the reference's hot path, pinned bit-for-bit against
 * the reference's own sources compiled in oracle/_ref (tests/test_oracle_vs_ref.py, and the
 * committed vectors in tests/golden/ that were generated from oracle/_ref).
 * Every function cites the reference file:line it follows.
 */
#ifndef SE_ORACLE_H
#define SE_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---------- math primitives ---------- */
void so_rfft(float *x, int n, int m);                 /* etsi/cpp/rfft.c:45 (C: double cos/sin)              */
void so_rfft_cxx(float *x, int n, int m);             /* aurora_etsi/rfft.cpp:45 (C++: float cos/sin)        */

/* ---------- NoiseSup stream (both profiles) ---------- */
#define SO_PROFILE_P8  8      /* etsi/cpp/NoiseSup.h:28-79                              */
#define SO_PROFILE_P16 16     /* function/.../aurora_etsi/NoiseSup.h:22-72              */
#define SO_MAX_SPEC 129
#define SO_MEL 25

typedef struct so_ns so_ns;

typedef struct {
    int produced;       /* 1 if a denoised shift was output by this block                          */
    int stage0_ran, stage1_ran;
    int flag_vad, hang_over, nb_speech, nb_frame0, nb_frame1;
    int sp_var, sp_spec, sp_mel, sp_vadns;
    float mean_en, alfa_gf, low_snr, den_en;
    float gains[SO_MEL];  /* stage-2 band gains after gain factorisation (the ".wiener" row)       */
} so_ns_info;

so_ns *so_ns_create(int profile);
void so_ns_destroy(so_ns *s);
int so_ns_shift(const so_ns *s);                      /* 80 or 160                                 */
int so_ns_spec(const so_ns *s);                       /* 65 or 129                                 */
/* One DoNoiseSup call (etsi/cpp/NoiseSup.c:1061) / one loop iteration of func_Wiener
 * (aurora_etsi/NoiseSup.cpp:1156-1405, after the zero-block test). */
int so_ns_block(so_ns *s, const float *in, float *out, so_ns_info *info);
/* state views for traces: which = 0 nSigSE1, 1 noiseSE1, 2 denSigSE1, 3 nSigSE2, 4 noiseSE2, 5 denSigSE2,
 * 6 / 7 sqrt(PSDMeaned) of stage 1 / 2 as FilterCalc leaves it (NoiseSup.c:522-526) */
const float *so_ns_spectrum(const so_ns *s, int which);
/* init-time tables */
void so_ns_tables(const so_ns *s, float *sig_window, float *ir_window, float *bank /*25*spec-or-128*/,
                  int *bank_start, int *bank_len, float *idct /*25*25*/);

/* ---------- P8 whole-utterance entry points ---------- */
/* etsi_denoise (etsi/cpp/AdvFrontEnd.c:125) plus, when ceps!=NULL, the CompCeps step the reference
 * ships commented out (etsi/cpp/ParmInterface.c:274-293).  den_f32 (optional) receives the float
 * denoised stream aligned like den_i16.  Returns number of 80-sample calls. */
long so_etsi_denoise(const int16_t *in, long n, int16_t *den_i16, float *den_f32, float *ceps,
                     long *n_ceps, so_ns_info *per_call /* optional, [n/80] */);
void so_comp_ceps(const float *frame201, float *coef14);   /* etsi/cpp/CompCeps.c:368-549          */
void so_cc_tables(float *hamming100, float *dct12x23, int *start23, int *len23, float *tri23x129);

/* ---------- P8 back half: WaveProc / PostProc / VAD (etsi/cpp/ParmInterface.c:274-311) ---------- */
typedef struct { float buf[7][16]; int focus, hang_over, h_count, v_count, flush_focus; } so_vad;
void so_wave_proc(float *frame200);                          /* etsi/cpp/WaveProc.c:386-453          */
void so_post_proc(float *coef14, float *lms12);              /* etsi/cpp/PostProc.c:117-150          */
void so_vad_init(so_vad *v);
int so_vad_proc(so_vad *v, float *fb15, int speech_found, int frame_counter);   /* VAD.c:214-330    */
/* NoiseSup -> WaveProc -> CompCeps -> PostProc -> VAD; feat[row][15] = c1..c12,c0,logE,VAD flag. */
long so_etsi_features(const int16_t *in, long n, float *feat, int16_t *den_i16);
int so_vad_flush(so_vad *v, float *fb15, int *frame_counter);                   /* VAD.c:342-431    */
/* flags: bit 0 = 16 kHz mode, bit 1 = FlushAdvProcess at the end (up to 6 more rows) */
long so_etsi_features_ex(const int16_t *in, long n, float *feat, int16_t *den_i16, int flags);
/* 16 kHz mode (etsi/cpp/16kHzProcessing.c + its hooks in NoiseSup.c / CompCeps.c) */
int so_ns_attach_hp16k(so_ns *s);
void so_qmf_taps(float *dp118, float *hp118);
void so_cc_dct26(float *dct12x26);
void so_hp16k_tables(const so_ns *s, int *start5, int *len5, float *fb5x65);
void so_qmf_split(const int16_t *in, long n, float *lp, float *hp);
long so_etsi_features_16k(const int16_t *in, long n, float *feat, int16_t *den_i16);

/* ---------- P16 plugin entry point ---------- */
/* etsi_denoise_mapping_func_Wiener (aurora_etsi/NoiseSup.cpp:1140-1407) on an existing stream.
 * out / flag arrays follow the reference contract (caller zeroes; untouched where nothing ran).
 * gains[row][25] gets one row per stage-2 frame; returns number of rows. */
int so_p16_func_wiener(so_ns *s, const float *in, int n, float *out, int *sp_var, int *sp_spec,
                       int *sp_mel, int *sp_vadns, int *frame_counter, float *gains);

/* ---------- G64: gammatone / hair cell / sub-bands / masks / resynthesis ---------- */
void so_channel_params(float *cf64, float *bw64, float *mid64);  /* extractwav.cpp:41-54,134-167,259-279 */
void so_gammatone(const float *in, float *out, float cf, float bw, float mid, long n); /* :168-212 */
void so_haircell(const float *in, float *out, long n);                                 /* :213-258 */
void so_subband(const int16_t *wav, long n, int16_t *out64xn);  /* subband .../extractwav.cpp:36-97 */
void so_addnoise(const int16_t *pure, int16_t *noise, int16_t *noisy, long n, int db); /* :6-35    */
#define SO_RESYNTH_IBM  0     /* resyth_64sub_IBM/cpp/extractwav.cpp:95-108 : gate m>0.5, weight 1 */
#define SO_RESYNTH_ORI  1     /* resyth_64sub_ori/cpp/extractwav.cpp:95-108 : gate m>0, weight m   */
#define SO_RESYNTH_1DNN 2     /* 1dnn_resynth/extractwav.cpp:67        : ORI with numFrame=L/160   */
int so_resynth_num_frames(long n, int mode);
void so_resynth(const int16_t *wav, long n, float *mask /*[numFrame][64], IBM mode mutates*/,
                int num_frame, int mode, int16_t *out, float *out_f32 /*optional*/);
/* make_single_IBM restated (enhancement_extract_subband_linux/cpp/show_IBM.cpp:92-213) on top of OUR
 * restatement of the un-vendored asdk::SpecInfo — PARITY UNPINNED at that boundary.
 * pure64/noise64 are channel-major [64][n] int16 sub-band signals; irm/ibm are [frames][64]. */
long so_ibm_num_frames(long n);
void so_single_ibm(const int16_t *pure64, const int16_t *noise64, long n, float *irm, uint8_t *ibm);

#ifdef __cplusplus
}
#endif
#endif
