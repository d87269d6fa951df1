/* TEST INFRASTRUCTURE — not part of the product.
 *
 * oracle/_ref driver #1: the reference's own ETSI front end (etsi/cpp/*.c), compiled IN PLACE from
 * /root/reference (never copied), exposed to the tests through a few plain-C entry points.
 *
 * This translation unit textually includes the reference's NoiseSup.c so that the private state
 * struct (etsi/cpp/NoiseSup.c:49-143) and the static helpers (FFTtoPSD :249, PSDMean :289, ...) are
 * visible to the trace driver below.  Everything that computes is reference code; the only code of
 * ours here is the calling loop, which restates
 *   - etsi_denoise's read/process/copy loop          (etsi/cpp/AdvFrontEnd.c:173-191), and
 *   - the CompCeps step of DoAdvProcess that the reference ships commented out
 *                                                     (etsi/cpp/ParmInterface.c:274-311, minus
 *                                                      WaveProc / PostProc / VAD),
 * so the NoiseSup -> CompCeps chain of north_star can be pinned against the reference's functions.
 */
#include "NoiseSup.c" /* resolved through -I/root/reference/etsi/cpp */
#include "AdvFrontEnd.h"
#include "CompCepsExports.h"
#include <string.h>
#include <unistd.h>
#include <fcntl.h>

/* ---- plain pass-throughs to reference entry points ---------------------------------------- */

/* etsi/cpp/AdvFrontEnd.c:125.  The reference prints "NO SPEECH DETECTED" on every call
 * (AdvFrontEnd.c:202-203); silence stderr around it when quiet!=0 so timing loops stay clean. */
int ref_etsi_denoise(short *in, short *out, long n, int quiet)
{
    int saved = -1, rc;
    if (quiet) {
        fflush(stderr);
        saved = dup(2);
        int nul = open("/dev/null", O_WRONLY);
        dup2(nul, 2);
        close(nul);
    }
    rc = etsi_denoise(in, out, n);
    if (quiet) {
        fflush(stderr);
        dup2(saved, 2);
        close(saved);
    }
    return rc;
}

int ref_etsi_denoise_synchronization(short *in, short *out, long n)
{
    return etsi_denoise_synchronization(in, out, n);
}

/* etsi/cpp/rfft.c:45 */
void ref_rfft(float *x, int n, int m) { rfft(x, n, m); }

/* static helpers of NoiseSup.c, for known-answer tests */
void ref_fft_to_psd(float *fft_in, float *psd_out) { FFTtoPSD(fft_in, psd_out, NS_FFT_LENGTH); }
void ref_apply_wf(float *data, float *predata, float *filter, float *result)
{
    ApplyWF(data, predata, filter, result, NS_FRAME_SHIFT, NS_HALF_FILTER_LENGTH);
}
void ref_dc_offset(float *data, float *state2, int n)
{
    DC_FILTER f;
    f.lastSampleIn = state2[0];
    f.lastDCOut = state2[1];
    DCOffsetFil(data, &f, (X_INT16)n);
    state2[0] = f.lastSampleIn;
    state2[1] = f.lastDCOut;
}

/* Tables the reference builds at init time (NoiseSup.c:974-1008, MelProc.c:128-337,
 * CompCeps.c:87-94,153-173, MelProc.c:402-522), dumped so the restatement's tables can be pinned. */
int ref_etsi_tables(float *sig_window /*200*/, float *ir_window /*17*/, int *mel_start /*25*/,
                    int *mel_len /*25*/, float *mel_data /*25*65*/, float *idct /*25*25*/,
                    float *hamming /*100*/, float *dct /*12*23*/, int *cc_start /*23*/,
                    int *cc_len /*23*/, float *cc_data /*23*129*/)
{
    FEParamsX *fe = AdvProcessAlloc(8000);
    int i, j;
    MelFB_Window *w;
    AdvProcessInit(fe);
    memcpy(sig_window, fe->NSX->nsTmp.sigWindow, sizeof(float) * NS_FRAME_LENGTH);
    memcpy(ir_window, fe->NSX->nsTmp.IRWindow, sizeof(float) * NS_FILTER_LENGTH);
    for (w = fe->NSX->nsTmp.FirstWindow, i = 0; w; w = w->Next, i++) {
        mel_start[i] = w->StartingPoint;
        mel_len[i] = w->Length;
        for (j = 0; j < 65; j++) mel_data[i * 65 + j] = 0.0f;
        for (j = 0; j < w->Length; j++) mel_data[i * 65 + j] = w->Data[j];
    }
    for (i = 0; i < WF_MEL_ORDER; i++)
        for (j = 0; j < WF_MEL_ORDER; j++) idct[i * WF_MEL_ORDER + j] = fe->NSX->nsTmp.melIDCTbasis[i][j];
    {
        /* CompCepsStructX is private to CompCeps.c; mirror its leading fields (CompCeps.c:45-54). */
        struct cc_view {
            X_INT16 SamplingFrequency;
            X_FLOAT32 StartingFrequency;
            X_FLOAT32 HammingWindow[200];
            X_FLOAT32 *pDCTMatrix;
            MelFB_Window *FirstWindow;
        } *cc = (struct cc_view *)fe->CCX;
        memcpy(hamming, cc->HammingWindow, sizeof(float) * 100);
        memcpy(dct, cc->pDCTMatrix, sizeof(float) * 12 * 23);
        for (w = cc->FirstWindow, i = 0; w; w = w->Next, i++) {
            cc_start[i] = w->StartingPoint;
            cc_len[i] = w->Length;
            for (j = 0; j < 129; j++) cc_data[i * 129 + j] = 0.0f;
            for (j = 0; j < w->Length; j++) cc_data[i * 129 + j] = w->Data[j];
        }
    }
    AdvProcessDelete(&fe);
    return 0;
}

/* ---- trace driver -------------------------------------------------------------------------- */

typedef struct {
    /* per 80-sample call, length n_calls */
    int *ret;            /* 1 if DoNoiseSup produced a frame this call, 0 if latency, -1 zero frame */
    int *flag_vad, *hang_over, *nb_speech, *nb_frame0, *nb_frame1;
    int *sp_var, *sp_spec, *sp_mel, *sp_vadns;
    float *mean_en, *alfa_gf, *low_snr, *den_en;       /* den_en: denEn1[2]                       */
    float *nsig1, *noise1, *den1, *nsig2, *noise2, *den2; /* [n_calls][65], state after the call  */
    float *out_f32;      /* [n_calls][80] float denoised shift (valid where ret==1)                */
} ref_etsi_trace_t;

/* Runs the reference front end over `n` samples exactly like etsi_denoise (AdvFrontEnd.c:173-191),
 * but through our own copy of the DoAdvProcess body with the commented-out CompCeps step re-enabled
 * (ParmInterface.c:274-293) when ceps!=NULL.  Returns the number of 80-sample calls made;
 * *n_ceps receives the number of cepstral frames written to ceps[frame][14].                    */
long ref_etsi_trace(const short *in, long n, short *den_i16, float *ceps, long *n_ceps,
                    ref_etsi_trace_t *tr)
{
    FEParamsX *fe = AdvProcessAlloc(8000);
    X_FLOAT32 frameBuf[FRAME_BUF_SIZE + 8];
    X_FLOAT32 feat[NUM_CEP_COEFF + 2];
    short den_buf[80];
    long calls = 0, start = 0, nc = 0;
    int i;
    fe->Noc0 = 0;
    AdvProcessInit(fe);
    memset(den_buf, 0, sizeof den_buf); /* calloc'ed in the reference, AdvFrontEnd.c:161 */
    while (start + 80 <= n) {
        /* --- DoAdvProcess body, ParmInterface.c:232-328 --- */
        X_FLOAT32 *pt = BufInShiftToPut(fe->denoisedBuf, fe->FrameShift);
        float check = 0.0f;
        int ret;
        for (i = 0; i < 80; i++) {
            fe->CurFrame[i] = (X_FLOAT32)in[start + i];
            check += in[start + i] * in[start + i];
        }
        if (((int)check != 0) || (fe->NonZeroFrameOnset != 0)) {
            fe->NonZeroFrameOnset = 1;
            ret = fe->DoNoiseSup(fe->CurFrame, pt, fe) ? 1 : 0;
            if (ret) {
                for (i = 0; i < 80; i++) den_buf[i] = (FILE_TYPE)pt[i];
                if (ceps) {
                    if (fe->offsetDenoisedFrame < 0) fe->offsetDenoisedFrame += fe->FrameShift;
                    if (fe->offsetDenoisedFrame >= 0) {
                        BufInGetLast(fe->denoisedBuf, frameBuf,
                                     fe->FrameLength + fe->offsetDenoisedFrame + 1);
                        fe->DoCompCeps(frameBuf + 1, feat, fe);
                        memcpy(ceps + nc * 14, feat, sizeof(float) * 14);
                        nc++;
                    }
                }
            }
        } else {
            fe->ZeroFrameCounter++;
            ret = -1;
        }
        /* --- etsi_denoise copies DenoiseBuffer out on every call, AdvFrontEnd.c:186-190 --- */
        if (den_i16) memcpy(den_i16 + start, den_buf, sizeof den_buf);
        if (tr) {
            NoiseSupStructX *N = fe->NSX;
            long c = calls;
            tr->ret[c] = ret;
            tr->flag_vad[c] = N->nsVar.vadNS.flagVAD;
            tr->hang_over[c] = N->nsVar.vadNS.hangOver;
            tr->nb_speech[c] = N->nsVar.vadNS.nbSpeechFrames;
            tr->nb_frame0[c] = N->nsVar.vadNS.nbFrame[0];
            tr->nb_frame1[c] = N->nsVar.vadNS.nbFrame[1];
            tr->sp_var[c] = fe->SpeechFoundVar;
            tr->sp_spec[c] = fe->SpeechFoundSpec;
            tr->sp_mel[c] = fe->SpeechFoundMel;
            tr->sp_vadns[c] = fe->SpeechFoundVADNS;
            tr->mean_en[c] = N->nsVar.vadNS.meanEn;
            tr->alfa_gf[c] = N->nsVar.gainFact.alfaGF;
            tr->low_snr[c] = N->nsVar.gainFact.lowSNRtrack;
            tr->den_en[c] = N->nsVar.gainFact.denEn1[2];
            memcpy(tr->nsig1 + c * 65, N->nsVar.spectrum.nSigSE1, 65 * sizeof(float));
            memcpy(tr->noise1 + c * 65, N->nsVar.spectrum.noiseSE1, 65 * sizeof(float));
            memcpy(tr->den1 + c * 65, N->nsVar.spectrum.denSigSE1, 65 * sizeof(float));
            memcpy(tr->nsig2 + c * 65, N->nsVar.spectrum.nSigSE2, 65 * sizeof(float));
            memcpy(tr->noise2 + c * 65, N->nsVar.spectrum.noiseSE2, 65 * sizeof(float));
            memcpy(tr->den2 + c * 65, N->nsVar.spectrum.denSigSE2, 65 * sizeof(float));
            if (ret == 1) memcpy(tr->out_f32 + c * 80, pt, 80 * sizeof(float));
            else memset(tr->out_f32 + c * 80, 0, 80 * sizeof(float));
        }
        calls++;
        start += 80;
    }
    AdvProcessDelete(&fe);
    if (n_ceps) *n_ceps = nc;
    return calls;
}

/* Stand-alone CompCeps on one 200-sample frame (Data[-1] must be valid), CompCeps.c:309 */
void ref_comp_ceps(const float *frame201, float *coef14)
{
    FEParamsX *fe = AdvProcessAlloc(8000);
    float tmp[201];
    AdvProcessInit(fe);
    memcpy(tmp, frame201, sizeof tmp);
    fe->DoCompCeps(tmp + 1, coef14, fe);
    AdvProcessDelete(&fe);
}

/* ---- full feature pipeline ------------------------------------------------------------------ */
#include "WaveProcExports.h"

/* DoWaveProc on one 200-sample frame, in place (etsi/cpp/WaveProc.c:386) */
void ref_wave_proc(float *frame200)
{
    FEParamsX *fe = AdvProcessAlloc(8000);
    AdvProcessInit(fe);
    fe->DoWaveProc(frame200, fe);
    AdvProcessDelete(&fe);
}

/* The DoAdvProcess body with the WHOLE commented-out block re-enabled (ParmInterface.c:274-311:
 * DoWaveProc -> DoCompCeps -> DoPostProc -> DoVADProc) and the zero-frame branch (:314-328), driven
 * like etsi_denoise's loop (AdvFrontEnd.c:173-191).  FEParamsX.FrameCounter, which gates DoVADProc
 * (VAD.c:256), is maintained by DoNoiseSup itself (NoiseSup.c:1366: stage-1 VAD frame count).
 * feat: [n/80][15] = c1..c12, c0, logE (post-processed), VAD flag.  Returns the number of rows written. */
static int g_flush = 0;      /* ref_etsi_features_ex sets it: run FlushAdvProcess after the last call */
long ref_etsi_features(const short *in, long n, float *feat, short *den_i16)
{
    FEParamsX *fe = AdvProcessAlloc(8000);
    X_FLOAT32 frameBuf[FRAME_BUF_SIZE + 8];
    X_FLOAT32 fb[NUM_CEP_COEFF + 3];
    short den_buf[80];
    long start = 0, rows = 0;
    int i;
    fe->Noc0 = 0;
    AdvProcessInit(fe);
    memset(den_buf, 0, sizeof den_buf);
    memset(fb, 0, sizeof fb);
    while (start + 80 <= n) {
        X_FLOAT32 *pt = BufInShiftToPut(fe->denoisedBuf, fe->FrameShift);
        float check = 0.0f;
        int out = 0;
        for (i = 0; i < 80; i++) {
            fe->CurFrame[i] = (X_FLOAT32)in[start + i];
            check += in[start + i] * in[start + i];
        }
        if (((int)check != 0) || (fe->NonZeroFrameOnset != 0)) {
            fe->NonZeroFrameOnset = 1;
            if (fe->DoNoiseSup(fe->CurFrame, pt, fe)) {
                for (i = 0; i < 80; i++) den_buf[i] = (FILE_TYPE)pt[i];
                if (fe->offsetDenoisedFrame < 0) fe->offsetDenoisedFrame += fe->FrameShift;
                if (fe->offsetDenoisedFrame >= 0) {
                    BufInGetLast(fe->denoisedBuf, frameBuf, fe->FrameLength + fe->offsetDenoisedFrame + 1);
                    fe->DoWaveProc(frameBuf + 1, fe);
                    fe->DoCompCeps(frameBuf + 1, fb, fe);
                    fe->DoPostProc(fb, fe);
                    out = fe->DoVADProc(fb, fe) ? 1 : 0;
                }
            }
        } else {
            fe->ZeroFrameCounter++;
            for (i = 0; i <= NUM_CEP_COEFF; i++) fb[i] = 0.0;
            fe->VAD = NON_SPEECH_FRAME;
            out = 1;
        }
        if (out) {
            memcpy(feat + rows * 15, fb, sizeof(float) * 15);
            rows++;
        }
        if (den_i16) memcpy(den_i16 + start, den_buf, sizeof den_buf);
        start += 80;
    }
    while (g_flush && FlushAdvProcess(fb, fe)) { memcpy(feat + rows * 15, fb, sizeof(float) * 15); rows++; }
    AdvProcessDelete(&fe);
    return rows;
}

/* ---- 16 kHz mode ----------------------------------------------------------- */
#include "16kHzProcExports.h"

/* Same loop at SamplingFrequency 16000 (AdvProcessAlloc :99-109): 160 samples per call, QMF split by
 * Do16kProcessing (ParmInterface.c:259-260), NoiseSup on the 0-4 kHz half, HP-band coding inside DoNoiseSup
 * and decoding/merging inside DoCompCeps (26 log bands).  den_i16: the 80 denoised LOW-BAND samples of each
 * call, [n/160][80] (what etsi_denoise_16k's DenoiseBuffer really holds, AdvFrontEnd.c:264).
 * feat as above.  Returns the number of rows. */
long ref_etsi_features_16k(const short *in, long n, float *feat, short *den_i16)
{
    FEParamsX *fe = AdvProcessAlloc(16000);
    X_FLOAT32 frameBuf[FRAME_BUF_SIZE + 8];
    X_FLOAT32 fb[NUM_CEP_COEFF + 3];
    short den_buf[80];
    long start = 0, rows = 0, call = 0;
    int i;
    fe->Noc0 = 0;
    AdvProcessInit(fe);
    memset(den_buf, 0, sizeof den_buf);
    memset(fb, 0, sizeof fb);
    while (start + 160 <= n) {
        X_FLOAT32 *pt = BufInShiftToPut(fe->denoisedBuf, fe->FrameShift);
        float check = 0.0f;
        int out = 0;
        for (i = 0; i < 160; i++) {
            fe->CurFrame[i] = (X_FLOAT32)in[start + i];
            check += in[start + i] * in[start + i];
        }
        if (((int)check != 0) || (fe->NonZeroFrameOnset != 0)) {
            fe->NonZeroFrameOnset = 1;
            fe->Do16kProcessing(fe->CurFrame, fe->pData16k, 160);
            if (fe->DoNoiseSup(fe->CurFrame, pt, fe)) {
                for (i = 0; i < 80; i++) den_buf[i] = (FILE_TYPE)pt[i];
                if (fe->offsetDenoisedFrame < 0) fe->offsetDenoisedFrame += fe->FrameShift;
                if (fe->offsetDenoisedFrame >= 0) {
                    BufInGetLast(fe->denoisedBuf, frameBuf, fe->FrameLength + fe->offsetDenoisedFrame + 1);
                    fe->DoWaveProc(frameBuf + 1, fe);
                    fe->DoCompCeps(frameBuf + 1, fb, fe);
                    fe->DoPostProc(fb, fe);
                    out = fe->DoVADProc(fb, fe) ? 1 : 0;
                }
            }
        } else {
            fe->ZeroFrameCounter++;
            for (i = 0; i <= NUM_CEP_COEFF; i++) fb[i] = 0.0;
            fe->VAD = NON_SPEECH_FRAME;
            out = 1;
        }
        if (out) {
            memcpy(feat + rows * 15, fb, sizeof(float) * 15);
            rows++;
        }
        if (den_i16) memcpy(den_i16 + call * 80, den_buf, sizeof den_buf);
        start += 160;
        call++;
    }
    while (g_flush && FlushAdvProcess(fb, fe)) { memcpy(feat + rows * 15, fb, sizeof(float) * 15); rows++; }
    AdvProcessDelete(&fe);
    return rows;
}

/* flags: bit 0 = 16 kHz mode, bit 1 = flush (the ETSI driver loop: while (FlushAdvProcess(FeatureBuffer, fe)) write) */
long ref_etsi_features_ex(const short *in, long n, float *feat, short *den_i16, int flags)
{
    long rows;
    g_flush = (flags >> 1) & 1;
    rows = (flags & 1) ? ref_etsi_features_16k(in, n, feat, den_i16) : ref_etsi_features(in, n, feat, den_i16);
    g_flush = 0;
    return rows;
}

/* Do16kProcessing alone (16kHzProcessing.c:711): x[n] (n a multiple of 160) -> lp[n/2], hp[n/2], streaming. */
void ref_qmf_split(const short *in, long n, float *lp, float *hp)
{
    FEParamsX *fe = AdvProcessAlloc(16000);
    long start;
    int i;
    AdvProcessInit(fe);
    for (start = 0; start + 160 <= n; start += 160) {
        float buf[160];
        for (i = 0; i < 160; i++) buf[i] = (float)in[start + i];
        Do16kProcessing(buf, fe->pData16k, 160);
        for (i = 0; i < 80; i++) {
            lp[start / 2 + i] = buf[i];
            hp[start / 2 + i] = Get16k_dataHP(fe->pData16k, i);
        }
    }
    AdvProcessDelete(&fe);
}

/* ---- DoAdvProcess / FlushAdvProcess exactly as shipped (the feature block of ParmInterface.c:274-311 stays commented
 * out): one call per 80-sample frame, then the ETSI driver's flush loop.  rets[n_calls]: DoAdvProcess's return value
 * per call (the success path falls off the end of the function without a return, ParmInterface.c:262-313, so -- like
 * the reference's own callers -- only "did it write DenoiseBuffer" is recorded: out80 is pre-filled with a sentinel);
 * out[n_calls*80] / feat[n_calls*15] hold what the call left in DenoiseBuffer / FeatureBuffer (feat pre-filled with
 * the sentinel 7.0f).  flush_rets[8] / flush_feat[8*15]: return value and FeatureBuffer of up to 8 FlushAdvProcess
 * calls.  Returns the number of flush calls made (the first FALSE included). */
int ref_adv_process_as_shipped(const short *in, long n, short *out, float *feat, int *wrote, int *flush_rets,
                               float *flush_feat)
{
    FEParamsX *fe = AdvProcessAlloc(8000);
    long calls = n / 80, c;
    int k, i;
    short frame[80], den[80];
    float fb[15];
    AdvProcessInit(fe);
    for (c = 0; c < calls; c++) {
        memcpy(frame, in + c * 80, sizeof frame);
        for (i = 0; i < 80; i++) den[i] = (short)0x7a7a;
        for (i = 0; i < 15; i++) fb[i] = 7.0f;
        (void)DoAdvProcess(frame, den, fb, fe);
        wrote[c] = 0;
        for (i = 0; i < 80; i++) if (den[i] != (short)0x7a7a) wrote[c] = 1;
        memcpy(out + c * 80, den, sizeof den);
        memcpy(feat + c * 15, fb, sizeof fb);
    }
    for (k = 0; k < 8; k++) {
        for (i = 0; i < 15; i++) fb[i] = 7.0f;
        flush_rets[k] = FlushAdvProcess(fb, fe) ? 1 : 0;
        memcpy(flush_feat + k * 15, fb, sizeof fb);
        if (!flush_rets[k]) { k++; break; }
    }
    AdvProcessDelete(&fe);
    return k;
}
