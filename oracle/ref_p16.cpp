/* TEST INFRASTRUCTURE — not part of the product.
 *
 * oracle/_ref driver #2: the reference's C++ plugin-API NoiseSup at 16 kHz frame sizes
 * (function/20141106_speech_enhancement/aurora_etsi/{NoiseSup,MelProc,rfft}.cpp), compiled IN PLACE
 * from /root/reference with oracle/shims/aurora/aurora_include.h on the include path.
 *
 * NoiseSup.cpp is textually included so the private NoiseSupStructX is visible for state dumps;
 * MelProc.cpp and rfft.cpp are compiled as their own translation units (oracle/Makefile).
 * All arithmetic is reference code; ours is only the marshalling below.
 */
#include "NoiseSup.cpp" /* resolved through -I.../aurora_etsi */
#include <string.h>
#include <iostream>
#include <stdlib.h>
#include <unistd.h>
#include <fcntl.h>

extern "C" {

typedef struct {
    INSTANCE glb;
    INSTANCE thd;
} ref_p16_t;

/* etsi_denoise_mapping_global_init + _thread_init, NoiseSup.cpp:913-1084.
 * InitGammaIDCTbasis prints the centre frequencies to stdout (MelProc.cpp:483-485): silenced. */
void *ref_p16_create(void)
{
    ref_p16_t *h = (ref_p16_t *)calloc(1, sizeof *h);
    fflush(stdout);
    int saved = dup(1);
    int nul = open("/dev/null", O_WRONLY);
    dup2(nul, 1);
    close(nul);
    etsi_denoise_mapping_global_init(&h->glb, NULL);
    etsi_denoise_mapping_thread_init(&h->thd, h->glb);
    std::cout.flush();
    fflush(stdout);
    dup2(saved, 1);
    close(saved);
    return h;
}

void ref_p16_destroy(void *hh)
{
    ref_p16_t *h = (ref_p16_t *)hh;
    etsi_denoise_mapping_thread_release(&h->thd);
    etsi_denoise_mapping_global_release(&h->glb);
    free(h);
}

/* etsi_denoise_mapping_func_Wiener, NoiseSup.cpp:1140-1407.
 * `out` and the five int arrays must be zeroed by the caller (reference contract, NoiseSup.cpp:1136).
 * The 25 stage-2 gains per processed block that the reference fprintf's with "%f " are captured
 * through open_memstream and parsed back into gains[row][25]; *n_rows = rows written.
 * If text_out!=NULL the raw text (at most text_cap-1 bytes) is also returned for byte comparison. */
int ref_p16_func_wiener(void *hh, float *in, int n, float *out, int *sp_var, int *sp_spec,
                        int *sp_mel, int *sp_vadns, int *frame_counter, float *gains, int *n_rows,
                        char *text_out, long text_cap)
{
    ref_p16_t *h = (ref_p16_t *)hh;
    esti_denoise_in di;
    esti_denoise_out dout;
    char *buf = NULL;
    size_t len = 0;
    FILE *fp = open_memstream(&buf, &len);
    di.inData = in;
    di.dataNum = n;
    dout.outData = out;
    dout.pSpeechFoundVar = sp_var;
    dout.pSpeechFoundSpec = sp_spec;
    dout.pSpeechFoundMel = sp_mel;
    dout.pSpeechFoundVADNS = sp_vadns;
    dout.pFrameCounter = frame_counter;
    int rc = etsi_denoise_mapping_func_Wiener(h->glb, h->thd, &di, &dout, (INSTANCE)fp);
    fclose(fp);
    int rows = 0;
    if (gains) {
        char *p = buf;
        while (p && *p) {
            char *e = strchr(p, '\n');
            if (!e) break;
            for (int k = 0; k < 25; k++) gains[rows * 25 + k] = strtof(p, &p);
            rows++;
            p = e + 1;
        }
    } else {
        for (size_t i = 0; i < len; i++) rows += (buf[i] == '\n');
    }
    if (n_rows) *n_rows = rows;
    if (text_out && text_cap > 0) {
        size_t c = len < (size_t)(text_cap - 1) ? len : (size_t)(text_cap - 1);
        memcpy(text_out, buf, c);
        text_out[c] = 0;
    }
    free(buf);
    return rc;
}

/* Dump of the per-stream state after the last call, for per-block traces (call with n=160). */
void ref_p16_state(void *hh, float *nsig1, float *noise1, float *den1, float *nsig2, float *noise2,
                   float *den2, float *scal /*8*/, int *iscal /*8*/)
{
    NoiseSupStructX *N = (NoiseSupStructX *)((ref_p16_t *)hh)->thd;
    memcpy(nsig1, N->nsVar.spectrum.nSigSE1, 129 * sizeof(float));
    memcpy(noise1, N->nsVar.spectrum.noiseSE1, 129 * sizeof(float));
    memcpy(den1, N->nsVar.spectrum.denSigSE1, 129 * sizeof(float));
    memcpy(nsig2, N->nsVar.spectrum.nSigSE2, 129 * sizeof(float));
    memcpy(noise2, N->nsVar.spectrum.noiseSE2, 129 * sizeof(float));
    memcpy(den2, N->nsVar.spectrum.denSigSE2, 129 * sizeof(float));
    scal[0] = N->nsVar.vadNS.meanEn;
    scal[1] = N->nsVar.gainFact.alfaGF;
    scal[2] = N->nsVar.gainFact.lowSNRtrack;
    scal[3] = N->nsVar.gainFact.denEn1[0];
    scal[4] = N->nsVar.gainFact.denEn1[1];
    scal[5] = N->nsVar.gainFact.denEn1[2];
    scal[6] = N->nsVar.prevSamples.lastSampleIn;
    scal[7] = N->nsVar.prevSamples.lastDCOut;
    iscal[0] = N->nsVar.vadNS.flagVAD;
    iscal[1] = N->nsVar.vadNS.hangOver;
    iscal[2] = N->nsVar.vadNS.nbSpeechFrames;
    iscal[3] = N->nsVar.vadNS.nbFrame[0];
    iscal[4] = N->nsVar.vadNS.nbFrame[1];
    iscal[5] = N->nsVar.buffers.nbFramesInFirstStage;
    iscal[6] = N->nsVar.buffers.nbFramesInSecondStage;
    iscal[7] = N->nsVar.buffers.nbFramesOutSecondStage;
}

/* Tables built by etsi_denoise_mapping_thread_init (NoiseSup.cpp:1030-1080, MelProc.cpp:268-341,
 * 463-504): Hanning-480, Hanning-17, Gamma bank 25x128 (+StartingPoint), Gamma-IDCT 25x25. */
void ref_p16_tables(void *hh, float *sig_window /*480*/, float *ir_window /*17*/,
                    float *gamma /*25*128*/, int *gamma_start /*25*/, float *idct /*25*25*/)
{
    NoiseSupStructX *N = (NoiseSupStructX *)((ref_p16_t *)hh)->thd;
    memcpy(sig_window, N->nsTmp.sigWindow, 480 * sizeof(float));
    memcpy(ir_window, N->nsTmp.IRWindow, 17 * sizeof(float));
    int i = 0;
    for (Gamma_Window *w = N->nsTmp.FirstWindow; w; w = w->Next, i++) {
        gamma_start[i] = w->StartingPoint;
        memcpy(gamma + i * 128, w->Data, 128 * sizeof(float));
    }
    for (i = 0; i < 25; i++) memcpy(idct + i * 25, N->nsTmp.GammaIDCTbasis[i], 25 * sizeof(float));
}

/* rfft.cpp:45 — same split-radix code; called with (512, 8) by the plugin (Q1 in SURVEY.md). */
void ref_p16_rfft(float *x, int n, int m) { rfft(x, n, m); }

} /* extern "C" */
