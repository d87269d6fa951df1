/* TEST INFRASTRUCTURE — the parity oracle (see se_oracle.h).  NOT part of the product.
 *
 * Plain-C CPU restatement of the reference hot path.  It is written to evaluate every float /
 * double sub-expression with the same promotions and in the same order as the reference sources
 * (SURVEY.md Q5), so that — compiled with -O2 -ffp-contract=off like oracle/_ref — it reproduces
 * the reference's outputs bit for bit.  tests/test_oracle_vs_ref.py checks exactly that wherever
 * /root/reference is present, and tests/golden/ holds vectors generated from oracle/_ref for the
 * GPU box, where it is not.
 *
 * PARITY STATUS: pinned against oracle/_ref (the reference's own code) for everything except
 * so_single_ibm(), whose spectral front end (asdk::SpecInfo) is un-vendored: "parity unpinned"
 * for the IRM values; the frame count and the bins summed are pinned by show_IBM.cpp:110,156.
 */
#include "se_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define SO_PI   3.14159265358979323846
#define SO_PIx2 6.28318530717958647692
#define SO_SQRT2 1.41421356237309504880

/* =====================================================================================
 * Split-radix real FFT, in place.  Follows etsi/cpp/rfft.c:45-180 (Sorensen et al. 1987):
 * output order Re(0..n/2), Im(n/2-1..1).  `m` is the number of butterfly levels actually run;
 * the P16 plugin calls it with n=512, m=8 (one level short) and that behaviour is
 * preserved because the level loop is bounded by m, not by log2(n).
 * ===================================================================================== */
static void so_bitrev(float *x, int n)
{
    int i, j = 0, k;
    for (i = 0; i < n - 1; i++) {
        if (i < j) { float t = x[j]; x[j] = x[i]; x[i] = t; }
        k = n >> 1;
        while (k <= j) { j -= k; k >>= 1; }
        j += k;
    }
}

/* The ETSI files are C (cos/sin evaluate in double, rfft.c:136-139); the plugin copy rfft.cpp is
 * C++, where cos(float) resolves to the float overload.  float_trig selects the latter. */
static void so_rfft_impl(float *x, int n, int m, int float_trig)
{
    int level, is, id, i, j;
    so_bitrev(x, n);                                   /* rfft.c:56-75  */
    for (is = 0, id = 4; is < n - 1; is = 2 * id - 2, id *= 4)   /* rfft.c:77-92 length-2 */
        for (i = is; i < n; i += id) {
            float a = x[i];
            x[i] = a + x[i + 1];
            x[i + 1] = a - x[i + 1];
        }
    for (level = 1; level < m; level++) {              /* rfft.c:94-179 L-shaped      */
        const int n2 = 2 << level, n4 = n2 >> 2, n8 = n2 >> 3;
        const float e = (float)((SO_PI * 2) / n2);
        for (is = 0, id = 2 * n2; is < n; is = 2 * id - n2, id *= 4)
            for (i = is; i < n; i += id) {
                float *p1 = x + i, *p2 = p1 + n4, *p3 = p2 + n4, *p4 = p3 + n4;
                float t1 = *p4 + *p3;
                *p4 -= *p3;
                *p3 = *p1 - t1;
                *p1 += t1;
                if (n4 != 1) {
                    float t2;
                    p1 += n8; p2 += n8; p3 += n8; p4 += n8;
                    t1 = (float)((*p3 + *p4) / SO_SQRT2);
                    t2 = (float)((*p3 - *p4) / SO_SQRT2);
                    *p4 = *p2 - t1;
                    *p3 = -*p2 - t1;
                    *p2 = *p1 - t2;
                    *p1 = *p1 + t2;
                }
            }
        for (j = 1; j < n8; j++) {
            const float a = j * e, a3 = 3 * a;
            const float c1 = float_trig ? cosf(a) : (float)cos(a), s1 = float_trig ? sinf(a) : (float)sin(a);
            const float c3 = float_trig ? cosf(a3) : (float)cos(a3), s3 = float_trig ? sinf(a3) : (float)sin(a3);
            for (is = 0, id = 2 * n2; is < n; is = 2 * id - n2, id *= 4)
                for (i = is; i < n; i += id) {
                    float *q1 = x + i + j, *q2 = q1 + n4, *q3 = q2 + n4, *q4 = q3 + n4;
                    float *q5 = x + i + n4 - j, *q6 = q5 + n4, *q7 = q6 + n4, *q8 = q7 + n4;
                    float t1 = *q3 * c1 + *q7 * s1;
                    float t2 = *q7 * c1 - *q3 * s1;
                    float t3 = *q4 * c3 + *q8 * s3;
                    float t4 = *q8 * c3 - *q4 * s3;
                    float t5 = t1 + t3, t6 = t2 + t4;
                    t3 = t1 - t3;
                    t4 = t2 - t4;
                    t2 = *q6 + t6;
                    *q3 = t6 - *q6;
                    *q8 = t2;
                    t2 = *q2 - t3;
                    *q7 = -*q2 - t3;
                    *q4 = t2;
                    t1 = *q1 + t5;
                    *q6 = *q1 - t5;
                    *q1 = t1;
                    t1 = *q5 + t4;
                    *q5 = *q5 - t4;
                    *q2 = t1;
                }
        }
    }
}

void so_rfft(float *x, int n, int m) { so_rfft_impl(x, n, m, 0); }
void so_rfft_cxx(float *x, int n, int m) { so_rfft_impl(x, n, m, 1); }

/* =====================================================================================
 * NoiseSup stream
 * ===================================================================================== */
#define NS_FILT 17
#define NS_HALF 8
#define MAXBUF 640
#define MAXFRAME 480
#define MAXFFT 512

/* State of the 16 kHz extension (DataFor16kProc + QMF_FIR, etsi/cpp/16kHzProcessing.c:37-84), attached to a
 * P8 stream by so_ns_attach_hp16k(). */
#define HP_QMF 118
typedef struct so_hp16k {
    float dp[HP_QMF], hp[HP_QMF], hist[HP_QMF - 1];   /* QMF_FIR: low/high-pass taps, delay line */
    float data_hp[80];                               /* dataHP: this call's 4-8 kHz half, shifted to 0-4 kHz */
    float buf[480];                                  /* bufferData16k */
    float snap[256];                                 /* tmpWorkSpace16k (frame copied at stage 0, used at stage 1) */
    float hp_bands[9], code[9], code_buf[27];        /* hpBands, CodeForBands16k, bufferCodeForBands16k */
    float lp_log[9];                                 /* BandsForCoding16k: 3 frames x 3 log band energies */
    float fb16k[3];                                  /* this call's spectrally-subtracted HP bands */
    float noise[3], mean_en;
    int nb_speech, hang_over;
    int fb_start[5], fb_len[5];
    float fb[5][65];                                 /* FirstWindow16k: 5 un-normalised mel triangles over 65 bins */
} so_hp16k;

struct so_ns {
    struct so_hp16k *hp16k;
    /* profile constants: etsi/cpp/NoiseSup.h:28-79 vs aurora_etsi/NoiseSup.h:22-72 */
    int profile, S, FL, NFFT, M, SPEC, BUF, DIN, ANA, fs;
    /* BUFFERS (NoiseSup.c:92-100) */
    int nb_in1, nb_in2, nb_out2;
    float buf1[MAXBUF], buf2[MAXBUF];
    /* DC_FILTER, GAIN_FACT, VAD_DATA_NS, VAD_DATA_FD (NoiseSup.c:49-90) */
    float dc_x, dc_y;
    float den_en[3], low_snr, alfa_gf;
    int nb_frame[2];
    short flag_vad, hang_over, nb_speech;
    float mean_en;
    float fd_mel_mean, fd_var_mean, fd_acc_test, fd_spec_mean, fd_mel_val[2], fd_spec_val, fd_in_vadq;
    /* SPECTRUM (NoiseSup.c:102-115) */
    float nsig[2][SO_MAX_SPEC], noise[2][SO_MAX_SPEC], den[2][SO_MAX_SPEC];
    float psdm[2][SO_MAX_SPEC];     /* trace only: sqrt(PSDMeaned) as FilterCalc leaves it (NoiseSup.c:522-526) */
    short idx[2];
    float psd_hist[2][SO_MAX_SPEC * 2];
    /* NS_TMP tables (NoiseSup.c:130-137) */
    float sig_window[MAXFRAME], ir_window[NS_FILT];
    int bank_start[SO_MEL], bank_len[SO_MEL];
    float bank[SO_MEL][SO_MAX_SPEC];
    float idct[SO_MEL][SO_MEL];
    float eps;
};

/* --- mel bank for the Wiener filter: etsi/cpp/MelProc.c:128-230 (InitMelFBwindows, normalize=1) --- */
static void init_mel_bank(so_ns *s, float st_freq, float smpl_freq, int fft_len, int nch)
{
    int c[SO_MEL], i, j, k;
    float start_mel = 2595.0 * log10(1.0 + (float)st_freq / 700.0);
    float half_mel = 2595.0 * log10(1.0 + (smpl_freq / 2) / 700.0);
    for (i = 0; i < nch; i++) {
        float f = 700 * (pow(10, (start_mel + (float)i / (nch - 1) * (half_mel - start_mel)) / 2595.0) - 1.0);
        c[i] = (int)(fft_len * f / smpl_freq + 0.5);
    }
    for (i = 0; i < nch; i++) {
        float norm = 0.0, *d = s->bank[i];
        int len;
        memset(d, 0, sizeof s->bank[i]);
        if (i == 0) {
            s->bank_start[i] = c[0];
            len = c[1] - c[0];
            for (j = 0; j < len; j++) { d[j] = 1.0 - (float)j / (float)len; norm += d[j]; }
        } else if (i < nch - 1) {
            const int up = c[i] - c[i - 1];
            s->bank_start[i] = c[i - 1] + 1;
            len = c[i + 1] - c[i - 1] - 1;
            for (j = 0; j < up; j++) { d[j] = (float)(j + 1) / (float)up; norm += d[j]; }
            for (j = up, k = 0; j < len; j++, k++) { d[j] = 1.0 - (k + 1) / (float)(c[i + 1] - c[i]); norm += d[j]; }
        } else {
            s->bank_start[i] = c[nch - 2] + 1;
            len = c[nch - 1] - c[nch - 2];
            for (j = 0; j < len; j++) { d[j] = (float)(j + 1) / (float)len; norm += d[j]; }
        }
        s->bank_len[i] = len;
        for (j = 0; j < len; j++) d[j] /= norm;
    }
}

/* --- mel IDCT basis: etsi/cpp/MelProc.c:283-337 --- */
static void init_mel_idct(so_ns *s, int order, int samp_freq, int fft_len)
{
    float cf[SO_MEL], df[SO_MEL];
    float lin = samp_freq / (float)fft_len;
    int i, j;
    for (j = 0; j < order; j++) {
        if (j == 0) cf[j] = s->bank_start[0] * lin;
        else if (j == order - 1) cf[j] = (s->bank_start[j] + s->bank_len[j] - 1) * lin;
        else {
            float st = s->bank_start[j] * lin, sum = 0.0;
            cf[j] = 0.0;
            for (i = 0; i < s->bank_len[j]; i++) {
                cf[j] += s->bank[j][i] * (st + i * lin);
                sum += s->bank[j][i];
            }
            cf[j] /= sum;
        }
    }
    for (j = 0; j < order; j++) {
        if (j == 0) df[j] = (cf[1] - cf[0]) / samp_freq;
        else if (j == order - 1) df[j] = (cf[j] - cf[j - 1]) / samp_freq;
        else df[j] = (cf[j + 1] - cf[j - 1]) / samp_freq;
    }
    for (i = 0; i < order; i++)
        for (j = 0; j < order; j++) s->idct[i][j] = df[j] * cos(SO_PIx2 * i * cf[j] / samp_freq);
}

/* --- Gamma bank: aurora_etsi/MelProc.cpp:268-341 (+ HzToERBRate/ERBRateToHz :505-513) --- */
static float hz_to_erb(float hz) { return (21.4 * log10(hz * 0.00437 + 1.0)); }
static float erb_to_hz(float e) { return ((pow(10, e / 21.4) - 1) / 0.00437); }

static void init_gamma_bank(so_ns *s, float st_freq, float smpl_freq, int fft_len, int nch)
{
    float cfs[SO_MEL], erb[SO_MEL];
    float lo = hz_to_erb(st_freq), hi = hz_to_erb(smpl_freq / 2);
    float step = (hi - lo) / (nch - 1);
    int i, j;
    for (i = 0; i < nch; i++) {
        cfs[i] = erb_to_hz(lo + step * i);
        erb[i] = 1.019 * 24.7 * (4.37 * cfs[i] / 1000 + 1);
    }
    for (i = 0; i < nch; i++) {
        float norm = 0.0, *d = s->bank[i];
        memset(d, 0, sizeof s->bank[i]);
        s->bank_start[i] = (int)cfs[i] * fft_len / smpl_freq;
        s->bank_len[i] = fft_len / 2;
        for (j = 0; j < s->bank_len[i]; j++) {
            d[j] = 1.0 / pow((1.0 + (j * 1.0 / fft_len * smpl_freq - cfs[i]) / erb[i] *
                                    (j * 1.0 / fft_len * smpl_freq - cfs[i]) / erb[i]), 2);
            norm += d[j];
        }
        for (j = 0; j < s->bank_len[i]; j++) d[j] /= norm;
    }
}

/* --- Gamma IDCT basis: aurora_etsi/MelProc.cpp:463-504 --- */
static void init_gamma_idct(so_ns *s, int order, int samp_freq, int fft_len)
{
    float cf[SO_MEL], df[SO_MEL];
    float lin = samp_freq / (float)fft_len;
    int i, j;
    for (j = 0; j < order; j++) cf[j] = s->bank_start[j] * lin;
    for (j = 0; j < order; j++) {
        if (j == 0) df[j] = (cf[1] - cf[0]) / samp_freq;
        else if (j == order - 1) df[j] = (cf[j] - cf[j - 1]) / samp_freq;
        else df[j] = (cf[j + 1] - cf[j - 1]) / samp_freq;
    }
    for (i = 0; i < order; i++)
        for (j = 0; j < order; j++) s->idct[i][j] = df[j] * cos(SO_PIx2 * i * cf[j] / samp_freq);
}

/* DoNoiseSupInit (etsi/cpp/NoiseSup.c:884-1009) / etsi_denoise_mapping_thread_init
 * (aurora_etsi/NoiseSup.cpp:937-1084) */
so_ns *so_ns_create(int profile)
{
    so_ns *s = (so_ns *)calloc(1, sizeof *s);
    int i;
    if (!s) return NULL;
    s->profile = profile;
    if (profile == SO_PROFILE_P8) {
        s->S = 80; s->FL = 200; s->NFFT = 256; s->M = 8; s->SPEC = 65; s->BUF = 320; s->DIN = 240; s->ANA = 60;
        s->fs = 8000;
    } else {
        s->S = 160; s->FL = 480; s->NFFT = 512; s->M = 8; s->SPEC = 129; s->BUF = 640; s->DIN = 480; s->ANA = 80;
        s->fs = 16000;
    }
    s->eps = (float)exp((double)-10.0);
    s->alfa_gf = 0.8;
    for (i = 0; i < s->SPEC; i++) s->noise[0][i] = s->noise[1][i] = s->eps;
    for (i = 0; i < s->FL; i++)
        s->sig_window[i] = 0.5 - 0.5 * cos((SO_PIx2 * ((float)i + 0.5)) / (float)(s->FL));
    for (i = 0; i < NS_FILT; i++)
        s->ir_window[i] = 0.5 - 0.5 * cos((SO_PIx2 * ((float)i + 0.5)) / (float)(NS_FILT));
    if (profile == SO_PROFILE_P8) {
        init_mel_bank(s, 0.0, (float)s->fs, 2 * (s->SPEC - 1), SO_MEL);
        init_mel_idct(s, SO_MEL, s->fs, 2 * (s->SPEC - 1));
    } else {
        init_gamma_bank(s, 80.0, (float)s->fs, 2 * (s->SPEC - 1), SO_MEL);
        init_gamma_idct(s, SO_MEL, s->fs, 2 * (s->SPEC - 1));
    }
    return s;
}

void so_ns_destroy(so_ns *s) { if (s) free(s->hp16k); free(s); }
int so_ns_shift(const so_ns *s) { return s->S; }
int so_ns_spec(const so_ns *s) { return s->SPEC; }

const float *so_ns_spectrum(const so_ns *s, int which)
{
    switch (which) {
    case 0: return s->nsig[0];
    case 1: return s->noise[0];
    case 2: return s->den[0];
    case 3: return s->nsig[1];
    case 4: return s->noise[1];
    case 5: return s->den[1];
    case 6: return s->psdm[0];
    default: return s->psdm[1];
    }
}

void so_ns_tables(const so_ns *s, float *sig_window, float *ir_window, float *bank, int *bank_start,
                  int *bank_len, float *idct)
{
    int i, w = (s->profile == SO_PROFILE_P8) ? 65 : 128;
    memcpy(sig_window, s->sig_window, sizeof(float) * s->FL);
    memcpy(ir_window, s->ir_window, sizeof(float) * NS_FILT);
    for (i = 0; i < SO_MEL; i++) {
        memcpy(bank + i * w, s->bank[i], sizeof(float) * w);
        bank_start[i] = s->bank_start[i];
        bank_len[i] = s->bank_len[i];
        memcpy(idct + i * SO_MEL, s->idct[i], sizeof(float) * SO_MEL);
    }
}

/* VAD for noise suppression: etsi/cpp/NoiseSup.c:359-430 */
static void ns_vad(so_ns *s, int fstage, const float *cur)
{
    int i, nbf = s->nb_frame[fstage];
    short flag = s->flag_vad, hang = s->hang_over, nsp = s->nb_speech;
    float mean_en = s->mean_en, frame_en, lambda;
    if (nbf < 2147483647) nbf++;
    s->nb_frame[fstage] = nbf;
    if (fstage == 1) return;
    if (nbf < 10) lambda = 1 - 1 / (float)nbf;
    else lambda = (float)0.97;
    frame_en = 64.0;
    for (i = 0; i < s->S; i++) frame_en += cur[i] * cur[i];
    frame_en = (float)(0.5 + (log(frame_en / 64.0) / log(2.0)) * 16.0);
    if (((frame_en - mean_en) < 20) || (nbf < 10)) {
        if ((frame_en < mean_en) || (nbf < 10)) mean_en += (1 - lambda) * (frame_en - mean_en);
        else mean_en += (1 - (float)0.99) * (frame_en - mean_en);
        if (mean_en < (float)80.0) mean_en = (float)80.0;
    }
    if (nbf > 4) {
        if ((frame_en - mean_en) > 15) {
            flag = 1;
            nsp++;
        } else {
            if (nsp > 4) hang = 15;
            nsp = 0;
            if (hang != 0) { hang--; flag = 1; }
            else flag = 0;
        }
    }
    s->mean_en = mean_en; s->flag_vad = flag; s->hang_over = hang; s->nb_speech = nsp;
}

/* FilterCalc: etsi/cpp/NoiseSup.c:449-563.  nbFrame/flagVAD are read through an int16 copy (Q2). */
static void ns_filter_calc(so_ns *s, int fstage, float *psd_mean, float *W)
{
    float *nsig = s->nsig[fstage], *noise = s->noise[fstage], *den = s->den[fstage];
    const short nbf = (short)s->nb_frame[fstage], flag = s->flag_vad;
    const int n = s->SPEC;
    const float beta = (float)0.98, rsb_min = (float)0.079432823;
    float lambda;
    int i;
    if (fstage == 1) {
        for (i = 0; i < n; i++) noise[i] *= noise[i];
        if (nbf < 11) {
            lambda = 1 - 1 / (float)nbf;
            for (i = 0; i < n; i++) noise[i] = lambda * noise[i] + (1 - lambda) * psd_mean[i];
        } else {
            for (i = 0; i < n; i++) {
                float up = 0.9 + 0.1 * (psd_mean[i] / (psd_mean[i] + noise[i])) *
                                     (1.0 + 1.0 / (1.0 + 0.1 * (psd_mean[i] / noise[i])));
                noise[i] *= up;
            }
        }
        for (i = 0; i < n; i++) {
            noise[i] = (float)sqrt(noise[i]);
            if (noise[i] < s->eps) noise[i] = s->eps;
        }
    }
    for (i = 0; i < n; i++) {
        nsig[i] = (float)sqrt(nsig[i]);
        psd_mean[i] = (float)sqrt(psd_mean[i]);
        s->psdm[fstage][i] = psd_mean[i];
    }
    if (fstage == 0) {
        if (nbf < 100) lambda = 1 - 1 / (float)nbf;
        else lambda = (float)0.99;
        if (flag == 0)
            for (i = 0; i < n; i++) {
                noise[i] = lambda * noise[i] + (1 - lambda) * psd_mean[i];
                if (noise[i] < s->eps) noise[i] = s->eps;
            }
    }
    for (i = 0; i < n; i++) {
        float post = (psd_mean[i] / noise[i]) - 1;
        float prio = beta * (den[i] / noise[i]) + (1 - beta) * ((0 > post) ? (0) : (post));
        W[i] = prio / (1 + prio);
        prio = W[i] * psd_mean[i] / noise[i];
        prio = (prio > rsb_min) ? prio : rsb_min;
        W[i] = prio / (1 + prio);
        den[i] = W[i] * nsig[i];
    }
}

/* DoGainFact: etsi/cpp/NoiseSup.c:581-642 (DoGainFact_IBM, aurora_etsi/NoiseSup.cpp:634-698, same math) */
static void ns_gain_fact(so_ns *s, int fstage, float *W)
{
    int i;
    if (fstage == 0) {
        s->den_en[0] = s->den_en[1];
        s->den_en[1] = s->den_en[2];
        s->den_en[2] = 0.0;
        for (i = 0; i < s->SPEC; i++) s->den_en[2] += s->den[0][i];
    } else {
        float noise_en = 0.0, aver, lambda;
        for (i = 0; i < s->SPEC; i++) noise_en += s->noise[1][i];
        aver = (s->den_en[0] * s->den_en[1] * s->den_en[2]) / (noise_en * noise_en * noise_en);
        /* NoiseSup.cpp is C++: log10(float) is the float overload there */
        if (aver > 0.00001) aver = (s->profile == SO_PROFILE_P16) ? (20 * log10f(aver)) / 3.0 : (20 * log10(aver)) / 3.0;
        else aver = -100.0 / 3.0;
        if (((aver - s->low_snr) < 10.0) || (s->nb_frame[1] < 10)) {
            if (s->nb_frame[1] < 10) lambda = 1.0 - 1.0 / (float)s->nb_frame[1];
            else if (aver < s->low_snr) lambda = 0.95;
            else lambda = 0.99;
            s->low_snr += (1.0 - lambda) * (aver - s->low_snr);
        }
        if (s->den_en[2] > 100) {
            if (aver < (s->low_snr + 3.5)) {
                s->alfa_gf += 0.15;
                if (s->alfa_gf > 0.8) s->alfa_gf = 0.8;
            } else {
                s->alfa_gf -= 0.3;
                if (s->alfa_gf < 0.1) s->alfa_gf = 0.1;
            }
        }
        for (i = 0; i < SO_MEL; i++) W[i] = s->alfa_gf * W[i] + (1.0 - s->alfa_gf) * 1.0;
    }
}

/* SpeechQVar / SpeechQSpec / SpeechQMel: etsi/cpp/NoiseSup.c:686-839 */
static int ns_q_var(so_ns *s, const float *W)
{
    const short ssize = s->NFFT / 4, fc = (short)s->nb_frame[0];
    float var = 0.0, mean = 0.0, sv;
    int i;
    for (i = 0; i < ssize; i++) { mean += W[i]; var += W[i] * W[i]; }
    sv = (var / ssize) - mean * mean / (ssize * ssize);
    if (fc < 15) s->fd_var_mean = (s->fd_var_mean > sv) ? s->fd_var_mean : sv;
    if (sv < s->fd_var_mean * 1.5 && sv > s->fd_var_mean * 0.85) s->fd_var_mean = (s->fd_var_mean * 0.8 + sv * 0.2);
    if (sv <= s->fd_var_mean * 0.25) s->fd_var_mean = (s->fd_var_mean * 0.97 + sv * 0.03);
    return (sv > s->fd_var_mean * 1.65) ? 1 : 0;
}

static int ns_q_spec(so_ns *s)
{
    const short fc = (short)s->nb_frame[0];
    if (fc == 1) s->fd_spec_mean = s->fd_spec_val;
    if (fc < 15) {
        float acc;
        s->fd_acc_test = 1.1 * (s->fd_acc_test * (float)(fc - 1) + s->fd_spec_val) / (float)(fc);
        acc = s->fd_spec_val / s->fd_acc_test;
        if (acc > 2.5) s->fd_in_vadq = 1;
        if (s->fd_in_vadq == 0) s->fd_spec_mean = (s->fd_spec_mean > s->fd_spec_val) ? s->fd_spec_mean : s->fd_spec_val;
    }
    if (s->fd_spec_val < s->fd_spec_mean * 1.5 && s->fd_spec_val > s->fd_spec_mean * 0.75)
        s->fd_spec_mean = (s->fd_spec_mean * 0.8 + s->fd_spec_val * 0.2);
    if (s->fd_spec_val <= s->fd_spec_mean * 0.5) s->fd_spec_mean = (s->fd_spec_mean * 0.97 + s->fd_spec_val * 0.03);
    return (s->fd_spec_val > s->fd_spec_mean * 1.65) ? 1 : 0;
}

static int ns_q_mel(so_ns *s)
{
    const short fc = (short)s->nb_frame[0];
    float sm = 0.75 * s->fd_mel_val[1] + 0.25 * s->fd_mel_val[0];
    s->fd_mel_val[0] = s->fd_mel_val[1];
    if (fc < 15) s->fd_mel_mean = (s->fd_mel_mean > sm) ? s->fd_mel_mean : sm;
    if (sm < s->fd_mel_mean * 1.5 && sm > s->fd_mel_mean * 0.75) s->fd_mel_mean = (s->fd_mel_mean * 0.8 + sm * 0.2);
    if (sm <= s->fd_mel_mean * 0.5) s->fd_mel_mean = (s->fd_mel_mean * 0.97 + sm * 0.03);
    return (sm > s->fd_mel_mean * 3.25) ? 1 : 0;
}

/* One block: DoNoiseSup (etsi/cpp/NoiseSup.c:1061-1440) == body of the func_Wiener loop
 * (aurora_etsi/NoiseSup.cpp:1172-1403). */
/* =====================================================================================
 * 16 kHz extension of the P8 stream: etsi/cpp/16kHzProcessing.c and its hooks in
 * DoNoiseSup (etsi/cpp/NoiseSup.c:1104-1121, 1166-1172, 1216-1235, 1307-1327, 1378-1382, 1402-1435)
 * ===================================================================================== */

/* Half-band prototype (DP_HP_filters, 16kHzProcessing.c:559-608): 118 symmetric taps, integers / 2^23; only the
 * first 59 are listed, tap[117-i] = tap[i].  The high-pass is the time-reversed prototype with alternating signs. */
static const int qmf_half[59] = {
    1584, 805, -4192, -8985, -5987, 2583, 4657, -3035, -7004, 1542, 8969, 567, -10924, -3757, 12320, 7951, -12793,
    -13048, 11923, 18793, -9331, -24802, 4694, 30570, 2233, -35439, -11526, 38680, 23114, -39474, -36701, 36999,
    51797, -30419, -67658, 18962, 83318, -1927, -97566, -21284, 108971, 51215, -115837, -88430, 116130, 133716,
    -107253, -188497, 85497, 255795, -44643, -342699, -28185, 468096, 167799, -696809, -519818, 1446093, 3562497};

void so_qmf_taps(float *dp118, float *hp118)
{
    int i, sign = 1;
    for (i = 0; i < HP_QMF; i++) dp118[i] = (float)((double)qmf_half[i < 59 ? i : HP_QMF - 1 - i] / (float)0x00800000);
    for (i = 0; i < HP_QMF; i++) {
        hp118[HP_QMF - 1 - i] = (float)sign * dp118[i];
        sign = -sign;
    }
}

/* Do16kProcAlloc/Init (16kHzProcessing.c:178-271): everything zero, 5 un-normalised mel triangles from 80 Hz over
 * the 65-bin PSD (InitMelFBwindows(.., 80.0, 8000, 128, 5, 0)). */
static void hp16k_init(so_hp16k *h)
{
    int c[5], i, j, k;
    const float st_freq = 80.0, smpl = 8000.0;
    const int fft_len = 128, nch = 5;
    float start_mel = 2595.0 * log10(1.0 + (float)st_freq / 700.0);
    float half_mel = 2595.0 * log10(1.0 + (smpl / 2) / 700.0);
    memset(h, 0, sizeof *h);
    so_qmf_taps(h->dp, h->hp);
    for (i = 0; i < nch; i++) {
        float f = 700 * (pow(10, (start_mel + (float)i / (nch - 1) * (half_mel - start_mel)) / 2595.0) - 1.0);
        c[i] = (int)(fft_len * f / smpl + 0.5);
    }
    for (i = 0; i < nch; i++) {
        float *d = h->fb[i];
        int len;
        if (i == 0) {
            h->fb_start[i] = c[0];
            len = c[1] - c[0];
            for (j = 0; j < len; j++) d[j] = 1.0 - (float)j / (float)len;
        } else if (i < nch - 1) {
            const int up = c[i] - c[i - 1];
            h->fb_start[i] = c[i - 1] + 1;
            len = c[i + 1] - c[i - 1] - 1;
            for (j = 0; j < up; j++) d[j] = (float)(j + 1) / (float)up;
            for (j = up, k = 0; j < len; j++, k++) d[j] = 1.0 - (k + 1) / (float)(c[i + 1] - c[i]);
        } else {
            h->fb_start[i] = c[nch - 2] + 1;
            len = c[nch - 1] - c[nch - 2];
            for (j = 0; j < len; j++) d[j] = (float)(j + 1) / (float)len;
        }
        h->fb_len[i] = len;
    }
}

int so_ns_attach_hp16k(so_ns *s)
{
    if (!s || s->profile != SO_PROFILE_P8 || s->hp16k) return -1;
    s->hp16k = (so_hp16k *)malloc(sizeof(so_hp16k));
    if (!s->hp16k) return -1;
    hp16k_init(s->hp16k);
    return 0;
}

void so_hp16k_tables(const so_ns *s, int *start5, int *len5, float *fb5x65)
{
    memcpy(start5, s->hp16k->fb_start, sizeof s->hp16k->fb_start);
    memcpy(len5, s->hp16k->fb_len, sizeof s->hp16k->fb_len);
    memcpy(fb5x65, s->hp16k->fb, sizeof s->hp16k->fb);
}

/* Do16kProcessing (16kHzProcessing.c:711-787) on one 160-sample call: x[160] (float 16 kHz samples) ->
 * lp[80] (0-4 kHz half, what NoiseSup sees) and h->data_hp[80] (4-8 kHz half, spectrally flipped to 0-4 kHz). */
static void hp16k_split(so_hp16k *h, const float *x, float *lp)
{
    float line[HP_QMF - 1 + 160];
    int i, j, k, sign = 1;
    memcpy(line, h->hist, sizeof h->hist);
    memcpy(line + HP_QMF - 1, x, sizeof(float) * 160);
    for (i = 0, k = 0; i < 160; i += 2, k++) {
        float a = 0.0, b = 0.0;
        for (j = 0; j < HP_QMF; j++) {
            a += line[i + j] * h->dp[j];
            b += line[i + j] * h->hp[j];
        }
        lp[k] = a;
        h->data_hp[k] = (float)sign * b;
        sign = -sign;
    }
    memcpy(h->hist, x + 160 - (HP_QMF - 1), sizeof h->hist);
}

/* Stage-0 hook (NoiseSup.c:1166-1172, 1216-1235): remember the HP frame that lines up with the cepstral frame,
 * and log-energies of three low-band PSD regions (GetBandsForCoding16k, 16kHzProcessing.c:497-511) pushed into a
 * 3-frame history. */
static void hp16k_stage0(so_hp16k *h, const float *nsig)
{
    float aux[3] = {0.0, 0.0, 0.0};
    const float floor_v = (float)exp((double)-10.0);
    int i;
    for (i = 0; i < 200; i++) h->snap[i] = h->buf[80 + i];             /* NS_ANALYSIS_WINDOW_16K */
    for (i = 33; i < 39; i++) aux[0] += nsig[i];
    for (i = 39; i < 49; i++) aux[1] += nsig[i];
    for (i = 49; i < 65; i++) aux[2] += nsig[i];
    for (i = 0; i < 6; i++) h->lp_log[i] = h->lp_log[i + 3];
    for (i = 0; i < 3; i++) {
        if (aux[i] > floor_v) h->lp_log[6 + i] = (float)log((double)aux[i]);
        else h->lp_log[6 + i] = (float)-10.0;
    }
}

/* DoSpecSub16k (16kHzProcessing.c:388-480): energy VAD on the 3 HP bands, noise tracking, over-subtraction 1.5
 * with a floor at 10 %. */
static void hp16k_spec_sub(so_hp16k *h, float *fb, long frame_no)
{
    const float eps = (float)exp((double)-10.0);
    int i, flag, nbf = frame_no;
    float lambda, diff, floor_v, frame_en, mean_en = h->mean_en;
    if (nbf < 100) lambda = 1.0 - 1.0 / (float)nbf;
    else lambda = 0.99;
    frame_en = 0.0;
    for (i = 0; i < 3; i++) frame_en += fb[i];
    if (frame_en > 0.001) frame_en = (float)log((double)frame_en);
    else frame_en = (float)log((double)0.001);
    if (((frame_en - mean_en) < 1.2) || (nbf < 10)) {
        if (nbf < 10) mean_en += (1 - lambda) * (frame_en - mean_en);
        else if (frame_en < mean_en) mean_en += (1 - 0.98) * (frame_en - mean_en);
        else mean_en += (1 - 0.995) * (frame_en - mean_en);
    }
    if ((frame_en - mean_en) > 2.2) {
        flag = 1;
        h->nb_speech++;
    } else {
        if (h->nb_speech > 4) h->hang_over = 15;
        h->nb_speech = 0;
        if (h->hang_over != 0) { h->hang_over--; flag = 1; }
        else flag = 0;
    }
    if (flag == 0) {
        for (i = 0; i < 3; i++) {
            if (nbf < 10) h->noise[i] = lambda * h->noise[i] + (1 - lambda) * fb[i];
            else if (fb[i] < h->noise[i]) h->noise[i] = lambda * h->noise[i] + (1 - lambda) * fb[i];
            else h->noise[i] = 0.995 * h->noise[i] + (1 - 0.995) * fb[i];
            if (h->noise[i] < eps) h->noise[i] = eps;
        }
    }
    for (i = 0; i < 3; i++) {
        floor_v = 0.1 * fb[i];
        diff = fb[i] - 1.5 * h->noise[i];
        fb[i] = (diff > floor_v) ? diff : floor_v;
    }
    h->mean_en = mean_en;
}

/* Stage-1 hook (NoiseSup.c:1307-1327): PSD of the remembered HP frame -> mel bands 1..3 of 5 -> log ->
 * code = (oldest low-band log energies) - (HP log band) (CodeBands16k, :530-537) -> spectral subtraction. */
static void hp16k_stage1(so_hp16k *h, const float *sig_window, long frame_no)
{
    float t[256], psd[65], sum[5], aux[3];
    const float floor_v = (float)exp((double)-10.0);
    int i, j;
    for (i = 0; i < 200; i++) t[i] = h->snap[i] * sig_window[i];
    for (i = 200; i < 256; i++) t[i] = 0.0;
    so_rfft(t, 256, 8);
    t[0] = t[0] * t[0];
    for (i = 1, j = 255; i < 128; i++, j--) t[i] = (t[i] * t[i] + t[j] * t[j]);
    t[i] = t[i] * t[i];
    for (i = 0, j = 0; i < 64; i++, j += 2) psd[i] = (t[j] + t[j + 1]) / 2.0;
    psd[i] = t[j];
    for (j = 0; j < 5; j++) {
        sum[j] = 0.0;
        for (i = 0; i < h->fb_len[j]; i++) sum[j] += psd[h->fb_start[j] + i] * h->fb[j][i];
    }
    for (i = 0; i < 3; i++) h->fb16k[i] = sum[i + 1];
    for (i = 0; i < 3; i++) {
        if (h->fb16k[i] > floor_v) aux[i] = log(h->fb16k[i]);
        else aux[i] = (float)-10.0;
    }
    for (i = 0; i < 3; i++)
        for (j = 0; j < 3; j++) h->code[3 * i + j] = h->lp_log[j] - aux[i];
    hp16k_spec_sub(h, h->fb16k, frame_no);
}

/* After a denoised shift is out (NoiseSup.c:1402-1435): 3-frame delay lines of the HP bands and their codes. */
static void hp16k_push(so_hp16k *h)
{
    int i;
    for (i = 3; i < 9; i++) h->hp_bands[i - 3] = h->hp_bands[i];
    for (i = 0; i < 3; i++) h->hp_bands[6 + i] = h->fb16k[i];
    for (i = 0; i < 18; i++) h->code_buf[i] = h->code_buf[i + 9];
    for (i = 0; i < 9; i++) h->code_buf[18 + i] = h->code[i];
}

int so_ns_block(so_ns *s, const float *in, float *out, so_ns_info *info)
{
    float tmp[MAXFFT + SO_MAX_SPEC]; /* the reference's tmpMem aliases, kept apart here */
    float W[2 * SO_MAX_SPEC], psd_mean[SO_MAX_SPEC], fir[NS_FILT], sig_out[MAXFRAME];
    const int S = s->S, n = s->SPEC;
    int fstage, i, j;
    if (info) { memset(info, 0, sizeof *info); }
    for (i = 0; i < S; i++) s->buf1[s->DIN + i] = in[i];
    s->nb_in1++;
    if (s->hp16k) memcpy(s->hp16k->buf + 480 - 80, s->hp16k->data_hp, sizeof(float) * 80);   /* NoiseSup.c:1119-1120 */
    for (fstage = 0; fstage < 2; fstage++) {
        float *buf, *nsig, *hist, *prv, *cur;
        short *idx;
        if (fstage == 0 && (s->nb_in1 - s->nb_in2) > 2) buf = s->buf1;
        else if (fstage == 1 && (s->nb_in2 - s->nb_out2) > 2) buf = s->buf2;
        else continue;
        nsig = s->nsig[fstage];
        hist = s->psd_hist[fstage];
        idx = &s->idx[fstage];
        prv = buf;
        cur = buf + S;
        /* DoSigWindowing :218-231, rfft, FFTtoPSD :249-270 */
        for (i = 0; i < s->FL; i++) tmp[i] = buf[s->ANA + i] * s->sig_window[i];
        for (i = s->FL; i < s->NFFT; i++) tmp[i] = 0.0;
        so_rfft_impl(tmp, s->NFFT, s->M, s->profile == SO_PROFILE_P16);
        tmp[0] = tmp[0] * tmp[0];
        for (i = 1, j = s->NFFT - 1; i < s->NFFT / 2; i++, j--) tmp[i] = (tmp[i] * tmp[i] + tmp[j] * tmp[j]);
        tmp[i] = tmp[i] * tmp[i];
        for (i = 0, j = 0; i < n - 1; i++, j += 2) nsig[i] = (tmp[j] + tmp[j + 1]) / 2.0;
        nsig[i] = tmp[j];
        if (s->hp16k && fstage == 0) hp16k_stage0(s->hp16k, nsig);
        /* PSDMean :289-303 */
        *idx = 1 - *idx;
        for (i = 0; i < n; i++) {
            int k = i * 2 + *idx;
            hist[k] = nsig[i];
            k += 1 - 2 * *idx;
            psd_mean[i] = (hist[k] + nsig[i]) / 2;
        }
        ns_vad(s, fstage, cur);
        ns_filter_calc(s, fstage, psd_mean, W);
        if (fstage == 0 && info) info->sp_var = ns_q_var(s, W);
        else if (fstage == 0) ns_q_var(s, W);
        /* DoMelFB (etsi/cpp/MelProc.c:82-104) / DoGamma (aurora_etsi/MelProc.cpp:119-141, Q3) */
        {
            float sum[SO_MEL];
            for (j = 0; j < SO_MEL; j++) {
                const int st = (s->profile == SO_PROFILE_P8) ? s->bank_start[j] : 0;
                sum[j] = 0.0;
                for (i = 0; i < s->bank_len[j]; i++) sum[j] += W[st + i] * s->bank[j][i];
            }
            for (j = SO_MEL - 1; j >= 0; j--) W[j] = sum[j];
        }
        if (fstage == 0) {
            float te = 0.0;
            int a, b;
            for (i = 0; i < SO_MEL; i++) te += (float)W[i];
            s->fd_spec_val = te * te - 3.0;
            a = ns_q_spec(s);
            s->fd_mel_val[1] = (float)(W[1] + W[2] + W[3]) / 3.0;
            b = ns_q_mel(s);
            if (info) { info->sp_spec = a; info->sp_mel = b; }
        }
        ns_gain_fact(s, fstage, W);
        if (fstage == 1 && info) memcpy(info->gains, W, sizeof(float) * SO_MEL);
        /* DoMelIDCT (MelProc.c:357-378), mirror (NoiseSup.c:1292), DoFilterWindowing (:660-669) */
        {
            float o[SO_MEL];
            int t, f;
            for (t = 0; t < SO_MEL; t++) {
                o[t] = 0.0;
                for (f = 0; f < SO_MEL; f++) o[t] += W[f] * s->idct[t][f];
            }
            for (t = 0; t < SO_MEL; t++) W[t] = o[t];
            for (i = 1; i < SO_MEL; i++) W[2 * SO_MEL - 1 - i] = W[i];
            for (i = NS_HALF, j = 0; i < NS_FILT; i++, j++) {
                fir[NS_HALF + j] = W[j] * s->ir_window[i];
                fir[NS_HALF - j] = fir[NS_HALF + j];
            }
        }
        /* ApplyWF :324-340 */
        for (i = 0; i < S; i++) {
            sig_out[i] = 0.0;
            for (j = -NS_HALF; j <= (i <= NS_HALF ? i : NS_HALF); j++) sig_out[i] += (fir[j + NS_HALF] * cur[i - j]);
            for (j = i + 1; j <= NS_HALF; j++) sig_out[i] += (fir[j + NS_HALF] * prv[S - j + i]);
        }
        if (s->hp16k && fstage == 1) hp16k_stage1(s->hp16k, s->sig_window, s->nb_frame[1]);
        if (fstage == 0) {
            for (i = 0; i < S; i++) s->buf2[s->DIN + i] = sig_out[i];
            s->nb_in2++;
            if (info) { info->stage0_ran = 1; info->sp_vadns = (s->nb_speech > 4) ? 1 : 0; }
        } else {
            for (i = 0; i < S; i++) out[i] = sig_out[i];
            s->nb_out2++;
            if (info) info->stage1_ran = 1;
        }
    }
    if (s->nb_in1) for (i = 0; i < s->DIN; i++) s->buf1[i] = s->buf1[i + S];
    if (s->nb_in1 && s->hp16k) memmove(s->hp16k->buf, s->hp16k->buf + 80, sizeof(float) * 400);   /* :1378-1382 */
    if (s->nb_in2) for (i = 0; i < s->DIN; i++) s->buf2[i] = s->buf2[i + S];
    if (info) {
        info->flag_vad = s->flag_vad; info->hang_over = s->hang_over; info->nb_speech = s->nb_speech;
        info->nb_frame0 = s->nb_frame[0]; info->nb_frame1 = s->nb_frame[1];
        info->mean_en = s->mean_en; info->alfa_gf = s->alfa_gf; info->low_snr = s->low_snr;
        info->den_en = s->den_en[2];
    }
    if (s->nb_out2 > 0) {
        /* DCOffsetFil :182-198 */
        for (i = 0; i < S; i++) {
            float aux = out[i];
            out[i] = out[i] - s->dc_x + 0.9990234375 * s->dc_y;
            s->dc_x = aux;
            s->dc_y = out[i];
        }
        if (info) info->produced = 1;
        if (s->hp16k) hp16k_push(s->hp16k);
        return 1;
    }
    return 0;
}

/* =====================================================================================
 * CompCeps (P8): etsi/cpp/CompCeps.c
 * ===================================================================================== */
typedef struct {
    float hamming[100], dct[12 * 23], dct26[12 * 26];
    int start[23], len[23];
    float tri[23][129];
    int ready;
} so_cc_tab;
static so_cc_tab g_cc;

static void cc_init(void)
{
    int i, j, upper = 0;
    const float st_freq = 64.0, smpl = 8000.0;
    const int fft_len = 256, nch = 23;
    float start_mel, half_mel;
    if (g_cc.ready) return;
    for (i = 0; i < 100; i++) g_cc.hamming[i] = 0.54 - 0.46 * cos(SO_PIx2 * (i + 0.5) / 200); /* CompCeps.c:87-94 */
    for (i = 1; i < 13; i++)                                                                   /* :153-173 */
        for (j = 0; j < nch; j++)
            g_cc.dct[(i - 1) * nch + j] = cos(SO_PI * (float)i / (float)nch * ((float)j + 0.5));
    for (i = 1; i < 13; i++)                                                                   /* 16 kHz: 23 + 3 bands */
        for (j = 0; j < 26; j++)
            g_cc.dct26[(i - 1) * 26 + j] = cos(SO_PI * (float)i / (float)26 * ((float)j + 0.5));
    /* InitFFTWindows, etsi/cpp/MelProc.c:402-463 */
    start_mel = 2595.0 * log10(1.0 + (float)st_freq / 700.0);
    half_mel = 2595.0 * log10(1.0 + (smpl / 2) / 700.0);
    for (i = 0; i < nch; i++) {
        float f = 700 * (pow(10, (start_mel + (float)i / (nch + 1) * (half_mel - start_mel)) / 2595.0) - 1.0);
        int lo = (int)(fft_len * f / smpl + 0.5);
        g_cc.start[i] = lo;
        f = 700 * (pow(10, (start_mel + (float)(i + 2) / (nch + 1) * (half_mel - start_mel)) / 2595.0) - 1.0);
        g_cc.len[i] = (int)(fft_len * f / smpl + 0.5) - lo + 1;
    }
    /* ComputeTriangle, etsi/cpp/MelProc.c:482-522 */
    for (i = 0; i < nch; i++) {
        int low = (i < nch - 1) ? g_cc.start[i + 1] - g_cc.start[i] + 1 : upper - g_cc.start[i] + 1;
        int hgh = g_cc.len[i] - low + 1;
        memset(g_cc.tri[i], 0, sizeof g_cc.tri[i]);
        for (j = 0; j < low; j++) g_cc.tri[i][j] = (float)(j + 1) / low;
        for (j = 1; j < hgh; j++) g_cc.tri[i][low + j - 1] = (float)(hgh - j) / hgh;
        upper = g_cc.start[i] + g_cc.len[i] - 1;
    }
    g_cc.ready = 1;
}

void so_cc_tables(float *hamming100, float *dct12x23, int *start23, int *len23, float *tri23x129)
{
    cc_init();
    memcpy(hamming100, g_cc.hamming, sizeof g_cc.hamming);
    memcpy(dct12x23, g_cc.dct, sizeof g_cc.dct);
    memcpy(start23, g_cc.start, sizeof g_cc.start);
    memcpy(len23, g_cc.len, sizeof g_cc.len);
    memcpy(tri23x129, g_cc.tri, sizeof g_cc.tri);
}

void so_cc_dct26(float *dct12x26)
{
    cc_init();
    memcpy(dct12x26, g_cc.dct26, sizeof g_cc.dct26);
}

/* WI8CompCeps, etsi/cpp/CompCeps.c:368-549.  frame201[0] is sample -1.  hp == NULL: the 8 kHz branch (23 bands);
 * otherwise the 16 kHz branch: 3 decoded + spectrally-subtracted HP bands are appended (26 bands, DCT over 26) and
 * their energy is added to logE. */
static void cc_core(const float *frame201, float *coef14, const so_hp16k *hp)
{
    const float *cur = frame201 + 1;
    float fb[256 + 16], sum[23], log_e, coded[3];
    const float floor_fb = (float)exp((double)-10.0), floor_loge = (float)exp((double)-50.0);
    const int nb = hp ? 26 : 23;
    int i, j;
    cc_init();
    log_e = 0.0;
    for (i = 0; i < 200; i++) log_e += cur[i] * cur[i];
    if (!hp) {
        if (log_e < floor_loge) log_e = -50.0;
        else log_e = (float)log((double)log_e);
    }
    for (i = 0; i < 200; i++) fb[i] = cur[i] - 0.90 * cur[i - 1];
    for (i = 0; i < 100; i++) fb[i] *= g_cc.hamming[i];
    for (i = 100; i < 200; i++) fb[i] *= g_cc.hamming[200 - 1 - i];
    for (i = 200; i < 256; i++) fb[i] = 0.0;
    so_rfft(fb, 256, 8);
    fb[0] = (double)fb[0] * fb[0];
    for (i = 1; i < 128; i++)
        fb[i] = (float)((double)fb[i] * (double)fb[i] + (double)fb[256 - i] * (double)fb[256 - i]);
    fb[128] = (double)fb[128] * fb[128];
    if (hp) {
        /* GetBandsForDecoding16k + DecodeBands16k (16kHzProcessing.c:318-366) with the fixed weights of
         * CompCeps.c:475-477 and the OLDEST of the three buffered codes */
        float aux[3] = {0.0, 0.0, 0.0}, w[3];
        for (i = 66; i < 77; i++) aux[0] += fb[i];
        for (i = 78; i < 97; i++) aux[1] += fb[i];
        for (i = 98; i < 129; i++) aux[2] += fb[i];
        for (i = 0; i < 3; i++) {
            aux[i] /= 2.0;
            if (aux[i] > floor_fb) aux[i] = (float)log((double)aux[i]);
            else aux[i] = -10.0;
        }
        w[0] = 0.1; w[1] = 0.2; w[2] = 0.7;
        for (i = 0; i < 3; i++) {
            coded[i] = 0.0;
            for (j = 0; j < 3; j++) coded[i] += (w[j] * (aux[j] - hp->code_buf[3 * i + j]));
        }
    }
    for (j = 0; j < 23; j++) {      /* DoMelFB, MelProc.c:82-104 */
        sum[j] = 0.0;
        for (i = 0; i < g_cc.len[j]; i++) sum[j] += fb[g_cc.start[j] + i] * g_cc.tri[j][i];
    }
    for (j = 22; j >= 0; j--) fb[j] = sum[j];
    if (hp) {                       /* CompCeps.c:488-505: rough pre-emphasis correction of both HP estimates */
        const float corr = log(1.0 + 0.90);
        for (i = 0; i < 3; i++) fb[23 + i] = (1.0 + 0.90) * hp->hp_bands[i];
        for (i = 0; i < 3; i++) coded[i] += corr;
    }
    for (i = 0; i < nb; i++) {
        if (fb[i] < floor_fb) fb[i] = -10.0;
        else fb[i] = (float)log((double)fb[i]);
    }
    if (hp) {
        /* MergeSSandCoded (16kHzProcessing.c:106-125): 70 % coded + 30 % subtracted, then bands 22|23 pulled together */
        const float perc = 0.7;
        float avg, e_hp = 0.0;
        const float preem = 0.90;                 /* CorrectEnergy takes the coefficient as a float argument */
        const float corr = log(1.0 + preem);
        for (i = 0; i < 3; i++) fb[23 + i] = perc * coded[i] + (1.0 - perc) * fb[23 + i];
        avg = 0.5 * fb[22] + 0.5 * fb[23];
        fb[22] = 0.6 * fb[22] + 0.4 * avg;
        fb[23] = 0.6 * fb[23] + 0.4 * avg;
        /* CorrectEnergy (:145-158) */
        for (i = 0; i < 3; i++) e_hp += exp(fb[23 + i] - corr);
        log_e += e_hp;
        if (log_e < floor_loge) log_e = -50.0;
        else log_e = (float)log((double)log_e);
    }
    for (i = 1; i < 13; i++) {      /* DCT, CompCeps.c:203-227 */
        const float *mx = hp ? g_cc.dct26 : g_cc.dct;
        fb[nb + (i - 1)] = 0.0;
        for (j = 0; j < nb; j++) fb[nb + (i - 1)] += fb[j] * mx[(i - 1) * nb + j];
    }
    fb[nb + 12] = 0.0;
    for (i = 0; i < nb; i++) fb[nb + 12] += fb[i];
    fb[nb + 13] = log_e;            /* Noc0 == FALSE: logE appended after c0 (:539) */
    for (i = 0; i < 14; i++) coef14[i] = fb[nb + i];
}

void so_comp_ceps(const float *frame201, float *coef14) { cc_core(frame201, coef14, NULL); }

/* etsi_denoise (etsi/cpp/AdvFrontEnd.c:125-210) driving DoAdvProcess (etsi/cpp/ParmInterface.c:208-330),
 * with the commented CompCeps step (:274-293) when ceps != NULL. */
long so_etsi_denoise(const int16_t *in, long n, int16_t *den_i16, float *den_f32, float *ceps,
                     long *n_ceps, so_ns_info *per_call)
{
    so_ns *s = so_ns_create(SO_PROFILE_P8);
    float den_buf[241], cur[80];
    int16_t out_buf[80];
    long start = 0, calls = 0, nc = 0;
    int onset = 0, offset = -200, i;
    memset(den_buf, 0, sizeof den_buf);   /* BufInAlloc calloc, BufferIn.c:37-48 */
    memset(out_buf, 0, sizeof out_buf);   /* DenoiseBuffer calloc, AdvFrontEnd.c:161 */
    while (start + 80 <= n) {
        float check = 0.0, *pt = den_buf + 241 - 80;
        so_ns_info info;
        int produced = 0;
        memmove(den_buf, den_buf + 80, sizeof(float) * (241 - 80));   /* BufInShiftToPut */
        for (i = 0; i < 80; i++) {
            cur[i] = (float)in[start + i];
            check += in[start + i] * in[start + i];
        }
        memset(&info, 0, sizeof info);
        if (((int)check != 0) || onset) {
            onset = 1;
            produced = so_ns_block(s, cur, pt, &info);
            if (produced) {
                for (i = 0; i < 80; i++) out_buf[i] = (int16_t)pt[i];
                if (ceps) {
                    if (offset < 0) offset += 80;
                    if (offset >= 0) {
                        so_comp_ceps(den_buf + 241 - (200 + offset + 1), ceps + nc * 14);
                        nc++;
                    }
                }
            }
        } else {
            info.produced = -1;
        }
        if (den_i16) memcpy(den_i16 + start, out_buf, sizeof out_buf);
        if (den_f32) {
            if (produced) memcpy(den_f32 + start, pt, sizeof(float) * 80);
            else memset(den_f32 + start, 0, sizeof(float) * 80);
        }
        if (per_call) per_call[calls] = info;
        calls++;
        start += 80;
    }
    so_ns_destroy(s);
    if (n_ceps) *n_ceps = nc;
    return calls;
}

/* -------------------------------------------------------------------------------------
 * Back half of the ETSI feature pipeline: waveform processing,
 * blind equalisation and the frame-dropping VAD, as DoAdvProcess would chain them
 * (etsi/cpp/ParmInterface.c:274-311, shipped commented out).
 * ------------------------------------------------------------------------------------- */

/* argmax over t[lo..hi] (inclusive, ascending scan) where later equal values win (>=), as the
 * neighbour searches of GetMaximaPositions do (WaveProc.c:107-114, 126-133; they scan offsets 25..79
 * away from the current maximum, so "later" means farther away).  -1 when every candidate is negative
 * (only possible after int32 wrap-around of the 9-point sum on full-scale input). */
static int wp_far_peak(const int32_t *t, int from, int dir, int len)
{
    int best = -1, d;
    int32_t top = 0;
    for (d = 25; d < 80; d++) {
        const int p = from + dir * d;
        if (p < 0 || p >= len) continue;
        if (t[p] >= top) { top = t[p]; best = p; }
    }
    return best;
}

/* DoWaveProc, etsi/cpp/WaveProc.c:386-453 (TeagerEng :173-184, GetTeagerFilter :204-291,
 * GetMaximaPositions :74-153).  d = 200 samples, modified in place. */
void so_wave_proc(float *d)
{
    enum { L = 200 };
    float energy = 0.0f, w[L];
    int32_t q[L], sm[L];
    int pos[24], n_pos = 0, i, k;
    for (i = 0; i < L; i++) energy += d[i] * d[i];
    if (!(energy >= 100.0)) return;                                   /* :403-404 */
    /* Teager operator with replicated edges, then round(0.25 * |.|) to int (:222-227) */
    for (i = 0; i < L; i++) {
        const float a = d[i > 0 ? i - 1 : 0], b = d[i < L - 1 ? i + 1 : L - 1];
        const float t = (float)fabs(d[i] * d[i] - a * b);
        q[i] = (int)floor(t * 0.25 + 0.5);
    }
    /* 9-point running sum with replicated edges (:222-254); unsigned = the reference's wrap-around */
    for (i = 0; i < L; i++) {
        uint32_t acc = 0;
        for (k = -4; k <= 4; k++) {
            int j = i + k;
            j = j < 0 ? 0 : (j > L - 1 ? L - 1 : j);
            acc += (uint32_t)q[j];
        }
        sm[i] = (int32_t)acc;
    }
    /* global maximum (first strict), then neighbours 25..79 samples away on either side (:86-135) */
    {
        int32_t top = 0;
        int g = -1, m, left[12], nl = 0;
        for (i = 0; i < L; i++)
            if (sm[i] > top) { top = sm[i]; g = i; }
        if (g >= 0) {
            m = g;
            while (m - 25 > 0) { const int p = wp_far_peak(sm, m, -1, L); if (p < 0) break; left[nl++] = m = p; }
            for (i = nl - 1; i >= 0; i--) pos[n_pos++] = left[i];
            pos[n_pos++] = g;
            m = g;
            while (m + 25 < L) { const int p = wp_far_peak(sm, m, +1, L); if (p < 0) break; pos[n_pos++] = m = p; }
        }
    }
    /* window: 0.4 everywhere, 0.6 over the first 80 % of every inter-peak gap, starting 4 samples
     * before the peak (:262-289).  pos[] ascending (sort_it, :44-66). */
    {
        const float eps = (float)0.2;
        const float lo = (float)((1 - eps) / 2.0), hi = (float)((1 + eps) / 2.0);
        for (i = 0; i < L; i++) w[i] = lo;
        if (n_pos > 1) {
            for (k = 0; k < n_pos; k++) {
                const int gap = (k < n_pos - 1) ? pos[k + 1] - pos[k] : pos[k] - pos[k - 1];
                const int a = pos[k] - 4, b = a + (80 * gap + 99) / 100;
                for (i = a; i < b; i++)
                    if (i >= 0 && i < L) w[i] = hi;
            }
        }
    }
    for (i = 0; i < L; i++) d[i] *= w[i] + w[i < L - 1 ? i + 1 : i];       /* :430-447 */
}

/* DoPostProc, etsi/cpp/PostProc.c:117-150 (Noc0 = 0: the energy term is coef[13] = logE) */
static const float pp_target[12] = {(float)-6.618909, (float)0.198269, (float)-0.740308, (float)0.055132,
    (float)-0.227086, (float)0.144280, (float)-0.112451, (float)-0.146940, (float)-0.327466, (float)0.134571,
    (float)0.027884, (float)-0.114905};

void so_post_proc(float *coef14, float *lms12)
{
    const float lambda = (float)0.0087890625;
    float wp = (coef14[13] * 64.0f - 211.0f) / 64.0f;
    int i;
    if (wp < 0) wp = 0;
    else if (wp > 1) wp = lambda;
    else wp *= lambda;
    for (i = 0; i < 12; i++) {
        const float dif = (coef14[i] - lms12[i]) - pp_target[i];
        coef14[i] = coef14[i] - lms12[i];
        lms12[i] += dif * wp;
    }
}

/* DoVADProc, etsi/cpp/VAD.c:214-330: 7-frame circular delay line + hangover logic.
 * fb15 in: 14 coefficients; out (when 1 is returned): the frame 6 calls back + its VAD flag. */
void so_vad_init(so_vad *v)
{
    memset(v, 0, sizeof *v);
    v->flush_focus = -1;                                                /* :142 */
    v->hang_over = 23;                                                  /* HANGOVERV, :113 */
}

int so_vad_proc(so_vad *v, float *fb15, int speech_found, int frame_counter)
{
    int i, run = 0, longest = 0;
    v->focus = (v->focus + 1) % 7;
    memcpy(v->buf[v->focus], fb15, sizeof(float) * 14);
    v->buf[v->focus][14] = speech_found ? 1.0f : 0.0f;
    if (frame_counter <= 10) return 0;                                  /* BUFFER_SIZE + 3, :256 */
    for (i = 1; i <= 7; i++) {                                          /* oldest -> newest */
        if (v->buf[(v->focus + i) % 7][14] != 0.0f) { if (++run > longest) longest = run; }
        else run = 0;
    }
    if (longest >= 4) {                                                 /* H_TRIGGER */
        v->h_count = v->hang_over;
        if (frame_counter <= 35) v->hang_over = 50;                     /* SANGOVER */
    }
    if (v->h_count && longest < 3) v->h_count--;
    if (longest >= 3) v->v_count = 5;                                   /* V_TRIGGER / VANGOVER */
    if (v->v_count && longest < 3) v->v_count--;
    memcpy(fb15, v->buf[(v->focus + 1) % 7], sizeof(float) * 14);
    fb15[14] = (v->v_count || v->h_count || longest >= 3) ? 1.0f : 0.0f;
    return 1;
}

/* DoVADFlush, etsi/cpp/VAD.c:342-431: one flushing step.  The delay line is not written, only the focus moves, so
 * the scan order of the 7 speech flags rotates.  Returns 1 while rows are due (6 times); fb15 is rewritten only when
 * the incremented frame counter is past 10 -- otherwise the caller's FeatureBuffer keeps what it held and the
 * reference's driver loop (while (FlushAdvProcess(...)) write) emits that stale row. */
int so_vad_flush(so_vad *v, float *fb15, int *frame_counter)
{
    int i, run = 0, longest = 0;
    if (v->flush_focus < 0) v->flush_focus = v->focus;
    v->focus = (v->focus + 1) % 7;
    if (v->focus == v->flush_focus) return 0;
    (*frame_counter)++;
    if (*frame_counter > 10) {
        for (i = 1; i <= 7; i++) {
            if (v->buf[(v->focus + i) % 7][14] != 0.0f) { if (++run > longest) longest = run; }
            else run = 0;
        }
        if (longest >= 4) {
            v->h_count = v->hang_over;
            if (*frame_counter <= 35) v->hang_over = 50;
        }
        if (v->h_count && longest < 3) v->h_count--;
        if (longest >= 3) v->v_count = 5;
        if (v->v_count && longest < 3) v->v_count--;
        memcpy(fb15, v->buf[(v->focus + 1) % 7], sizeof(float) * 14);
        fb15[14] = (v->v_count || v->h_count || longest >= 3) ? 1.0f : 0.0f;
    }
    return 1;
}

/* The full per-frame chain, driven like oracle/ref_etsi.c:ref_etsi_features.  The VAD's FrameCounter
 * is the stage-1 frame count DoNoiseSup publishes (NoiseSup.c:1366).  feat: [n/80][15]; returns rows. */
static long features_impl(const int16_t *in, long n, float *feat, int16_t *den_i16, int flush)
{
    int frame_counter = 0;
    so_ns *s = so_ns_create(SO_PROFILE_P8);
    so_vad vad;
    float den_buf[241], cur[80], fb[16], lms[12], frame[201];
    int16_t out_buf[80];
    long start = 0, rows = 0;
    int onset = 0, offset = -200, i;
    memset(den_buf, 0, sizeof den_buf);
    memset(out_buf, 0, sizeof out_buf);
    memset(fb, 0, sizeof fb);
    memset(lms, 0, sizeof lms);
    so_vad_init(&vad);
    while (start + 80 <= n) {
        float check = 0.0, *pt = den_buf + 241 - 80;
        so_ns_info info;
        int out = 0;
        memmove(den_buf, den_buf + 80, sizeof(float) * (241 - 80));
        for (i = 0; i < 80; i++) {
            cur[i] = (float)in[start + i];
            check += in[start + i] * in[start + i];
        }
        memset(&info, 0, sizeof info);
        if (((int)check != 0) || onset) {
            onset = 1;
            const int produced = so_ns_block(s, cur, pt, &info);
            if (info.stage0_ran || info.stage1_ran) frame_counter = info.nb_frame0;     /* NoiseSup.c:1366 */
            if (produced) {
                for (i = 0; i < 80; i++) out_buf[i] = (int16_t)pt[i];
                if (offset < 0) offset += 80;
                if (offset >= 0) {
                    memcpy(frame, den_buf + 241 - (200 + offset + 1), sizeof frame);
                    so_wave_proc(frame + 1);
                    so_comp_ceps(frame, fb);
                    so_post_proc(fb, lms);
                    out = so_vad_proc(&vad, fb, info.sp_spec || info.sp_mel || info.sp_var || info.sp_vadns,
                                      info.nb_frame0);
                }
            }
        } else {
            for (i = 0; i < 14; i++) fb[i] = 0.0f;                      /* ParmInterface.c:320-321 */
            out = 1;
        }
        if (out) { memcpy(feat + rows * 15, fb, sizeof(float) * 15); rows++; }
        if (den_i16) memcpy(den_i16 + start, out_buf, sizeof out_buf);
        start += 80;
    }
    /* FlushAdvProcess (ParmInterface.c:347-355) the way the ETSI driver loop uses it: while it returns TRUE, write */
    while (flush && so_vad_flush(&vad, fb, &frame_counter)) { memcpy(feat + rows * 15, fb, sizeof(float) * 15); rows++; }
    so_ns_destroy(s);
    return rows;
}

/* The same chain at SamplingFrequency 16000 (AdvProcessAlloc, ParmInterface.c:99-109): 160 samples per call, QMF
 * split, NoiseSup on the low band, HP-band coding inside NoiseSup, decoding/merging inside CompCeps.  Driven like
 * oracle/ref_etsi.c:ref_etsi_features_16k.  feat: [n/160][15]; den_i16: [n/160][80] (low band); returns rows. */
static long features16_impl(const int16_t *in, long n, float *feat, int16_t *den_i16, int flush)
{
    int frame_counter = 0;
    so_ns *s = so_ns_create(SO_PROFILE_P8);
    so_vad vad;
    float den_buf[241], cur[160], lp[80], fb[16], lms[12], frame[201];
    int16_t out_buf[80];
    long start = 0, rows = 0, call = 0;
    int onset = 0, offset = -200, i;
    so_ns_attach_hp16k(s);
    memset(den_buf, 0, sizeof den_buf);
    memset(out_buf, 0, sizeof out_buf);
    memset(fb, 0, sizeof fb);
    memset(lms, 0, sizeof lms);
    so_vad_init(&vad);
    while (start + 160 <= n) {
        float check = 0.0, *pt = den_buf + 241 - 80;
        so_ns_info info;
        int out = 0;
        memmove(den_buf, den_buf + 80, sizeof(float) * (241 - 80));
        for (i = 0; i < 160; i++) {
            cur[i] = (float)in[start + i];
            check += in[start + i] * in[start + i];
        }
        memset(&info, 0, sizeof info);
        if (((int)check != 0) || onset) {
            onset = 1;
            hp16k_split(s->hp16k, cur, lp);
            const int produced = so_ns_block(s, lp, pt, &info);
            if (info.stage0_ran || info.stage1_ran) frame_counter = info.nb_frame0;
            if (produced) {
                for (i = 0; i < 80; i++) out_buf[i] = (int16_t)pt[i];
                if (offset < 0) offset += 80;
                if (offset >= 0) {
                    memcpy(frame, den_buf + 241 - (200 + offset + 1), sizeof frame);
                    so_wave_proc(frame + 1);
                    cc_core(frame, fb, s->hp16k);
                    so_post_proc(fb, lms);
                    out = so_vad_proc(&vad, fb, info.sp_spec || info.sp_mel || info.sp_var || info.sp_vadns,
                                      info.nb_frame0);
                }
            }
        } else {
            for (i = 0; i < 14; i++) fb[i] = 0.0f;
            out = 1;
        }
        if (out) { memcpy(feat + rows * 15, fb, sizeof(float) * 15); rows++; }
        if (den_i16) memcpy(den_i16 + call * 80, out_buf, sizeof out_buf);
        start += 160;
        call++;
    }
    /* FlushAdvProcess (ParmInterface.c:347-355) the way the ETSI driver loop uses it: while it returns TRUE, write */
    while (flush && so_vad_flush(&vad, fb, &frame_counter)) { memcpy(feat + rows * 15, fb, sizeof(float) * 15); rows++; }
    so_ns_destroy(s);
    return rows;
}

long so_etsi_features(const int16_t *in, long n, float *feat, int16_t *den_i16) { return features_impl(in, n, feat, den_i16, 0); }
long so_etsi_features_16k(const int16_t *in, long n, float *feat, int16_t *den_i16) { return features16_impl(in, n, feat, den_i16, 0); }
/* flags: bit 0 = 16 kHz mode, bit 1 = flush the VAD delay line at the end; feat needs calls + 6 rows */
long so_etsi_features_ex(const int16_t *in, long n, float *feat, int16_t *den_i16, int flags)
{
    return (flags & 1) ? features16_impl(in, n, feat, den_i16, (flags >> 1) & 1) : features_impl(in, n, feat, den_i16, (flags >> 1) & 1);
}

/* Do16kProcessing as a stream: in[n] -> lp[n/2], hp[n/2] (n rounded down to a multiple of 160). */
void so_qmf_split(const int16_t *in, long n, float *lp, float *hp)
{
    so_hp16k *h = (so_hp16k *)malloc(sizeof *h);
    long start;
    int i;
    hp16k_init(h);
    for (start = 0; start + 160 <= n; start += 160) {
        float x[160];
        for (i = 0; i < 160; i++) x[i] = (float)in[start + i];
        hp16k_split(h, x, lp + start / 2);
        memcpy(hp + start / 2, h->data_hp, sizeof h->data_hp);
    }
    free(h);
}

/* etsi_denoise_mapping_func_Wiener, aurora_etsi/NoiseSup.cpp:1140-1407 */
int so_p16_func_wiener(so_ns *s, const float *in, int n, float *out, int *sp_var, int *sp_spec,
                       int *sp_mel, int *sp_vadns, int *frame_counter, float *gains)
{
    const int S = s->S;
    long nblk = n / S, b;
    int rows = 0, i;
    for (b = 0; b < nblk; b++, in += S, out += S) {
        float check = 0.0;
        so_ns_info info;
        for (i = 0; i < S; i++) check += in[i] * in[i];
        if (0 == (int)check) continue;
        so_ns_block(s, in, out, &info);
        if (info.stage0_ran) {
            sp_var[b] = info.sp_var; sp_spec[b] = info.sp_spec; sp_mel[b] = info.sp_mel; sp_vadns[b] = info.sp_vadns;
        }
        if (info.stage0_ran || info.stage1_ran) frame_counter[b] = info.nb_frame0;
        if (info.stage1_ran) {
            if (gains) memcpy(gains + rows * SO_MEL, info.gains, sizeof(float) * SO_MEL);
            rows++;
        }
    }
    return rows;
}

/* =====================================================================================
 * G64: gammatone / hair cell / sub-bands / resynthesis
 * (resyth_64sub_IBM/cpp/extractwav.cpp and HuWang.h; identical helpers in the other variants)
 * ===================================================================================== */
#define G_NCH 64
#define G_FS 16000
#define G_WINDOW 320
#define G_OFFSET 160
#define G_PI (3.1415926535897932384626433832795)

/* initMiddleEar (extractwav.cpp:134-167): BS3383 table 1 */
static const float g_ear_f[29] = {20.0, 25.0, 31.5, 40.0, 50.0, 63.0, 80.0, 100.0, 125.0, 160.0, 200.0, 250.0, 315.0,
    400.0, 500.0, 630.0, 800.0, 1000.0, 1250.0, 1600.0, 2000.0, 2500.0, 3150.0, 4000.0, 5000.0, 6300.0, 8000.0,
    10000.0, 12500.0};
static const float g_ear_af[29] = {2.347, 2.190, 2.050, 1.879, 1.724, 1.579, 1.512, 1.466, 1.426, 1.394, 1.372, 1.344,
    1.304, 1.256, 1.203, 1.135, 1.062, 1.000, 0.967, 0.943, 0.932, 0.933, 0.937, 0.952, 0.974, 1.027, 1.135, 1.266,
    1.501};
static const float g_ear_bf[29] = {0.00561, 0.00527, 0.00481, 0.00404, 0.00383, 0.00286, 0.00259, 0.00257, 0.00256,
    0.00255, 0.00254, 0.00248, 0.00229, 0.00201, 0.00162, 0.00111, 0.00052, 0.00000, -0.00039, -0.00067, -0.00092,
    -0.00105, -0.00104, -0.00088, -0.00055, 0.00000, 0.00089, 0.00211, 0.00488};
static const float g_ear_tf[29] = {74.3, 65.0, 56.3, 48.4, 41.7, 35.5, 29.8, 25.1, 20.7, 16.8, 13.8, 11.2, 8.9, 7.2, 6.0,
    5.0, 4.4, 4.2, 3.7, 2.6, 1.0, -1.2, -3.6, -3.9, -1.1, 6.6, 15.3, 16.4, 11.6};

/* loudnessLevelInPhons, extractwav.cpp:259-279 */
static float loudness_phons(float freq)
{
    int i = 0;
    float ratio, afy, bfy, tfy;
    while (g_ear_f[i] < freq) i++;
    ratio = (freq - g_ear_f[i - 1]) / (g_ear_f[i] - g_ear_f[i - 1]);
    afy = g_ear_af[i - 1] + ratio * (g_ear_af[i] - g_ear_af[i - 1]);
    bfy = g_ear_bf[i - 1] + ratio * (g_ear_bf[i] - g_ear_bf[i - 1]);
    tfy = g_ear_tf[i - 1] + ratio * (g_ear_tf[i] - g_ear_tf[i - 1]);
    return (4.2 + afy * (60.0 - tfy) / (1.0 + bfy * (60.0 - tfy)));
}

/* channel set-up inline in resynth()/subbband(): extractwav.cpp:41-54 */
void so_channel_params(float *cf64, float *bw64, float *mid64)
{
    float lower = 21.4 * log10(50 * 0.00437 + 1.0);
    float upper = 21.4 * log10(8000 * 0.00437 + 1.0);
    float space = (upper - lower) / (G_NCH - 1);
    int c;
    for (c = 0; c < G_NCH; c++) {
        float cf = (pow(10, (lower + c * space) / 21.4) - 1) / 0.00437;
        float phon;
        cf64[c] = cf;
        bw64[c] = 24.7 * (cf * 0.00437 + 1.0) * 1.019;
        phon = (loudness_phons(cf) - 60.0);
        mid64[c] = pow(10, phon / 20);
    }
}

/* gammaToneFilter, extractwav.cpp:168-212: starts from rest, output delayed one sample (Q7) */
void so_gammatone(const float *in, float *out, float cf, float bw, float mid, long n)
{
    float dt = 1 / (float)(G_FS);
    float two_pi_t = 2 * G_PI * dt;
    float gain = mid * pow(two_pi_t * bw, 4) / 3.0;
    /* extractwav.cpp is C++: exp/cos/sin of a float argument are the float overloads */
    float z = expf(-two_pi_t * bw);
    float f1 = cosf(cf * two_pi_t) * z;
    float f2 = sinf(cf * two_pi_t) * z;
    float p[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0}, x[4], y[4];
    long t;
    int i;
    for (t = 0; t < n; t++) {
        out[t] = p[3] * gain;
        for (i = 0; i < 4; i++) {
            x[i] = f1 * p[i] - f2 * q[i];
            y[i] = f2 * p[i] + f1 * q[i];
        }
        p[0] = in[t] * f1 + x[0];
        q[0] = in[t] * f2 + y[0];
        p[1] = p[0] + x[1];
        q[1] = q[0] + y[1];
        p[2] = p[1] + x[1] + x[2];
        q[2] = q[1] + y[1] + y[2];
        p[3] = p[2] + x[1] + 2 * x[2] + x[3];
        q[3] = q[2] + y[1] + 2 * y[2] + y[3];
    }
}

/* hairCell, extractwav.cpp:213-258 (Meddis 1988) */
void so_haircell(const float *in, float *out, long n)
{
    const double Y = 5.05, G = 2000.0, L = 2500.0, R = 6580.0, X = 66.31, A = 3.0, B = 300.0, H = 48000.0, Mm = 1.0;
    float dt = 1 / (float)(G_FS);
    float ymdt = Y * Mm * dt, xdt = X * dt, ydt = Y * dt, lplusrdt = (L + R) * dt, rdt = R * dt, gdt = G * dt, hdt = H;
    float kt = G * A / (A + B);
    float c = Mm * Y * kt / (L * kt + Y * (L + R));
    float q = c * (L + R) / kt;
    float w = c * R / X;
    long t;
    for (t = 0; t < n; t++) {
        float replenish, eject, loss, reuptake, reprocess;
        kt = ((in[t] + A) > 0.0) ? (gdt * (in[t] + A) / (in[t] + A + B)) : 0;
        replenish = (q < Mm) ? (ymdt - ydt * q) : 0;
        eject = kt * q;
        loss = lplusrdt * c;
        reuptake = rdt * c;
        reprocess = xdt * w;
        q = q + replenish - eject + reprocess;
        if (q < 0.0) q = 0.0;
        c = c + eject - loss;
        if (c < 0.0) c = 0.0;
        w = w + reuptake - reprocess;
        if (w < 0.0) w = 0.0;
        out[t] = hdt * c;
    }
}

/* subbband, enhancement_extract_subband_linux/cpp/extractwav.cpp:36-97 */
void so_subband(const int16_t *wav, long n, int16_t *out64xn)
{
    float cf[G_NCH], bw[G_NCH], mid[G_NCH];
    float *in = (float *)malloc(sizeof(float) * n), *g = (float *)malloc(sizeof(float) * n),
          *h = (float *)malloc(sizeof(float) * n);
    long i;
    int c;
    so_channel_params(cf, bw, mid);
    for (i = 0; i < n; i++) in[i] = (float)wav[i];
    for (c = 0; c < G_NCH; c++) {
        so_gammatone(in, g, cf[c], bw[c], mid[c], n);
        so_haircell(g, h, n);
        for (i = 0; i < n; i++) out64xn[(long)c * n + i] = (int16_t)h[i];
    }
    free(in); free(g); free(h);
}

/* addnoise, enhancement_extract_subband_linux/cpp/extractwav.cpp:6-35 (Q4: cast before multiply) */
void so_addnoise(const int16_t *pure, int16_t *noise, int16_t *noisy, long n, int db)
{
    float dB = pow(10.0, db / 10.0);
    float ps = 0, ns = 0, gain;
    long i;
    for (i = 0; i < n; i++) {
        ps += pure[i] * pure[i] * 1.0;
        ns += noise[i] * noise[i] * 1.0;
    }
    gain = sqrt(ps / ns / dB);
    for (i = 0; i < n; i++) {
        noise[i] = (short)noise[i] * gain;
        noisy[i] = pure[i] + noise[i];
    }
}

int so_resynth_num_frames(long n, int mode)
{
    return (mode == SO_RESYNTH_1DNN) ? (int)(n / G_OFFSET) : (int)((n - G_WINDOW) / G_OFFSET + 1);
}

/* resynth: resyth_64sub_IBM/cpp/extractwav.cpp:9-132 (+ the _ori / 1dnn diffs) */
void so_resynth(const int16_t *wav, long n, float *mask, int num_frame, int mode, int16_t *out, float *out_f32)
{
    float cf[G_NCH], bw[G_NCH], mid[G_NCH];
    float *in = (float *)malloc(sizeof(float) * n), *g = (float *)malloc(sizeof(float) * n),
          *rev = (float *)malloc(sizeof(float) * n), *wgt = (float *)malloc(sizeof(float) * n),
          *acc = (float *)calloc(n, sizeof(float));
    long i;
    int c, frame, k;
    so_channel_params(cf, bw, mid);
    for (i = 0; i < n; i++) in[i] = (float)wav[i];
    for (c = 0; c < G_NCH; c++) {
        so_gammatone(in, g, cf[c], bw[c], mid[c], n);
        for (i = 0; i < n; i++) rev[n - i - 1] = g[i] / mid[c];
        so_gammatone(rev, g, cf[c], bw[c], mid[c], n);
        for (i = 0; i < n; i++) rev[n - i - 1] = g[i] / mid[c];
        for (i = 0; i < n; i++) wgt[i] = 0.0;
        for (frame = 0; frame < num_frame; frame++) {
            float *m = &mask[frame * G_NCH + c];
            if ((mode == SO_RESYNTH_IBM) ? (*m > 0.5) : (*m > 0)) {
                if (mode == SO_RESYNTH_IBM) *m = 1.0;
                if (frame > 0)
                    for (k = 0; k < G_OFFSET; k++)
                        wgt[(frame - 1) * G_OFFSET + k] += 0.5 * (1.0 + cos(k * G_PI / (G_OFFSET) + G_PI)) * *m;
                for (k = G_OFFSET; k < G_WINDOW; k++)
                    wgt[(frame - 1) * G_OFFSET + k] += 0.5 * (1.0 + cos((k - G_OFFSET) * G_PI / (G_OFFSET))) * *m;
            }
        }
        for (i = 0; i < n; i++) acc[i] += wgt[i] * rev[i];
    }
    for (i = 0; i < n; i++) out[i] = (short)acc[i];
    if (out_f32) memcpy(out_f32, acc, sizeof(float) * n);
    free(in); free(g); free(rev); free(wgt); free(acc);
}

/* -------------------------------------------------------------------------------------
 * make_single_IBM, enhancement_extract_subband_linux/cpp/show_IBM.cpp:92-213.
 * asdk::SpecInfo(WINDOW=320, OFFSET=160), m_nFFT=512, GetSpecInfo() is UN-VENDORED.  Our
 * restatement of it (documented as ours): frame i = samples [160 i, 160 i + 320),
 * Hamming window 0.54-0.46cos(2 pi k/319), zero-pad to 512, unscaled |FFT|^2.  Pinned by the
 * reference: frame count (show_IBM.cpp:110), bins summed 0..m_nFFT/2-1 (:156), IRM formula (:168);
 * the IBM bit is the strict compare of function/.../show_IBM/show_IBM.cpp:138-148.
 * ------------------------------------------------------------------------------------- */
long so_ibm_num_frames(long n) { return (n - G_WINDOW) / G_OFFSET + 1; }

static float frame_half_power(const int16_t *x)
{
    float buf[512], sum = 0;
    int k;
    for (k = 0; k < G_WINDOW; k++) buf[k] = (float)x[k] * (float)(0.54 - 0.46 * cos(SO_PIx2 * k / (G_WINDOW - 1)));
    for (k = G_WINDOW; k < 512; k++) buf[k] = 0.0f;
    so_rfft(buf, 512, 9);
    sum += buf[0] * buf[0];
    for (k = 1; k < 256; k++) sum += buf[k] * buf[k] + buf[512 - k] * buf[512 - k];
    return sum;
}

void so_single_ibm(const int16_t *pure64, const int16_t *noise64, long n, float *irm, uint8_t *ibm)
{
    long frames = so_ibm_num_frames(n), f;
    int c;
    for (c = 0; c < G_NCH; c++)
        for (f = 0; f < frames; f++) {
            float sp = frame_half_power(pure64 + (long)c * n + f * G_OFFSET);
            float sn = frame_half_power(noise64 + (long)c * n + f * G_OFFSET);
            if (irm) irm[f * G_NCH + c] = sp / (sp + sn);
            if (ibm) ibm[f * G_NCH + c] = (sn < sp) ? 1 : 0;
        }
}
