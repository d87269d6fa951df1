/* Deterministic synthetic noisy-speech corpus.  Host-only helper used by the tests
 * and bench.py to make inputs; it is not on the hot path and has no reference counterpart (the
 * reference ships no audio).  The mixing recipe mirrors what the reference's drivers do to real
 * audio: additive noise at a chosen SNR (function/.../main_part/main.cpp:110-180) and RMS
 * normalisation to 3000 (function/.../aurora_speech_enhancement.cpp:172-184).
 *
 *   utterance u, seed 0x5EED0000+u, xorshift64* PRNG
 *   "speech": 5 harmonics (amplitude 1/k) of an f0 gliding 100->250 Hz, 4 Hz syllabic AM, gated on
 *             in 300-800 ms bursts separated by 200-500 ms gaps; the first 200 ms are noise only
 *   noise   : white Gaussian (u%6 < 3) or one-pole pink, a = 0.95 (u%6 >= 3); SNR = {-5,0,5}[u%3] dB
 *   5 % of utterances (u%20 == 7) start with 50 ms of exact zeros (leading zero-frame path)
 *   mixture scaled to RMS 3000, rounded to int16, clipped
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static inline uint64_t xs64(uint64_t *s)
{
    uint64_t x = *s;
    x ^= x >> 12;
    x ^= x << 25;
    x ^= x >> 27;
    *s = x;
    return x * 0x2545F4914F6CDD1DULL;
}
static inline double urand(uint64_t *s) { return ((xs64(s) >> 11) + 0.5) * (1.0 / 9007199254740992.0); }

static inline short clip16(double v)
{
    double r = nearbyint(v);
    if (r > 32767.0) r = 32767.0;
    if (r < -32768.0) r = -32768.0;
    return (short)r;
}

/* Length rule for the ragged configs: uniform in [min_s, max_s] seconds, forced to be a
 * NON-multiple of 160 samples so the reference's tail-drop paths are exercised. */
long se_synth_length(long u, int fs, double min_s, double max_s)
{
    uint64_t s = 0x9E3779B97F4A7C15ULL ^ (uint64_t)(0x5EED0000L + u);
    xs64(&s);
    long n = (long)((min_s + (max_s - min_s) * urand(&s)) * fs);
    if (n % 160 == 0) n += 37;
    return n;
}

/* Writes n samples of utterance u.  Any of noisy/clean/noise may be NULL.
 * clean[] and noise[] are the two components at their mixed (scaled) levels. */
int se_synth_utterance(long u, int fs, long n, short *noisy, short *clean, short *noise)
{
    uint64_t s = (uint64_t)(0x5EED0000L + u);
    double *sp = (double *)malloc(sizeof(double) * (size_t)n);
    double *nz = (double *)malloc(sizeof(double) * (size_t)n);
    if (!sp || !nz) { free(sp); free(nz); return -1; }
    for (int i = 0; i < 8; i++) xs64(&s);
    const double two_pi = 6.283185307179586476925287;
    const double dur = (double)n / fs;
    const int pink = (u % 6) >= 3;
    static const double snr_tab[3] = {-5.0, 0.0, 5.0};
    const double snr_db = snr_tab[u % 3];
    const long lead_zero = (u % 20 == 7) ? (long)(0.050 * fs) : 0;

    /* speech: gated harmonic complex */
    long i = 0, on_start = (long)(0.200 * fs), on_end = on_start;
    double phase = 0.0;
    for (i = 0; i < n; i++) sp[i] = 0.0;
    long pos = on_start;
    while (pos < n) {
        long burst = (long)((0.300 + 0.500 * urand(&s)) * fs);
        long gap = (long)((0.200 + 0.300 * urand(&s)) * fs);
        on_end = pos + burst;
        if (on_end > n) on_end = n;
        for (i = pos; i < on_end; i++) {
            double t = (double)i / fs;
            double f0 = 100.0 + 150.0 * (dur > 0 ? t / dur : 0.0);
            phase += two_pi * f0 / fs;
            if (phase > two_pi) phase -= two_pi;
            double am = 0.6 + 0.4 * sin(two_pi * 4.0 * t);
            /* 5 ms raised-cosine edges so bursts do not click */
            double e = 1.0;
            long edge = (long)(0.005 * fs);
            if (i - pos < edge) e = 0.5 - 0.5 * cos(3.14159265358979323846 * (double)(i - pos) / edge);
            if (on_end - 1 - i < edge) e *= 0.5 - 0.5 * cos(3.14159265358979323846 * (double)(on_end - 1 - i) / edge);
            double v = 0.0;
            for (int k = 1; k <= 5; k++) {
                if (k * f0 < 0.45 * fs) v += sin(k * phase) / k;
            }
            sp[i] = v * am * e;
        }
        pos = on_end + gap;
    }

    /* noise: Box-Muller white, optionally one-pole low-passed */
    double prev = 0.0;
    for (i = 0; i + 1 < n; i += 2) {
        double u1 = urand(&s), u2 = urand(&s);
        double r = sqrt(-2.0 * log(u1));
        nz[i] = r * cos(two_pi * u2);
        nz[i + 1] = r * sin(two_pi * u2);
    }
    if (i < n) {
        double u1 = urand(&s), u2 = urand(&s);
        nz[i] = sqrt(-2.0 * log(u1)) * cos(two_pi * u2);
    }
    if (pink)
        for (i = 0; i < n; i++) { prev = 0.95 * prev + nz[i]; nz[i] = prev; }

    for (i = 0; i < lead_zero && i < n; i++) { sp[i] = 0.0; nz[i] = 0.0; }

    double es = 0.0, en = 0.0;
    for (i = 0; i < n; i++) { es += sp[i] * sp[i]; en += nz[i] * nz[i]; }
    if (es <= 0.0) es = 1.0;
    if (en <= 0.0) en = 1.0;
    double g = sqrt(es / en / pow(10.0, snr_db / 10.0)); /* noise gain for the target SNR */
    double em = 0.0;
    for (i = 0; i < n; i++) { double m = sp[i] + g * nz[i]; em += m * m; }
    long nn = n - lead_zero;
    double scale = 3000.0 / sqrt(em / (nn > 0 ? nn : 1));
    for (i = 0; i < n; i++) {
        double c = scale * sp[i], z = scale * g * nz[i];
        if (clean) clean[i] = clip16(c);
        if (noise) noise[i] = clip16(z);
        if (noisy) noisy[i] = clip16(c + z);
    }
    free(sp);
    free(nz);
    return 0;
}

/* Synthetic "DNN" soft masks for config 4: m = sigmoid(4*(irm-0.5) + 0.3*g), g~N(0,1). */
int se_synth_soft_mask(long u, const float *irm, long count, float *mask)
{
    uint64_t s = 0xD1B54A32D192ED03ULL ^ (uint64_t)(0x5EED0000L + u);
    for (int i = 0; i < 8; i++) xs64(&s);
    for (long i = 0; i < count; i++) {
        double u1 = urand(&s), u2 = urand(&s);
        double g = sqrt(-2.0 * log(u1)) * cos(6.283185307179586476925287 * u2);
        double x = 4.0 * ((double)irm[i] - 0.5) + 0.3 * g;
        mask[i] = (float)(1.0 / (1.0 + exp(-x)));
    }
    return 0;
}
