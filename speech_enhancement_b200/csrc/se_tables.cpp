// Host-side constant tables (see se_tables.h).  All sub-expressions are written with explicit
// float/double types so they evaluate exactly like the reference's C (ETSI files) or C++
// (plugin / extractwav files) — where the reference's C++ picks a float overload (expf, cosf, sinf)
// the same call is made here.
#include "se_tables.h"
#include <cmath>
#include <cstring>

namespace se {

static const double kPIx2 = 6.28318530717958647692;   // etsi/cpp/ParmInterface.h:37
static const double kPI = 3.14159265358979323846;     // etsi/cpp/CompCeps.c:31

static void hanning(std::vector<float> &w, int n)
{
    // etsi/cpp/NoiseSup.c:974-979: 0.5 - 0.5*cos(PIx2*((float)i+0.5)/(float)N)
    w.resize(n);
    for (int i = 0; i < n; i++)
        w[i] = (float)(0.5 - 0.5 * std::cos((kPIx2 * ((double)(float)i + 0.5)) / (double)(float)n));
}

static void idct_from_centres(NsTables &t, const float *cf, int samp_freq)
{
    // etsi/cpp/MelProc.c:320-336 (same in aurora_etsi/MelProc.cpp:487-503)
    float df[kMel];
    for (int j = 0; j < kMel; j++) {
        if (j == 0) df[j] = (cf[1] - cf[0]) / (float)samp_freq;
        else if (j == kMel - 1) df[j] = (cf[j] - cf[j - 1]) / (float)samp_freq;
        else df[j] = (cf[j + 1] - cf[j - 1]) / (float)samp_freq;
    }
    t.idct.resize(kMel * kMel);
    for (int i = 0; i < kMel; i++)
        for (int j = 0; j < kMel; j++)
            t.idct[i * kMel + j] = (float)((double)df[j] * std::cos(kPIx2 * (double)i * (double)cf[j] / (double)samp_freq));
}

void build_ns_tables_p8(NsTables &t)
{
    // constants: etsi/cpp/NoiseSup.h:28-79
    t.shift = 80; t.frame = 200; t.nfft = 256; t.spec = 65; t.ana = 60; t.fs = 8000;
    t.bank_w = 65;
    t.eps = (float)std::exp(-10.0);
    hanning(t.sig_window, t.frame);
    hanning(t.ir_window, kFir);
    // InitMelFBwindows(FirstWindow, 0.0, 8000, 128, 25, normalize=1), etsi/cpp/MelProc.c:128-230
    const float st_freq = 0.0f, smpl = 8000.0f;
    const int fft_len = 2 * (t.spec - 1), nch = kMel;
    int c[kMel];
    const float start_mel = (float)(2595.0 * std::log10(1.0 + (double)st_freq / 700.0));
    const float half_mel = (float)(2595.0 * std::log10(1.0 + (double)(smpl / 2) / 700.0));
    for (int i = 0; i < nch; i++) {
        const float arg = start_mel + (float)i / (float)(nch - 1) * (half_mel - start_mel);
        const float f = (float)(700.0 * (std::pow(10.0, (double)arg / 2595.0) - 1.0));
        c[i] = (int)((double)((float)fft_len * f / smpl) + 0.5);
    }
    t.bank.assign(kMel * t.bank_w, 0.0f);
    for (int i = 0; i < nch; i++) {
        float *d = &t.bank[i * t.bank_w];
        float norm = 0.0f;
        int len;
        if (i == 0) {
            t.bank_start[i] = c[0];
            len = c[1] - c[0];
            for (int j = 0; j < len; j++) { d[j] = (float)(1.0 - (double)((float)j / (float)len)); norm += d[j]; }
        } else if (i < nch - 1) {
            const int up = c[i] - c[i - 1];
            t.bank_start[i] = c[i - 1] + 1;
            len = c[i + 1] - c[i - 1] - 1;
            for (int j = 0; j < up; j++) { d[j] = (float)(j + 1) / (float)up; norm += d[j]; }
            for (int j = up, k = 0; j < len; j++, k++) {
                d[j] = (float)(1.0 - (double)((float)(k + 1) / (float)(c[i + 1] - c[i])));
                norm += d[j];
            }
        } else {
            t.bank_start[i] = c[nch - 2] + 1;
            len = c[nch - 1] - c[nch - 2];
            for (int j = 0; j < len; j++) { d[j] = (float)(j + 1) / (float)len; norm += d[j]; }
        }
        t.bank_len[i] = len;
        for (int j = 0; j < len; j++) d[j] /= norm;
    }
    // InitMelIDCTbasis, etsi/cpp/MelProc.c:283-337
    float cf[kMel];
    const float lin = (float)t.fs / (float)fft_len;
    for (int j = 0; j < kMel; j++) {
        const float *d = &t.bank[j * t.bank_w];
        if (j == 0) cf[j] = (float)t.bank_start[0] * lin;
        else if (j == kMel - 1) cf[j] = (float)(t.bank_start[j] + t.bank_len[j] - 1) * lin;
        else {
            const float st = (float)t.bank_start[j] * lin;
            float sum = 0.0f;
            cf[j] = 0.0f;
            for (int i = 0; i < t.bank_len[j]; i++) {
                cf[j] += d[i] * (st + (float)i * lin);
                sum += d[i];
            }
            cf[j] /= sum;
        }
    }
    idct_from_centres(t, cf, t.fs);
}

void build_ns_tables_p16(NsTables &t)
{
    // constants: function/20141106_speech_enhancement/aurora_etsi/NoiseSup.h:22-72
    t.shift = 160; t.frame = 480; t.nfft = 512; t.spec = 129; t.ana = 80; t.fs = 16000;
    t.bank_w = 128;
    t.eps = (float)std::exp(-10.0);
    hanning(t.sig_window, t.frame);
    hanning(t.ir_window, kFir);
    // InitGammawindows(FirstWindow, 80.0, 16000, 256, 25, 1), aurora_etsi/MelProc.cpp:268-341
    const float st_freq = 80.0f, smpl = 16000.0f;
    const int fft_len = 2 * (t.spec - 1), nch = kMel;
    auto hz_to_erb = [](float hz) { return (float)(21.4 * std::log10((double)hz * 0.00437 + 1.0)); };
    auto erb_to_hz = [](float e) { return (float)((std::pow(10.0, (double)e / 21.4) - 1.0) / 0.00437); };
    const float lo = hz_to_erb(st_freq), hi = hz_to_erb(smpl / 2);
    const float step = (hi - lo) / (float)(nch - 1);
    float cfs[kMel], erb[kMel];
    for (int i = 0; i < nch; i++) {
        cfs[i] = erb_to_hz(lo + step * (float)i);
        erb[i] = (float)(1.019 * 24.7 * (4.37 * (double)cfs[i] / 1000.0 + 1.0));
    }
    t.bank.assign(kMel * t.bank_w, 0.0f);
    for (int i = 0; i < nch; i++) {
        float *d = &t.bank[i * t.bank_w];
        float norm = 0.0f;
        t.bank_start[i] = (int)((float)((int)cfs[i] * fft_len) / smpl);   // Q3: (int)cf*FFT/Fs
        t.bank_len[i] = fft_len / 2;
        for (int j = 0; j < t.bank_len[i]; j++) {
            const double dd = (double)j * 1.0 / (double)fft_len * (double)smpl - (double)cfs[i];
            d[j] = (float)(1.0 / std::pow(1.0 + dd / (double)erb[i] * dd / (double)erb[i], 2.0));
            norm += d[j];
        }
        for (int j = 0; j < t.bank_len[i]; j++) d[j] /= norm;
    }
    // InitGammaIDCTbasis, aurora_etsi/MelProc.cpp:463-504: centre = StartingPoint * linStep
    float cf[kMel];
    const float lin = (float)t.fs / (float)fft_len;
    for (int j = 0; j < kMel; j++) cf[j] = (float)t.bank_start[j] * lin;
    idct_from_centres(t, cf, t.fs);
}

void build_cc_tables(CcTables &t)
{
    std::memset(&t, 0, sizeof t);
    // InitializeHamming, etsi/cpp/CompCeps.c:87-94
    for (int i = 0; i < 100; i++) t.hamming[i] = (float)(0.54 - 0.46 * std::cos(kPIx2 * ((double)i + 0.5) / 200.0));
    // InitDCTMatrix, etsi/cpp/CompCeps.c:153-173
    for (int i = 1; i < 13; i++)
        for (int j = 0; j < 23; j++)
            t.dct[(i - 1) * 23 + j] = (float)std::cos(kPI * (double)(float)i / (double)(float)23 * ((double)(float)j + 0.5));
    // InitFFTWindows(64.0, 8000, 256, 23), etsi/cpp/MelProc.c:402-463
    const float st_freq = 64.0f, smpl = 8000.0f;
    const int fft_len = 256, nch = 23;
    const float start_mel = (float)(2595.0 * std::log10(1.0 + (double)st_freq / 700.0));
    const float half_mel = (float)(2595.0 * std::log10(1.0 + (double)(smpl / 2) / 700.0));
    auto edge = [&](int i) {
        const float arg = start_mel + (float)i / (float)(nch + 1) * (half_mel - start_mel);
        const float f = (float)(700.0 * (std::pow(10.0, (double)arg / 2595.0) - 1.0));
        return (int)((double)((float)fft_len * f / smpl) + 0.5);
    };
    for (int i = 0; i < nch; i++) {
        t.start[i] = edge(i);
        t.len[i] = edge(i + 2) - t.start[i] + 1;
    }
    // ComputeTriangle, etsi/cpp/MelProc.c:482-522
    int upper = 0;
    for (int i = 0; i < nch; i++) {
        const int low = (i < nch - 1) ? t.start[i + 1] - t.start[i] + 1 : upper - t.start[i] + 1;
        const int hgh = t.len[i] - low + 1;
        for (int j = 0; j < low; j++) t.tri[i][j] = (float)(j + 1) / (float)low;
        for (int j = 1; j < hgh; j++) t.tri[i][low + j - 1] = (float)(hgh - j) / (float)hgh;
        upper = t.start[i] + t.len[i] - 1;
    }
}

// First half of the 118-tap QMF prototype, integers over 2^23 (DP_HP_filters, etsi/cpp/16kHzProcessing.c:559-608);
// the prototype is symmetric.
static const int kQmfHalf[59] = {
    1584, 805, -4192, -8985, -5987, 2583, 4657, -3035, -7004, 1542, 8969, 567, -10924, -3757, 12320, 7951, -12793,
    -13048, 11923, 18793, -9331, -24802, 4694, 30570, 2233, -35439, -11526, 38680, 23114, -39474, -36701, 36999,
    51797, -30419, -67658, 18962, 83318, -1927, -97566, -21284, 108971, 51215, -115837, -88430, 116130, 133716,
    -107253, -188497, 85497, 255795, -44643, -342699, -28185, 468096, 167799, -696809, -519818, 1446093, 3562497};

void build_hp16_tables(Hp16Tables &t)
{
    std::memset(&t, 0, sizeof t);
    for (int i = 0; i < 118; i++) t.dp[i] = (float)((double)kQmfHalf[i < 59 ? i : 117 - i] / 8388608.0);
    // InitMelFBwindows(FirstWindow16k, 80.0, 8000, 128, 5, normalize=0), 16kHzProcessing.c:254-259
    const float st_freq = 80.0f, smpl = 8000.0f;
    const int fft_len = 128, nch = 5;
    int c[5];
    const float start_mel = (float)(2595.0 * std::log10(1.0 + (double)st_freq / 700.0));
    const float half_mel = (float)(2595.0 * std::log10(1.0 + (double)(smpl / 2) / 700.0));
    for (int i = 0; i < nch; i++) {
        const float arg = start_mel + (float)i / (float)(nch - 1) * (half_mel - start_mel);
        const float f = (float)(700.0 * (std::pow(10.0, (double)arg / 2595.0) - 1.0));
        c[i] = (int)((double)((float)fft_len * f / smpl) + 0.5);
    }
    for (int i = 0; i < nch; i++) {
        float *d = t.fb[i];
        int len;
        if (i == 0) {
            t.fb_start[i] = c[0];
            len = c[1] - c[0];
            for (int j = 0; j < len; j++) d[j] = (float)(1.0 - (double)((float)j / (float)len));
        } else if (i < nch - 1) {
            const int up = c[i] - c[i - 1];
            t.fb_start[i] = c[i - 1] + 1;
            len = c[i + 1] - c[i - 1] - 1;
            for (int j = 0; j < up; j++) d[j] = (float)(j + 1) / (float)up;
            for (int j = up, k = 0; j < len; j++, k++) d[j] = (float)(1.0 - (double)((float)(k + 1) / (float)(c[i + 1] - c[i])));
        } else {
            t.fb_start[i] = c[nch - 2] + 1;
            len = c[nch - 1] - c[nch - 2];
            for (int j = 0; j < len; j++) d[j] = (float)(j + 1) / (float)len;
        }
        t.fb_len[i] = len;
    }
    // InitDCTMatrix(13, 23 + 3), etsi/cpp/CompCeps.c:279-280
    for (int i = 1; i < 13; i++)
        for (int j = 0; j < 26; j++)
            t.dct26[(i - 1) * 26 + j] = (float)std::cos(kPI * (double)(float)i / (double)(float)26 * ((double)(float)j + 0.5));
}

// BS3383 table 1 as entered in initMiddleEar(), resyth_64sub_IBM/cpp/extractwav.cpp:134-167
static const float kEarF[29] = {20.0f, 25.0f, 31.5f, 40.0f, 50.0f, 63.0f, 80.0f, 100.0f, 125.0f, 160.0f, 200.0f,
    250.0f, 315.0f, 400.0f, 500.0f, 630.0f, 800.0f, 1000.0f, 1250.0f, 1600.0f, 2000.0f, 2500.0f, 3150.0f, 4000.0f,
    5000.0f, 6300.0f, 8000.0f, 10000.0f, 12500.0f};
static const float kEarAf[29] = {2.347f, 2.190f, 2.050f, 1.879f, 1.724f, 1.579f, 1.512f, 1.466f, 1.426f, 1.394f,
    1.372f, 1.344f, 1.304f, 1.256f, 1.203f, 1.135f, 1.062f, 1.000f, 0.967f, 0.943f, 0.932f, 0.933f, 0.937f, 0.952f,
    0.974f, 1.027f, 1.135f, 1.266f, 1.501f};
static const float kEarBf[29] = {0.00561f, 0.00527f, 0.00481f, 0.00404f, 0.00383f, 0.00286f, 0.00259f, 0.00257f,
    0.00256f, 0.00255f, 0.00254f, 0.00248f, 0.00229f, 0.00201f, 0.00162f, 0.00111f, 0.00052f, 0.00000f, -0.00039f,
    -0.00067f, -0.00092f, -0.00105f, -0.00104f, -0.00088f, -0.00055f, 0.00000f, 0.00089f, 0.00211f, 0.00488f};
static const float kEarTf[29] = {74.3f, 65.0f, 56.3f, 48.4f, 41.7f, 35.5f, 29.8f, 25.1f, 20.7f, 16.8f, 13.8f, 11.2f,
    8.9f, 7.2f, 6.0f, 5.0f, 4.4f, 4.2f, 3.7f, 2.6f, 1.0f, -1.2f, -3.6f, -3.9f, -1.1f, 6.6f, 15.3f, 16.4f, 11.6f};

static float loudness_phons(float freq)
{
    // loudnessLevelInPhons, extractwav.cpp:259-279, DB = 60.0
    int i = 0;
    while (kEarF[i] < freq) i++;
    const float ratio = (freq - kEarF[i - 1]) / (kEarF[i] - kEarF[i - 1]);
    const float afy = kEarAf[i - 1] + ratio * (kEarAf[i] - kEarAf[i - 1]);
    const float bfy = kEarBf[i - 1] + ratio * (kEarBf[i] - kEarBf[i - 1]);
    const float tfy = kEarTf[i - 1] + ratio * (kEarTf[i] - kEarTf[i - 1]);
    return (float)(4.2 + (double)afy * (60.0 - (double)tfy) / (1.0 + (double)bfy * (60.0 - (double)tfy)));
}

void build_rfft_plan(RfftPlan &p, int n, int m, bool float_trig)
{
    p.n = n;
    p.m = m;
    p.ops.clear();
    p.tw.clear();
    // length-two butterflies, rfft.c:77-92
    p.level_begin[0] = 0;
    p.tw_begin[0] = 0;
    for (int is = 0, id = 4; is < n - 1; is = 2 * id - 2, id *= 4)
        for (int i = is; i < n; i += id) p.ops.push_back((unsigned)i);
    // L-shaped butterflies, rfft.c:94-179: the blocks a level visits, once for the twiddle-free part and once per j
    for (int level = 1; level < m; level++) {
        const int n2 = 2 << level, n8 = n2 >> 3;
        p.level_begin[level] = (int)p.ops.size();
        p.tw_begin[level] = (int)(p.tw.size() / 4);
        const float e = (float)((3.14159265358979323846 * 2) / n2);
        for (int j = 0; j < (n8 > 1 ? n8 : 1); j++) {
            if (j >= 1) {
                const float a = (float)j * e, a3 = 3 * a;
                if (float_trig) {
                    p.tw.push_back(cosf(a)); p.tw.push_back(sinf(a)); p.tw.push_back(cosf(a3)); p.tw.push_back(sinf(a3));
                } else {
                    p.tw.push_back((float)std::cos((double)a)); p.tw.push_back((float)std::sin((double)a));
                    p.tw.push_back((float)std::cos((double)a3)); p.tw.push_back((float)std::sin((double)a3));
                }
            }
            for (int is = 0, id = 2 * n2; is < n; is = 2 * id - n2, id *= 4)
                for (int i = is; i < n; i += id) p.ops.push_back((unsigned)i | ((unsigned)j << 16));
        }
    }
    p.level_begin[m] = (int)p.ops.size();
    p.tw_begin[m] = (int)(p.tw.size() / 4);
}

void build_gt_tables(GtTables &t)
{
    // channel set-up inline in resynth()/subbband(), extractwav.cpp:41-54
    const float lower = (float)(21.4 * std::log10(50 * 0.00437 + 1.0));
    const float upper = (float)(21.4 * std::log10(8000 * 0.00437 + 1.0));
    const float space = (upper - lower) / (float)(kGtChannels - 1);
    // gammaToneFilter set-up, extractwav.cpp:176-183 (C++: exp/cos/sin of float are float overloads)
    const float dt = 1 / (float)16000;
    const float two_pi_t = (float)(2 * 3.1415926535897932384626433832795 * (double)dt);
    for (int c = 0; c < kGtChannels; c++) {
        const float cf = (float)((std::pow(10.0, (double)(lower + (float)c * space) / 21.4) - 1.0) / 0.00437);
        t.cf[c] = cf;
        t.bw[c] = (float)(24.7 * ((double)cf * 0.00437 + 1.0) * 1.019);
        const float phon = (float)((double)loudness_phons(cf) - 60.0);
        t.mid[c] = (float)std::pow(10.0, (double)(phon / 20));
        t.gain[c] = (float)((double)t.mid[c] * std::pow((double)(two_pi_t * t.bw[c]), 4.0) / 3.0);
        const float z = expf(-two_pi_t * t.bw[c]);
        t.f1[c] = cosf(cf * two_pi_t) * z;
        t.f2[c] = sinf(cf * two_pi_t) * z;
    }
    // hairCell constants, extractwav.cpp:220-234 (Meddis 1988, HuWang.h:38-46)
    const double Y = 5.05, G = 2000.0, L = 2500.0, R = 6580.0, X = 66.31, A = 3.0, B = 300.0, H = 48000.0, M = 1.0;
    const float ymdt = (float)(Y * M * (double)dt), xdt = (float)(X * (double)dt), ydt = (float)(Y * (double)dt);
    const float lplusrdt = (float)((L + R) * (double)dt), rdt = (float)(R * (double)dt), gdt = (float)(G * (double)dt);
    const float hdt = (float)H;
    const float kt = (float)(G * A / (A + B));
    const float c0 = (float)(M * Y * (double)kt / (L * (double)kt + Y * (L + R)));
    const float q0 = (float)((double)c0 * (L + R) / (double)kt);
    const float w0 = (float)((double)c0 * R / X);
    const float hc[16] = {ymdt, xdt, ydt, lplusrdt, rdt, gdt, hdt, c0, q0, w0, 0, 0, 0, 0, 0, 0};
    std::memcpy(t.hc, hc, sizeof hc);
    // raised-cosine halves of resynth(), extractwav.cpp:100-106 (double, PI from HuWang.h:7)
    const double PI = 3.1415926535897932384626433832795;
    for (int k = 0; k < 160; k++) {
        t.rise_d[k] = 0.5 * (1.0 + std::cos(k * PI / (160) + PI));
        t.fall_d[k] = 0.5 * (1.0 + std::cos(k * PI / (160)));
    }
}

}  // namespace se
