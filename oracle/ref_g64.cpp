/* TEST INFRASTRUCTURE — not part of the product.
 *
 * oracle/_ref driver #3: the reference's 64-channel gammatone code, compiled IN PLACE from
 * /root/reference with oracle/shims/{Wave.h,aurora/aurora_include.h} on the include path.
 * One source, four builds (oracle/Makefile), each including ONE reference extractwav.cpp:
 *   -DSE_REF_SUBBAND : enhancement_extract_subband_linux/cpp/extractwav.cpp  (addnoise, subbband)
 *   -DSE_REF_RESYNTH : resyth_64sub_IBM/cpp | resyth_64sub_ori/cpp | 1dnn_resynth  (resynth)
 * All arithmetic is reference code; ours is the marshalling and the CWave::Write capture sink.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <vector>
#include <string>
#include <unistd.h>
#include <fcntl.h>
#include "Wave.h"

struct captured_wave {
    std::string path;
    std::vector<short> data;
    int fs;
};
static std::vector<captured_wave> g_sink;

extern "C" void se_shim_wave_sink(const char *path, const short *data, long n, int fs)
{
    captured_wave w;
    w.path = path ? path : "";
    w.data.assign(data, data + n);
    w.fs = fs;
    g_sink.push_back(w);
}

#include "extractwav.cpp" /* resolved through -I<variant dir> */

static int silence_stdout(void)
{
    fflush(stdout);
    std::cout.flush();
    int saved = dup(1);
    int nul = open("/dev/null", O_WRONLY);
    dup2(nul, 1);
    close(nul);
    return saved;
}
static void restore_stdout(int saved)
{
    std::cout.flush();
    fflush(stdout);
    dup2(saved, 1);
    close(saved);
}

extern "C" {

/* gammaToneFilter, extractwav.cpp:168 (resynth dirs) / :133 (subband dir) */
void ref_gammatone(float *in, float *out, float cf, float bw, float mid, long n)
{
    gammaTone g;
    memset(&g, 0, sizeof g);
    g.cf = cf;
    g.bw = bw;
    g.midEarCoeff = mid;
    gammaToneFilter(in, out, g, n);
}

/* hairCell, extractwav.cpp:213 / :178 */
void ref_haircell(float *in, float *out, long n) { hairCell(in, out, n); }

/* The per-channel parameter block is inline in resynth()/subbband() (extractwav.cpp:41-54 /
 * :58-71), not a function.  Restated here 1:1 on top of the reference's own
 * initMiddleEar()/loudnessLevelInPhons() so the 64 (cf,bw,midEarCoeff) triples can be dumped. */
void ref_channel_params(float *cf_out, float *bw_out, float *mid_out)
{
    gammaTone fChan[NUMBER_CHANNEL];
    float lowerERB, upperERB, spaceERB, cf, phon;
    middleEar loudFunc = initMiddleEar();
    lowerERB = 21.4 * log10(MINCF * 0.00437 + 1.0);
    upperERB = 21.4 * log10(MAXCF * 0.00437 + 1.0);
    spaceERB = (NUMBER_CHANNEL > 1) ? (upperERB - lowerERB) / (NUMBER_CHANNEL - 1) : 0;
    for (int chan = 0; chan < NUMBER_CHANNEL; chan++) {
        cf = (pow(10, (lowerERB + chan * spaceERB) / 21.4) - 1) / 0.00437;
        fChan[chan].cf = cf;
        fChan[chan].bw = 24.7 * (cf * 0.00437 + 1.0) * BW_CORRECTION;
        phon = (loudnessLevelInPhons(cf, loudFunc) - DB);
        fChan[chan].midEarCoeff = pow(10, phon / 20);
        cf_out[chan] = fChan[chan].cf;
        bw_out[chan] = fChan[chan].bw;
        mid_out[chan] = fChan[chan].midEarCoeff;
    }
}

#ifdef SE_REF_SUBBAND
/* subbband, enhancement_extract_subband_linux/cpp/extractwav.cpp:36-97.
 * out64 receives the 64 captured channel wavs, channel-major [64][n]. */
int ref_subband(short *wav, long n, short *out64)
{
    asdk::CWave w(wav, n, 16000);
    char path[4] = "", tail[4] = "", id[4] = "u";
    g_sink.clear();
    subbband(&w, path, tail, id);
    if (g_sink.size() != 64) return -1;
    for (int c = 0; c < 64; c++) memcpy(out64 + (long)c * n, g_sink[c].data.data(), n * sizeof(short));
    g_sink.clear();
    return 0;
}

/* addnoise, extractwav.cpp:6-35.  noise[] is scaled in place (as in the reference), noisy[] is
 * written; the wav "file" the reference writes is captured and dropped. */
int ref_addnoise(short *pure, short *noise, short *noisy, long n, int db)
{
    asdk::CWave p(pure, n, 16000), q(noise, n, 16000), r(noisy, n, 16000);
    FILESNAME opts;
    memset(&opts, 0, sizeof opts);
    opts.addnoisedB = db;
    char id[4] = "u";
    g_sink.clear();
    addnoise(&p, &q, &r, id, opts);
    g_sink.clear();
    return 0;
}
#endif

#ifdef SE_REF_RESYNTH
/* resynth, <variant>/extractwav.cpp:9-132.  mask is [num_frame][64] row-major; the IBM variant
 * overwrites gated entries with 1.0 (extractwav.cpp:99) and that mutation is copied back. */
int ref_resynth(short *wav, long n, float *mask, int num_frame, short *out)
{
    asdk::CWave w(wav, n, 16000);
    FILESNAME opts;
    memset(&opts, 0, sizeof opts);
    std::vector<std::vector<float> > v(num_frame, std::vector<float>(64));
    for (int f = 0; f < num_frame; f++)
        for (int c = 0; c < 64; c++) v[f][c] = mask[f * 64 + c];
    char id[4] = "u";
    g_sink.clear();
    int saved = silence_stdout();
    resynth(&w, opts, v, id);
    restore_stdout(saved);
    if (g_sink.size() != 1 || (long)g_sink[0].data.size() != n) return -1;
    memcpy(out, g_sink[0].data.data(), n * sizeof(short));
    for (int f = 0; f < num_frame; f++)
        for (int c = 0; c < 64; c++) mask[f * 64 + c] = v[f][c];
    g_sink.clear();
    return 0;
}
#endif

} /* extern "C" */
