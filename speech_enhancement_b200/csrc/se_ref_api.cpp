// libse_b200_dropin.so — the reference's function names for the hot path, implemented as thin
// adapters over libse_b200.so (include/se_reference_api.h cites what each one replaces).
// Host C++ only; no arithmetic of the path lives here except int16<->float conversions and the
// text formatting the reference does around its kernels.
#include "se_b200.h"
#include "se_reference_api.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#define SE_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

std::mutex g_mu;
std::vector<se_ctx *> g_ctxs;                 // every context ever handed out (kept until the process ends)
thread_local se_ctx *t_ctx = nullptr;

// The reference's entry points carry no handle, and its batch tool calls them from N worker threads at once
// (aurora_speech_enhancement.cpp:110-120, 312-328).  A se_ctx serves one host thread at a time, so every calling
// thread gets its own (own stream, own staging buffers): calls from different threads run concurrently on the GPU.
// Contexts are not torn down at thread exit -- handles created from them (FEParamsX, plugin thread instances) may
// outlive the thread.  Fails loudly without a GPU.
se_ctx *default_ctx(const char *who)
{
    if (!t_ctx) {
        int dev = 0;
        if (const char *e = std::getenv("SE_B200_DEVICE")) dev = std::atoi(e);
        se_ctx *c = nullptr;
        int rc = se_ctx_create(dev, &c);
        if (rc != SE_OK) {
            std::fprintf(stderr, "%s: libse_b200 cannot run (%d): %s\n", who, rc, se_last_error());
            return nullptr;
        }
        std::lock_guard<std::mutex> lk(g_mu);
        g_ctxs.push_back(c);
        t_ctx = c;
    }
    return t_ctx;
}

// The sub-band adapters default to the bit-exact kernels (one serial work item per utterance).  SE_B200_FAST=1 selects
// the time-chunked default-mode kernels instead (int16 within 1 LSB of the reference, >= 60 dB; ~20x lower latency for
// a lone 10 s resynthesis); every adapter call carries ONE utterance, so resynthesis also asks for the short chunks of
// SE_FLAG_LATENCY.
int g64_flags()
{
    static const int fast = [] { const char *e = std::getenv("SE_B200_FAST"); return e && e[0] == '1'; }();
    return fast ? 0 : SE_FLAG_EXACT;
}

int report(const char *who, int rc)
{
    if (rc != SE_OK) std::fprintf(stderr, "%s: error %d: %s\n", who, rc, se_last_error());
    return rc;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// etsi/cpp/AdvFrontEnd.h
// ---------------------------------------------------------------------------------------------
SE_EXPORT int etsi_denoise(short *p_data, short *p_denoised, long i_frame)
{
    se_ctx *ctx = default_ctx("etsi_denoise");
    if (!ctx || !p_data || !p_denoised || i_frame < 0) return 1;
    const long body = i_frame - i_frame % 80;       // the tail is never read nor written (AdvFrontEnd.c:96-97)
    if (body == 0) return 0;
    const int64_t off[2] = {0, body};
    int rc = se_etsi_denoise_batch(ctx, p_data, off, 1, p_denoised, nullptr, nullptr, nullptr, nullptr, nullptr,
                                   nullptr, SE_MEM_HOST);
    if (report("etsi_denoise", rc) != SE_OK) return 1;
    return 0;
}

SE_EXPORT int etsi_denoise_synchronization(short *p_data, short *p_denoised, long i_frame)
{
    // AdvFrontEnd.c:213-226, quirks included: copies only when etsi_denoise returned TRUE (failure),
    // and the length is in BYTES.
    if (i_frame < 0) return 1;
    short *tmp = (short *)std::calloc((size_t)(i_frame > 0 ? i_frame : 1), sizeof(short));
    if (!tmp) return 1;
    const int flag = etsi_denoise(p_data, tmp, i_frame);
    if (flag && i_frame > 320) {
        std::memset(p_denoised, 0, (size_t)i_frame * sizeof(short));
        std::memcpy(p_denoised, tmp + 320, (size_t)(i_frame - 320));
    }
    std::free(tmp);
    return flag;
}

SE_EXPORT int etsi_denoise_16k(short *, short *, long)
{
    std::fprintf(stderr, "etsi_denoise_16k: unsupported (the reference reads 160 samples per call out of an 80-sample "
                         "buffer, AdvFrontEnd.c:264,290-293; use se_etsi16k_features_batch)\n");
    return 1;
}

SE_EXPORT int etsi_denoise_16k_synchronization(short *, short *, long)
{
    std::fprintf(stderr, "etsi_denoise_16k_synchronization: unsupported (see etsi_denoise_16k)\n");
    return 1;
}

// ---------------------------------------------------------------------------------------------
// aurora_speech_enhancement.cpp:25-79 (BufferDenoise) and :172-184 (RMS normalisation)
// ---------------------------------------------------------------------------------------------
namespace {
int buffer_denoise(const short *src, short *des, int samples, int bl, short *work)
{
    const int kOff = 320;                                    // etsi_offset_points, :23
    if (!src || !des || !work || bl <= 0 || samples < kOff) return SE_EINVAL;
    se_ctx *ctx = default_ctx("BufferDenoise");
    if (!ctx) return SE_ENODEVICE;
    const int times = samples / bl, left = samples % bl;
    // the windows the reference hands to etsi_denoise one after the other: (start, length)
    std::vector<int64_t> start, off(1, 0);
    auto add = [&](int64_t s, int64_t n) { start.push_back(s); off.push_back(off.back() + n); };
    if (times == 0) add(0, samples);
    else {
        add(0, bl);
        for (int i = 1; i < times; i++) add((int64_t)i * bl - kOff, bl + kOff);
        add((int64_t)times * bl - kOff, left + kOff);
    }
    const int nw = (int)start.size();
    std::vector<short> in((size_t)off.back()), out((size_t)off.back());
    for (int w = 0; w < nw; w++) std::memcpy(&in[off[w]], src + start[w], (size_t)(off[w + 1] - off[w]) * sizeof(short));
    int rc = se_etsi_denoise_batch(ctx, in.data(), off.data(), nw, out.data(), nullptr, nullptr, nullptr, nullptr, nullptr,
                                   nullptr, SE_MEM_HOST);
    if (report("BufferDenoise", rc) != SE_OK) return rc;
    // replay the reference's sequence of writes into `work` and copies out of it
    for (int w = 0; w < nw; w++) {
        const int64_t len = off[w + 1] - off[w];
        std::memcpy(work, &out[off[w]], (size_t)(len - len % 80) * sizeof(short));     // etsi_denoise leaves the tail
        if (times == 0) {
            std::memcpy(des, work + kOff, (size_t)(samples - kOff) * sizeof(short));
        } else if (w == 0) {
            std::memcpy(des, work + kOff, (size_t)(bl - kOff) * sizeof(short));
        } else if (w < times) {
            std::memcpy(des + (int64_t)w * bl - kOff, work + kOff, (size_t)bl * sizeof(short));
        } else {
            std::memcpy(des + (int64_t)times * bl - kOff, work + kOff, (size_t)left * sizeof(short));
        }
    }
    std::memset(des + samples - kOff, 0, kOff * sizeof(short));
    return SE_OK;
}
}  // namespace

SE_EXPORT void BufferDenoise(short *src_wave, short *des_wave, int samples, int batch_length, short *work)
{
    buffer_denoise(src_wave, des_wave, samples, batch_length, work);
}

SE_EXPORT int se_buffer_denoise(const short *src_wave, short *des_wave, int samples, int batch_length)
{
    if (batch_length == 0) batch_length = 8000 * 60 * 3;     // :22
    if (batch_length < 0) return SE_EINVAL;
    std::vector<short> work((size_t)batch_length + 320, 0);
    return buffer_denoise(src_wave, des_wave, samples, batch_length, work.data());
}

SE_EXPORT double se_rms_normalise(short *p_data, int samples, float target)
{
    double energy = 0.0;
    for (int t = 0; t < samples; t++) energy += p_data[t] * p_data[t];      // int product, double sum (:175-178)
    energy /= samples;
    energy = std::sqrt(energy);
    const double factor = target / energy;
    for (int t = 0; t < samples; t++) {
        const double v = (float)p_data[t] * factor;
        // (short) of a double on x86-64: cvttsd2si to int32 (0x80000000 when out of range or NaN), low 16 bits
        p_data[t] = (v > -2147483649.0 && v < 2147483648.0) ? (short)(int)v : (short)0;
    }
    return factor;
}

SE_EXPORT void rfft(float *x, int n, int m)
{
    se_ctx *ctx = default_ctx("rfft");
    if (!ctx) return;
    report("rfft", se_rfft_batch(ctx, x, n, m, 1, SE_MEM_HOST));
}

// ---------------------------------------------------------------------------------------------
// etsi/cpp/ParmInterface.h: AdvProcessAlloc / Init / DoAdvProcess / Delete on an opaque handle
// ---------------------------------------------------------------------------------------------
struct NoiseSupStructX { se_ns_stream *ns; };     // what DoNoiseSupAlloc hands out: one NoiseSup state on the GPU

struct FEParamsX {
    NoiseSupStructX *NSX;     // the member AdvProcessAlloc fills with DoNoiseSupAlloc() (ParmInterface.c:83)
    se_ns_stream *ns;         // == NSX->ns
    int feat_ready;
    int gate;          // zero-frame gate currently set on the stream (1 = DoAdvProcess, 0 = bare DoNoiseSup)
    // as-shipped DoAdvProcess never reaches DoVADProc (ParmInterface.c:274-311 is commented out), so all
    // FlushAdvProcess / DoVADFlush (VAD.c:342-431) ever see is the freshly initialised VAD structure:
    int ns_frames;            // vadNS.nbFrame[0]
    int frame_counter;        // FEParamsX.FrameCounter as DoNoiseSup last published it (NoiseSup.c:1366)
    int vad_focus, vad_flush_focus;
};

SE_EXPORT NoiseSupStructX *DoNoiseSupAlloc(void)
{
    se_ctx *ctx = default_ctx("DoNoiseSupAlloc");
    if (!ctx) return nullptr;
    NoiseSupStructX *n = new NoiseSupStructX();
    n->ns = nullptr;
    if (report("DoNoiseSupAlloc", se_ns_stream_create(ctx, 8, &n->ns)) != SE_OK) { delete n; return nullptr; }
    return n;
}

SE_EXPORT void DoNoiseSupDelete(NoiseSupStructX *NSX)
{
    if (!NSX) return;
    se_ns_stream_destroy(NSX->ns);
    delete NSX;
}

SE_EXPORT FEParamsX *AdvProcessAlloc(int SamplingFrequency)
{
    if (SamplingFrequency != 8000) {
        std::fprintf(stderr, "AdvProcessAlloc: only 8000 Hz is on the accelerated path (got %d)\n", SamplingFrequency);
        return nullptr;
    }
    se_ctx *ctx = default_ctx("AdvProcessAlloc");
    if (!ctx) return nullptr;
    FEParamsX *p = new FEParamsX();
    p->NSX = DoNoiseSupAlloc();                  // ParmInterface.c:83
    if (!p->NSX) { delete p; return nullptr; }
    p->ns = p->NSX->ns;
    p->feat_ready = 0;
    se_ns_stream_set_zero_gate(p->ns, 1);        // DoAdvProcess skips leading all-zero frames
    p->gate = 1;
    p->ns_frames = p->frame_counter = 0;
    p->vad_focus = 0;
    p->vad_flush_focus = -1;
    return p;
}

SE_EXPORT void AdvProcessInit(FEParamsX *This)
{
    if (!This) return;
    se_ns_stream_reset(This->ns);
    This->feat_ready = 0;
    This->ns_frames = This->frame_counter = 0;   // ParmInterface.c:184
    This->vad_focus = 0;                         // DoVADInit, VAD.c:137-142
    This->vad_flush_focus = -1;
}

SE_EXPORT int DoAdvProcess(short *CurrentFrame, short *OutBuf, float *FeatBuf, FEParamsX *This)
{
    if (!This || !CurrentFrame) return 0;
    float in[80], out[80], ceps[14];
    for (int i = 0; i < 80; i++) in[i] = (float)CurrentFrame[i];      // ParmInterface.c:239-242
    int32_t produced = 0, n_ceps = 0;
    This->feat_ready = 0;
    if (This->gate != 1) { se_ns_stream_set_zero_gate(This->ns, 1); This->gate = 1; }
    if (report("DoAdvProcess", se_ns_stream_push(This->ns, in, 1, out, &produced, nullptr, nullptr, nullptr,
                                                 FeatBuf ? ceps : nullptr, &n_ceps)) != SE_OK)
        return 0;
    if (produced & 4) {
        // leading all-zero frame (ParmInterface.c:314-327): null MFCC vector FeatureBuffer[0..NUM_CEP_COEFF],
        // VAD = NON_SPEECH_FRAME, return TRUE; DenoiseBuffer is not touched
        if (FeatBuf) std::memset(FeatBuf, 0, 14 * sizeof(float));
        return 1;
    }
    if (produced & 2) This->ns_frames++;                          // vadNS.nbFrame[0]: stage-1 frames so far
    if (produced & 3) This->frame_counter = This->ns_frames;      // published by every stage that ran (NoiseSup.c:1366)
    if (!(produced & 1)) return 0;
    if (OutBuf)
        for (int i = 0; i < 80; i++) OutBuf[i] = (short)(int)out[i];   // ParmInterface.c:264-268: C cast
    if (FeatBuf && n_ceps > 0) { std::memcpy(FeatBuf, ceps, sizeof ceps); This->feat_ready = 1; }
    return 1;
}

// FlushAdvProcess (ParmInterface.c:347-355) -> DoVADFlush (VAD.c:342-431) on the VAD structure as DoVADInit left it
// (the as-shipped DoAdvProcess never feeds it): the 7-slot delay line is all zeros and every counter is 0, so the
// flush walks Focus once around the ring -- TRUE six times, then FALSE -- incrementing FrameCounter, and from
// FrameCounter > BUFFER_SIZE + 3 = 10 on it copies an all-zero row with VAD flag 0 into FeatureBuffer[0..14].
SE_EXPORT int FlushAdvProcess(float *FeatBuf, FEParamsX *This)
{
    if (!This) return 0;
    if (This->vad_flush_focus == -1) This->vad_flush_focus = This->vad_focus;
    int focus = This->vad_focus + 1;
    if (focus == 7) focus = 0;
    if (focus == This->vad_flush_focus) return 0;
    This->frame_counter++;
    if (This->frame_counter > 7 + 3 && FeatBuf) std::memset(FeatBuf, 0, 15 * sizeof(float));
    This->vad_focus = focus;
    return 1;
}

SE_EXPORT int se_adv_feature_ready(FEParamsX *This) { return This ? This->feat_ready : 0; }

// etsi/cpp/NoiseSupExports.h:20-26 / CompCepsExports.h:17-20: the module entry points DoAdvProcess reaches
// through FEParamsX's function pointers, callable directly on the opaque handle.
SE_EXPORT void DoNoiseSupInit(FEParamsX *This) { AdvProcessInit(This); }

SE_EXPORT int DoNoiseSup(float *InData, float *OutData, FEParamsX *This)
{
    if (!This || !InData || !OutData) return 0;
    if (This->gate != 0) { se_ns_stream_set_zero_gate(This->ns, 0); This->gate = 0; }   // no zero-frame gate here
    int32_t produced = 0;
    float out[80];
    if (report("DoNoiseSup", se_ns_stream_push(This->ns, InData, 1, out, &produced, nullptr, nullptr, nullptr,
                                               nullptr, nullptr)) != SE_OK)
        return 0;
    if (!(produced & 1)) return 0;            // latency: OutData untouched, like the reference
    std::memcpy(OutData, out, sizeof out);
    return 1;
}

SE_EXPORT void DoCompCepsInit(FEParamsX *) {}   // tables are built with the context

SE_EXPORT int DoCompCeps(float *Data, float *Coef, FEParamsX *)
{
    if (!Data || !Coef) return 0;
    se_ctx *ctx = default_ctx("DoCompCeps");
    if (!ctx) return 0;
    // Data[-1] must be valid (pre-emphasis, CompCeps.c:427-429): the caller passes frameBuf + 1
    return report("DoCompCeps", se_comp_ceps_batch(ctx, Data - 1, 1, Coef, SE_MEM_HOST)) == SE_OK ? 1 : 0;
}

SE_EXPORT void AdvProcessDelete(FEParamsX **This)
{
    if (!This || !*This) return;
    DoNoiseSupDelete((*This)->NSX);              // ParmInterface.c:376
    delete *This;
    *This = nullptr;
}

// ---------------------------------------------------------------------------------------------
// function/.../aurora_etsi/NoiseSupExports.h: the plugin API
// ---------------------------------------------------------------------------------------------
namespace {
struct GlobalImpl { int SamplingFrequency; };
}

SE_EXPORT int32s etsi_denoise_mapping_global_init(PINSTANCE sm_glb_pins, INSTANCE)
{
    if (!sm_glb_pins) return 0;
    GlobalImpl *g = new GlobalImpl();
    g->SamplingFrequency = 16000;                  // NoiseSup.cpp:918
    *sm_glb_pins = g;
    return 1;
}

SE_EXPORT int32s etsi_denoise_mapping_thread_init(PINSTANCE sm_thd_pins, INSTANCE sm_glb_ins)
{
    if (!sm_thd_pins || !sm_glb_ins) return 0;     // NoiseSup.cpp:941-944
    se_ctx *ctx = default_ctx("etsi_denoise_mapping_thread_init");
    if (!ctx) return 0;
    se_ns_stream *s = nullptr;
    if (report("etsi_denoise_mapping_thread_init", se_ns_stream_create(ctx, 16, &s)) != SE_OK) return 0;
    *sm_thd_pins = s;
    return 1;
}

SE_EXPORT int32s etsi_denoise_mapping_func_Wiener(INSTANCE, INSTANCE sm_thd_ins, INSTANCE in_ins, INSTANCE out_ins,
                                                  INSTANCE fp_Wiener)
{
    se_ns_stream *s = (se_ns_stream *)sm_thd_ins;
    esti_denoise_in *in = (esti_denoise_in *)in_ins;
    esti_denoise_out *po = (esti_denoise_out *)out_ins;
    if (!s || !in || !po) return 1;
    const int nb = in->dataNum / 160;
    if (nb <= 0) return 0;
    std::vector<float> out((size_t)nb * 160), gains((size_t)nb * 25);
    std::vector<int32_t> produced(nb), flags((size_t)nb * 5);
    int32_t rows = 0;
    if (report("etsi_denoise_mapping_func_Wiener",
               se_ns_stream_push(s, in->inData, nb, out.data(), produced.data(), flags.data(), gains.data(), &rows,
                                 nullptr, nullptr)) != SE_OK)
        return 1;
    for (int b = 0; b < nb; b++) {
        const int p = produced[b];
        if (p & 4) continue;                                          // zero block: nothing touched (:1166-1170)
        if ((p & 1) && po->outData) std::memcpy(po->outData + (size_t)b * 160, out.data() + (size_t)b * 160, 160 * sizeof(float));
        if (p & 2) {
            if (po->pSpeechFoundVar) po->pSpeechFoundVar[b] = flags[0 * (size_t)nb + b];
            if (po->pSpeechFoundSpec) po->pSpeechFoundSpec[b] = flags[1 * (size_t)nb + b];
            if (po->pSpeechFoundMel) po->pSpeechFoundMel[b] = flags[2 * (size_t)nb + b];
            if (po->pSpeechFoundVADNS) po->pSpeechFoundVADNS[b] = flags[3 * (size_t)nb + b];
        }
        if ((p & 3) && po->pFrameCounter) po->pFrameCounter[b] = flags[4 * (size_t)nb + b];
    }
    if (FILE *fp = (FILE *)fp_Wiener) {
        for (int r = 0; r < rows; r++) {                              // NoiseSup.cpp:1318-1326
            for (int i = 0; i < 25; i++) std::fprintf(fp, "%f ", gains[(size_t)r * 25 + i]);
            std::fprintf(fp, "\n");
        }
    }
    return 0;
}

SE_EXPORT void etsi_denoise_mapping_thread_release(PINSTANCE sm_thd_pins)
{
    if (!sm_thd_pins || !*sm_thd_pins) return;
    se_ns_stream_destroy((se_ns_stream *)*sm_thd_pins);
    *sm_thd_pins = nullptr;
}

SE_EXPORT void etsi_denoise_mapping_global_release(PINSTANCE sm_glb_pins)
{
    if (!sm_glb_pins || !*sm_glb_pins) return;
    delete (GlobalImpl *)*sm_glb_pins;
    *sm_glb_pins = nullptr;
}

// ---------------------------------------------------------------------------------------------
// */cpp/extractwav.cpp C twins
// ---------------------------------------------------------------------------------------------
SE_EXPORT int se_subband(const short *wav, long n, short *out64xn)
{
    se_ctx *ctx = default_ctx("se_subband");
    if (!ctx) return SE_ENODEVICE;
    if (n <= 0) return SE_OK;
    const int64_t off[2] = {0, n};
    return report("se_subband", se_subband_batch(ctx, wav, off, 1, out64xn, SE_FLAG_EXACT, SE_MEM_HOST));
}

SE_EXPORT int se_resynth(const short *wav, long n, float *mask, int num_frame, int mode, short *out)
{
    se_ctx *ctx = default_ctx("se_resynth");
    if (!ctx) return SE_ENODEVICE;
    if (n <= 0) return SE_OK;
    const int64_t off[2] = {0, n}, fo[2] = {0, num_frame};
    const int flags = g64_flags() ? g64_flags() : SE_FLAG_LATENCY;
    int rc = report("se_resynth", se_resynth_batch(ctx, wav, off, 1, mask, fo, mode, flags, out, SE_MEM_HOST));
    if (rc == SE_OK && mode == SE_RESYNTH_IBM && mask)
        for (long i = 0; i < (long)num_frame * 64; i++)
            if (mask[i] > 0.5f) mask[i] = 1.0f;                       // resyth_64sub_IBM/cpp/extractwav.cpp:99
    return rc;
}

SE_EXPORT int se_single_ibm(const short *pure, const short *noise, long n, float *irm, unsigned char *ibm)
{
    se_ctx *ctx = default_ctx("se_single_ibm");
    if (!ctx) return SE_ENODEVICE;
    const int64_t frames = se_ibm_num_frames(n);
    if (frames <= 0) return SE_OK;
    const int64_t off[2] = {0, n}, fo[2] = {0, frames};
    std::vector<uint64_t> bits(ibm ? (size_t)frames : 0);
    int rc = report("se_single_ibm", se_ibm_batch(ctx, pure, noise, off, 1, irm, ibm ? bits.data() : nullptr, fo, SE_MEM_HOST));
    if (rc == SE_OK && ibm)
        for (int64_t f = 0; f < frames; f++)
            for (int c = 0; c < 64; c++) ibm[f * 64 + c] = (unsigned char)((bits[(size_t)f] >> c) & 1u);
    return rc;
}

SE_EXPORT int se_addnoise(const short *pure, short *noise, short *noisy, long n, int db)
{
    se_ctx *ctx = default_ctx("se_addnoise");
    if (!ctx) return SE_ENODEVICE;
    if (n <= 0) return SE_OK;
    const int64_t off[2] = {0, n};
    return report("se_addnoise", se_addnoise_batch(ctx, pure, noise, noisy, off, 1, db, SE_MEM_HOST));
}

SE_EXPORT int se_hair_cell(const float *input, float *output, long sigLength)
{
    se_ctx *ctx = default_ctx("se_hair_cell");
    if (!ctx) return SE_ENODEVICE;
    return report("se_hair_cell", se_haircell_batch(ctx, input, output, sigLength, 1, SE_MEM_HOST));
}

SE_EXPORT int se_gammatone_filter(const float *input, float *output, int channel, long sigLength)
{
    se_ctx *ctx = default_ctx("se_gammatone_filter");
    if (!ctx) return SE_ENODEVICE;
    const int32_t ch = channel;
    return report("se_gammatone_filter", se_gammatone_batch(ctx, input, output, &ch, sigLength, 1, SE_FLAG_EXACT, SE_MEM_HOST));
}

// gammaToneFilter with the channel passed by value (HuWang.h:49): any cf / bw / midEarCoeff.  The three derived
// coefficients follow extractwav.cpp:176-183 (C++: exp / cos / sin / pow of float arguments as the reference compiles
// them -- float overloads for exp/cos/sin, double pow).
SE_EXPORT int se_gammatone_filter_params(const float *input, float *output, float cf, float bw, float midEarCoeff,
                                         long sigLength)
{
    se_ctx *ctx = default_ctx("se_gammatone_filter_params");
    if (!ctx) return SE_ENODEVICE;
    const float dt = 1 / (float)16000;
    const float two_pi_t = (float)(2 * 3.1415926535897932384626433832795 * (double)dt);
    float coef[3];
    const float z = expf(-two_pi_t * bw);
    coef[0] = cosf(cf * two_pi_t) * z;
    coef[1] = sinf(cf * two_pi_t) * z;
    coef[2] = (float)((double)midEarCoeff * std::pow((double)(two_pi_t * bw), 4.0) / 3.0);
    return report("se_gammatone_filter_params",
                  se_gammatone_coef_batch(ctx, input, output, coef, sigLength, 1, SE_FLAG_EXACT, SE_MEM_HOST));
}

// make_single_IBM's arithmetic on the sub-band rows it reads back from disk (show_IBM.cpp:118-170)
SE_EXPORT int se_single_ibm_rows(const short *pure64xn, const short *noise64xn, long n, float *irm)
{
    se_ctx *ctx = default_ctx("se_single_ibm_rows");
    if (!ctx) return SE_ENODEVICE;
    const int64_t frames = se_ibm_num_frames(n);
    if (frames <= 0) return SE_OK;
    const int64_t off[2] = {0, n}, fo[2] = {0, frames};
    return report("se_single_ibm_rows", se_ibm_rows_batch(ctx, pure64xn, noise64xn, off, 1, irm, nullptr, fo, SE_MEM_HOST));
}
