// Test driver for include/se_reference_api.hpp: a stand-in for the reference's drivers (resyth_64sub_IBM/cpp/main.cpp,
// enhancement_extract_subband_linux/cpp/main.cpp) with a minimal wave class in place of the un-vendored asdk::CWave
// and the cfg struct fields those drivers pass.  Usage:
//   dropin_driver <in.i16> <noise.i16> <mask.f32> <numFrame> <mode> <outdir/> [<clean.i16>]
// With <clean.i16>: also drives make_single_IBM (sub-band wavs of clean and noise -> <outdir/>irm.txt) and
// gammaToneFilter / hairCell with a by-value channel (-> <outdir/>gt.f32, <outdir/>hc.f32).
// writes <outdir/>sub/<id>_t_<chan>.wav, <outdir/>res/<id>_e_resynth.wav, <outdir/>mix/<id>_noisy.wav as raw int16 files.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace asdk {
class CWave {
public:
    enum Channel { LEFT_CHANNEL, RIGHT_CHANNEL };
    enum Type { PCM };
    CWave() : p_(nullptr), n_(0), fs_(16000) {}
    CWave(short *p, long n, int fs) : p_(p), n_(n), fs_(fs) {}
    // raw int16 files stand in for RIFF here (Write below writes them); signature as the drivers call it
    // (enhancement_extract_subband_linux/cpp/show_IBM.cpp:95-107)
    int Read(const char *path, Channel, bool, Type, int nFs, int, int)
    {
        FILE *f = std::fopen(path, "rb");
        if (!f) return -1;
        std::fseek(f, 0, SEEK_END);
        own_.resize((size_t)std::ftell(f) / sizeof(short));
        std::fseek(f, 0, SEEK_SET);
        const size_t got = std::fread(own_.data(), sizeof(short), own_.size(), f);
        std::fclose(f);
        own_.resize(got);
        p_ = own_.data(); n_ = (long)own_.size(); fs_ = nFs;
        return 0;
    }
    short *GetDataPtr() { return p_; }
    long GetSampleNum() { return n_; }
    int GetFs() { return fs_; }
    int Write(const char *path)
    {
        FILE *f = std::fopen(path, "wb");
        if (!f) return -1;
        std::fwrite(p_, sizeof(short), (size_t)n_, f);
        std::fclose(f);
        return 0;
    }
private:
    short *p_; long n_; int fs_;
    std::vector<short> own_;
};
}  // namespace asdk

#define FILE_LEN 255
typedef struct _FILESNAME {
    char outputDictionary[FILE_LEN];
    char save_noisy_dir[FILE_LEN];
    char save_resynth_e_dir[FILE_LEN];
    int addnoisedB;
    char purewavDictionary[FILE_LEN];
    char save_subband_pure_wav_dir[FILE_LEN];
    char save_subband_noise_wav_dir[FILE_LEN];
} FILESNAME;

struct gammaTone {              // HuWang.h:20-26
    float cf, bw;
    float midEarCoeff;
    float p[4];
    float q[4];
};

#define SE_DROPIN_WAVE asdk::CWave
#define SE_DROPIN_HAVE_RESYNTH_OPTS
#define SE_DROPIN_HAVE_ADDNOISE_OPTS
#define SE_DROPIN_HAVE_IBM_OPTS
#define SE_DROPIN_HAVE_GAMMATONE
#include "se_reference_api.hpp"

template <class T> static std::vector<T> slurp(const char *path)
{
    std::vector<T> v;
    FILE *f = std::fopen(path, "rb");
    if (!f) { std::perror(path); std::exit(2); }
    std::fseek(f, 0, SEEK_END);
    v.resize((size_t)std::ftell(f) / sizeof(T));
    std::fseek(f, 0, SEEK_SET);
    if (std::fread(v.data(), sizeof(T), v.size(), f) != v.size()) std::exit(2);
    std::fclose(f);
    return v;
}

int main(int argc, char **argv)
{
    if (argc != 7 && argc != 8) return 2;
    std::vector<short> x = slurp<short>(argv[1]), nz = slurp<short>(argv[2]);
    std::vector<float> m = slurp<float>(argv[3]);
    const int num_frame = std::atoi(argv[4]), mode = std::atoi(argv[5]);
    FILESNAME opts;
    std::memset(&opts, 0, sizeof opts);
    std::snprintf(opts.outputDictionary, FILE_LEN, "%s", argv[6]);
    std::snprintf(opts.save_noisy_dir, FILE_LEN, "mix/");
    std::snprintf(opts.save_resynth_e_dir, FILE_LEN, "res/");
    opts.addnoisedB = 5;
    char id[] = "utt1", tail[] = "_t", path[600];
    std::snprintf(path, sizeof path, "%ssub/", argv[6]);

    asdk::CWave wav(x.data(), (long)x.size(), 16000);
    subbband(&wav, path, tail, id);                                   // reference name and signature

    std::vector<std::vector<float> > v_erm((size_t)num_frame, std::vector<float>(64));
    for (int f = 0; f < num_frame; f++)
        for (int c = 0; c < 64; c++) v_erm[f][c] = m[(size_t)f * 64 + c];
    if (mode == 0) resynth(&wav, opts, v_erm, id);                    // reference name and signature (IBM variant)
    else se_dropin::resynth(&wav, opts, v_erm, id, mode);
    std::snprintf(path, sizeof path, "%sres/mask_after.f32", argv[6]);
    FILE *f = std::fopen(path, "wb");
    for (int r = 0; r < num_frame; r++) std::fwrite(v_erm[r].data(), sizeof(float), 64, f);
    std::fclose(f);

    std::vector<short> noisy(x.size());
    asdk::CWave nw(nz.data(), (long)nz.size(), 16000), yw(noisy.data(), (long)noisy.size(), 16000);
    addnoise(&wav, &nw, &yw, id, opts);
    std::snprintf(path, sizeof path, "%smix/noise_scaled.i16", argv[6]);
    f = std::fopen(path, "wb");
    std::fwrite(nz.data(), sizeof(short), nz.size(), f);
    std::fclose(f);
    if (argc == 8) {
        // the training-data flow of enhancement_extract_subband_linux/cpp/main.cpp:236-255: subbband(pure),
        // subbband(noise) -> 2 x 64 wavs -> make_single_IBM appends the utterance's IRM matrix to one text file
        std::vector<short> clean = slurp<short>(argv[7]);
        std::vector<short> nraw = slurp<short>(argv[2]);
        std::snprintf(opts.purewavDictionary, FILE_LEN, "%spure/", argv[6]);
        std::snprintf(opts.save_subband_pure_wav_dir, FILE_LEN, "subp/");
        std::snprintf(opts.save_subband_noise_wav_dir, FILE_LEN, "subn/");
        asdk::CWave cw(clean.data(), (long)clean.size(), 16000), nw2(nraw.data(), (long)nraw.size(), 16000);
        std::snprintf(path, sizeof path, "%spure/%s.wav", argv[6], id);
        cw.Write(path);
        char none[] = "", ntail[] = "_noise";
        std::snprintf(path, sizeof path, "%ssubp/", argv[6]);
        subbband(&cw, path, none, id);
        std::snprintf(path, sizeof path, "%ssubn/", argv[6]);
        subbband(&nw2, path, ntail, id);
        std::snprintf(path, sizeof path, "%sirm.txt", argv[6]);
        FILE *fi = std::fopen(path, "at+");                           // main.cpp:253-255 opens it "at+" per utterance
        make_single_IBM(id, fi, opts);                                // reference name and signature (make_IBM.h:9)
        std::fclose(fi);
        // HuWang.h:49,62 by name: a channel that is not one of the 64 presets
        gammaTone ch;
        std::memset(&ch, 0, sizeof ch);
        ch.cf = 1234.5f; ch.bw = 150.0f; ch.midEarCoeff = 0.7f;
        ch.p[2] = 99.0f;                                              // state is zeroed on entry (extractwav.cpp:185-189)
        std::vector<float> in(x.size()), gt(x.size()), hc(x.size());
        for (size_t i = 0; i < x.size(); i++) in[i] = (float)x[i];
        gammaToneFilter(in.data(), gt.data(), ch, (long)in.size());
        hairCell(gt.data(), hc.data(), (long)gt.size());
        std::snprintf(path, sizeof path, "%sgt.f32", argv[6]);
        f = std::fopen(path, "wb");
        std::fwrite(gt.data(), sizeof(float), gt.size(), f);
        std::fclose(f);
        std::snprintf(path, sizeof path, "%shc.f32", argv[6]);
        f = std::fopen(path, "wb");
        std::fwrite(hc.data(), sizeof(float), hc.size(), f);
        std::fclose(f);
    }
    return 0;
}
