// Test driver for include/se_reference_api.hpp: a stand-in for the reference's drivers (resyth_64sub_IBM/cpp/main.cpp,
// enhancement_extract_subband_linux/cpp/main.cpp) with a minimal wave class in place of the un-vendored asdk::CWave
// and the cfg struct fields those drivers pass.  Usage:
//   dropin_driver <in.i16> <noise.i16> <mask.f32> <numFrame> <mode> <outdir/>
// writes <outdir/>sub/<id>_t_<chan>.wav, <outdir/>res/<id>_e_resynth.wav, <outdir/>mix/<id>_noisy.wav as raw int16 files.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace asdk {
class CWave {
public:
    CWave(short *p, long n, int fs) : p_(p), n_(n), fs_(fs) {}
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
};
}  // namespace asdk

#define FILE_LEN 255
typedef struct _FILESNAME {
    char outputDictionary[FILE_LEN];
    char save_noisy_dir[FILE_LEN];
    char save_resynth_e_dir[FILE_LEN];
    int addnoisedB;
} FILESNAME;

#define SE_DROPIN_WAVE asdk::CWave
#define SE_DROPIN_HAVE_RESYNTH_OPTS
#define SE_DROPIN_HAVE_ADDNOISE_OPTS
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
    if (argc != 7) return 2;
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
    return 0;
}
