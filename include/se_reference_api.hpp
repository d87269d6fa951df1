/* se_reference_api.hpp — the reference's C++ entry points that take asdk::CWave, as header-only adapters over the C
 * twins of se_reference_api.h (libse_b200_dropin.so).  asdk is not vendored in the reference tree, so
 * this header does not include it: the driver includes its own "Wave.h" and "struct.h" FIRST, as the reference's
 * extractwav.h does, then this file.  Any wave class with
 *     short *GetDataPtr();  long GetSampleNum();  int GetFs();  Write(const char *path);  Wave(short *, long, int fs)
 * works (usage: enhancement_extract_subband_linux/cpp/extractwav.cpp:8-11,33,88-90;
 * resyth_64sub_IBM/cpp/extractwav.cpp:28-29,123-125).
 *
 *   addnoise   enhancement_extract_subband_linux/cpp/extractwav.cpp:6-35     (extractwav.h:5)
 *   subbband   enhancement_extract_subband_linux/cpp/extractwav.cpp:36-97    (extractwav.h:6)
 *   resynth    resyth_64sub_IBM/cpp/extractwav.cpp:9-132                     (extractwav.h:4)
 *   make_single_IBM  enhancement_extract_subband_linux/cpp/show_IBM.cpp:92-213 (make_IBM.h:9)
 *   gammaToneFilter / hairCell  <tool>/cpp/extractwav.cpp:168-258            (HuWang.h:49,62)
 *              resyth_64sub_ori/cpp/extractwav.cpp, 1dnn_resynth/extractwav.cpp: same signature, other `mode`
 *
 * The templates live in namespace se_dropin.  Define SE_DROPIN_WAVE (e.g. asdk::CWave) and, for resynth/addnoise,
 * have FILESNAME declared, to also get the reference's exact non-template names:
 *     #include "Wave.h"
 *     #include "struct.h"
 *     #define SE_DROPIN_WAVE asdk::CWave
 *     #define SE_DROPIN_RESYNTH_MODE 0            // 0 resyth_64sub_IBM, 1 resyth_64sub_ori, 2 1dnn_resynth
 *     #include "se_reference_api.hpp"
 * File names and directory fields are the reference's.  All arithmetic runs on the GPU through the C twins; a
 * failure (no GPU, bad arguments) prints the reason on stderr and writes nothing, there is no CPU path. */
#ifndef SE_REFERENCE_API_HPP
#define SE_REFERENCE_API_HPP
#include <cstdio>
#include <vector>

#include "se_reference_api.h"

namespace se_dropin {

/* addnoise(): scales the noise IN PLACE, writes the mixture into noisywav's buffer and to
 * <outputDictionary><save_noisy_dir><id>_noisy.wav (extractwav.cpp:31-33). */
template <class Wave, class Opts>
inline void addnoise(Wave *purewav, Wave *temp_noisewav, Wave *noisywav, char *name_id, Opts opts)
{
    const long len = purewav->GetSampleNum();
    if (se_addnoise(purewav->GetDataPtr(), temp_noisewav->GetDataPtr(), noisywav->GetDataPtr(), len, opts.addnoisedB) != 0)
        return;
    char noisy_name[1024];
    std::snprintf(noisy_name, sizeof noisy_name, "%s%s%s_noisy.wav", opts.outputDictionary, opts.save_noisy_dir, name_id);
    noisywav->Write(noisy_name);
}

/* subbband(): 64 channel wavs <path><id><addtail>_<chan>.wav (extractwav.cpp:86-90). */
template <class Wave>
inline void subbband(Wave *wav, char *path, char *addtail, char *name_id)
{
    const long n = wav->GetSampleNum();
    std::vector<short> out((size_t)64 * (size_t)(n > 0 ? n : 0));
    if (se_subband(wav->GetDataPtr(), n, out.data()) != 0) return;
    char temp[1024];
    for (int chan = 0; chan < 64; chan++) {
        Wave chan_wav(out.data() + (size_t)chan * (size_t)n, n, wav->GetFs());
        std::snprintf(temp, sizeof temp, "%s%s%s_%d.wav", path, name_id, addtail, chan);
        chan_wav.Write(temp);
    }
}

/* resynth(): v_erm holds numFrame rows of 64 mask values; numFrame follows the variant ((n-320)/160+1, or n/160 for
 * 1dnn_resynth; the caller sized v_erm with the same formula, main.cpp:110).  In IBM mode entries > 0.5 come back
 * as 1.0 like the reference leaves them (extractwav.cpp:99).  Writes
 * <outputDictionary><save_resynth_e_dir><id>_e_resynth.wav (:123-125). */
template <class Wave, class Opts>
inline void resynth(Wave *wav, Opts opts, std::vector<std::vector<float> > &v_erm, char *name_id, int mode)
{
    const long n = wav->GetSampleNum();
    const int num_frame = (int)v_erm.size();
    std::vector<float> m((size_t)num_frame * 64);
    for (int f = 0; f < num_frame; f++)
        for (int c = 0; c < 64; c++) m[(size_t)f * 64 + c] = v_erm[f][c];
    std::vector<short> out((size_t)(n > 0 ? n : 0));
    if (se_resynth(wav->GetDataPtr(), n, m.data(), num_frame, mode, out.data()) != 0) return;
    for (int f = 0; f < num_frame; f++)
        for (int c = 0; c < 64; c++) v_erm[f][c] = m[(size_t)f * 64 + c];
    Wave resynth_wav_e(out.data(), n, wav->GetFs());
    char temp[1024];
    std::snprintf(temp, sizeof temp, "%s%s%s_e_resynth.wav", opts.outputDictionary, opts.save_resynth_e_dir, name_id);
    resynth_wav_e.Write(temp);
}

/* gammaToneFilter(): the channel is passed BY VALUE (cf, bw, midEarCoeff; its p / q state is zeroed on entry,
 * extractwav.cpp:168-189), so any parameters are legal, not only the 64 presets.  GammaTone is the driver's own
 * `struct gammaTone` (HuWang.h:20-26). */
template <class GammaTone>
inline void gammaToneFilter(float *input, float *output, GammaTone fChan, long sigLength)
{
    se_gammatone_filter_params(input, output, fChan.cf, fChan.bw, fChan.midEarCoeff, sigLength);
}

/* make_single_IBM(): reads the clean wav for its length and the 2 x 64 sub-band wavs subbband() wrote
 * (<out><save_subband_pure_wav_dir><id>_<c>.wav, <out><save_subband_noise_wav_dir><id>_noise_<c>.wav), computes
 * IRM = Sp / (Sp + Sn) per (frame, channel) on the GPU and appends the "<id>_noisy [ ... ]" text matrix to fp_IRM_W
 * (show_IBM.cpp:92-213).  Wave needs a default constructor and asdk::CWave's Read(path, channel, bHead, type, nFs,
 * nChannel, nMaxSample) with its LEFT_CHANNEL / PCM enumerators (:95-107).  Frame powers sit on OUR restatement of
 * the un-vendored asdk::SpecInfo (parity unpinned at that boundary). */
template <class Wave, class Opts>
inline void make_single_IBM(char *name_id, FILE *fp_IRM_W, Opts opts)
{
    char temp[1024];
    Wave speech;
    std::snprintf(temp, sizeof temp, "%s%s.wav", opts.purewavDictionary, name_id);
    speech.Read(temp, Wave::LEFT_CHANNEL, true, Wave::PCM, 16000, 1, -1);
    const long n = speech.GetSampleNum();
    const long frames = n >= 320 ? (n - 320) / 160 + 1 : 0;                 /* buffNum, :110 */
    if (frames <= 0 || !fp_IRM_W) return;
    std::vector<short> pure((size_t)64 * (size_t)n, 0), noise((size_t)64 * (size_t)n, 0);
    for (int c = 0; c < 64; c++) {
        Wave p, z;
        std::snprintf(temp, sizeof temp, "%s%s%s_%d.wav", opts.outputDictionary, opts.save_subband_pure_wav_dir, name_id, c);
        p.Read(temp, Wave::LEFT_CHANNEL, true, Wave::PCM, 16000, 1, -1);
        std::snprintf(temp, sizeof temp, "%s%s%s_noise_%d.wav", opts.outputDictionary, opts.save_subband_noise_wav_dir, name_id, c);
        z.Read(temp, Wave::LEFT_CHANNEL, true, Wave::PCM, 16000, 1, -1);
        const long np = p.GetSampleNum() < n ? p.GetSampleNum() : n, nz = z.GetSampleNum() < n ? z.GetSampleNum() : n;
        for (long i = 0; i < np; i++) pure[(size_t)c * (size_t)n + (size_t)i] = p.GetDataPtr()[i];
        for (long i = 0; i < nz; i++) noise[(size_t)c * (size_t)n + (size_t)i] = z.GetDataPtr()[i];
    }
    std::vector<float> irm((size_t)frames * 64);
    if (se_single_ibm_rows(pure.data(), noise.data(), n, irm.data()) != 0) return;
    std::fprintf(fp_IRM_W, "%s_noisy [\n", name_id);                         /* :194-209 */
    for (long f = 0; f < frames - 1; f++) {
        for (int c = 0; c < 64; c++) std::fprintf(fp_IRM_W, "%.7f ", irm[(size_t)f * 64 + c]);
        std::fprintf(fp_IRM_W, "\n ");
    }
    for (int c = 0; c < 64; c++) std::fprintf(fp_IRM_W, "%.7f ", irm[(size_t)(frames - 1) * 64 + c]);
    std::fprintf(fp_IRM_W, "]\n");
}

}  // namespace se_dropin

#ifdef SE_DROPIN_WAVE
#ifndef SE_DROPIN_RESYNTH_MODE
#define SE_DROPIN_RESYNTH_MODE 0
#endif
/* the reference's own names and signatures */
inline void subbband(SE_DROPIN_WAVE *wav, char *path, char *addhead, char *name_id)
{
    se_dropin::subbband(wav, path, addhead, name_id);
}
#ifdef SE_DROPIN_HAVE_RESYNTH_OPTS       /* FILESNAME of resyth_64sub_IBM/cpp/struct.h (save_resynth_e_dir) */
inline void resynth(SE_DROPIN_WAVE *wav, FILESNAME opts, std::vector<std::vector<float> > &v_erm, char *name_id)
{
    se_dropin::resynth(wav, opts, v_erm, name_id, SE_DROPIN_RESYNTH_MODE);
}
#endif
#ifdef SE_DROPIN_HAVE_ADDNOISE_OPTS      /* FILESNAME of enhancement_extract_subband_linux/cpp/struct.h (addnoisedB) */
inline void addnoise(SE_DROPIN_WAVE *purewav, SE_DROPIN_WAVE *temp_noisewav, SE_DROPIN_WAVE *noisywav, char *name_id,
                     FILESNAME opts)
{
    se_dropin::addnoise(purewav, temp_noisewav, noisywav, name_id, opts);
}
#endif
#ifdef SE_DROPIN_HAVE_IBM_OPTS           /* FILESNAME of enhancement_extract_subband_linux/cpp/struct.h (make_IBM.h:9) */
inline void make_single_IBM(char *name_id, FILE *fp_IRM_W, FILESNAME opts)
{
    se_dropin::make_single_IBM<SE_DROPIN_WAVE>(name_id, fp_IRM_W, opts);
}
#endif
#endif /* SE_DROPIN_WAVE */

#ifdef SE_DROPIN_HAVE_GAMMATONE          /* `struct gammaTone` of HuWang.h:20-26 is declared: HuWang.h:49,62 by name */
inline void gammaToneFilter(float *input, float *output, gammaTone fChan, long sigLength)
{
    se_dropin::gammaToneFilter(input, output, fChan, sigLength);
}
inline void hairCell(float *input, float *output, long sigLength) { se_hair_cell(input, output, sigLength); }
#endif

#endif
