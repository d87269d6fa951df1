/* se_io.h — host-side file formats either side of the hot path (SURVEY.md §8a row a24, §8b "on-disk
 * layouts", §8f rank 3): RIFF PCM16 wav, the Kaldi-style text mask matrices the resynthesis drivers
 * parse, and the text feature files the extraction drivers write.  Plain C, no GPU work, exported by
 * libse_b200_dropin.so.  These replace what the reference does through the un-vendored asdk::CWave and
 * inline fprintf/stringstream code in its main.cpp files, so its drivers keep their file layouts.
 * All functions return 0 on success (negative on error) unless stated otherwise; nothing exits.
 */
#ifndef SE_IO_H
#define SE_IO_H
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- RIFF/WAVE PCM16 (what asdk::CWave::Read / Write do for the drivers: etsi/cpp/main.cpp:48-64,
 *      resyth_64sub_IBM/cpp/extractwav.cpp:123-125).  Multi-channel files: `channel` selects one
 *      (0 = left, the drivers' CWave::LEFT_CHANNEL).  *data is malloc'd; free with se_io_free(). */
int se_wav_read(const char *path, int channel, short **data, long *n_samples, int *fs);
int se_wav_write(const char *path, const short *data, long n_samples, int fs);
void se_io_free(void *p);

/* ---- mask matrices: "<name> [" header line, then rows of 64 floats, "]" after the last value
 *      (parser: resyth_64sub_IBM/cpp/main.cpp:84-146 and 1dnn_resynth/main.cpp; writer:
 *      enhancement_extract_subband_linux/cpp/show_IBM.cpp:194-209).
 * The reference reads lines through a 64*15-char buffer, treats ANY line containing '[' as a header,
 * parses every other line as one row with operator>> (missing values read as 0) and takes the row
 * count from the wav length, not from the file; the reader reproduces exactly that. */
typedef struct se_mask_reader se_mask_reader;
se_mask_reader *se_mask_reader_open(const char *path);                 /* NULL on failure            */
/* Skips to the next header line.  Returns 1 (name = text before '[', trimmed) or 0 at end of file. */
int se_mask_reader_next(se_mask_reader *r, char *name, int name_cap);
/* Reads up to num_frame rows (64 floats each) after the current header; stops early at the next
 * header or EOF.  Returns the number of rows read. */
int se_mask_reader_rows(se_mask_reader *r, int num_frame, float *mask);
void se_mask_reader_close(se_mask_reader *r);
/* "<id>_noisy [\n" + rows of cols "%.7f " separated by "\n " + "]\n" (show_IBM.cpp:194-209).
 * m is [rows][cols] (frame-major, as se_ibm_batch / se_single_ibm produce it). */
int se_write_kaldi_matrix(FILE *fp, const char *id, const float *m, int rows, int cols);

/* ---- feature text files -------------------------------------------------------------------------
 * .wiener: one line of 25 "%f " per stage-2 frame (aurora_etsi/NoiseSup.cpp:1318-1326).           */
int se_write_wiener(FILE *fp, const float *gains /*[rows][25]*/, int rows);
/* .sIBM: "<id>_noisy_<subband> " + "%d " per frame + "\n" (function/.../show_IBM/show_IBM.cpp:150-162);
 * ibm is [frames][64] as se_single_ibm produces it, `subband` picks the column. */
int se_write_sibm(FILE *fp, const char *id, int subband, const unsigned char *ibm, int frames);
/* .IBM / .IRM: "%f " x n_chan per frame + "\n" (show_IBM.cpp:70-84). */
int se_write_frames_f(FILE *fp, const float *m /*[frames][n_chan]*/, int frames, int n_chan);

#ifdef __cplusplus
}
#endif
#endif
