// Host-side file formats around the hot path (include/se_io.h).  No GPU work here.
#include "se_io.h"
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#define SE_EXPORT extern "C" __attribute__((visibility("default")))

namespace {

bool read_exact(FILE *fp, void *p, size_t n) { return fread(p, 1, n, fp) == n; }
uint32_t le32(const unsigned char *b) { return (uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24); }
uint16_t le16(const unsigned char *b) { return (uint16_t)(b[0] | (b[1] << 8)); }
void put32(unsigned char *b, uint32_t v) { b[0] = v & 255; b[1] = (v >> 8) & 255; b[2] = (v >> 16) & 255; b[3] = (v >> 24) & 255; }
void put16(unsigned char *b, uint16_t v) { b[0] = v & 255; b[1] = (v >> 8) & 255; }

}  // namespace

SE_EXPORT int se_wav_read(const char *path, int channel, short **data, long *n_samples, int *fs)
{
    if (!path || !data || !n_samples) return -1;
    *data = nullptr;
    *n_samples = 0;
    FILE *fp = fopen(path, "rb");
    if (!fp) return -2;
    unsigned char hdr[12];
    if (!read_exact(fp, hdr, 12) || memcmp(hdr, "RIFF", 4) || memcmp(hdr + 8, "WAVE", 4)) { fclose(fp); return -3; }
    int n_chan = 0, bits = 0, rate = 0, fmt = 0;
    bool have_fmt = false;
    for (;;) {
        unsigned char ch[8];
        if (!read_exact(fp, ch, 8)) { fclose(fp); return -3; }
        const uint32_t sz = le32(ch + 4);
        if (!memcmp(ch, "fmt ", 4)) {
            unsigned char f[16];
            if (sz < 16 || !read_exact(fp, f, 16)) { fclose(fp); return -3; }
            fmt = le16(f); n_chan = le16(f + 2); rate = (int)le32(f + 4); bits = le16(f + 14);
            have_fmt = true;
            if (sz > 16) fseek(fp, (long)(sz - 16 + (sz & 1)), SEEK_CUR);
        } else if (!memcmp(ch, "data", 4)) {
            if (!have_fmt || (fmt != 1 && fmt != 0xFFFE) || bits != 16 || n_chan < 1 || channel < 0 || channel >= n_chan) {
                fclose(fp);
                return -4;      // only PCM16 is what the drivers read (CWave::PCM)
            }
            // tolerate a wrong/streamed size field: read what is there
            // (a streamed 0xFFFFFFFF size must not allocate 4 GB: clamp to what is left in the file)
            const long here = ftell(fp);
            long left = -1;
            if (here >= 0 && fseek(fp, 0, SEEK_END) == 0) { left = ftell(fp) - here; fseek(fp, here, SEEK_SET); }
            const size_t want = (left >= 0 && (unsigned long)left < (unsigned long)sz) ? (size_t)left : (size_t)sz;
            std::vector<unsigned char> raw;
            raw.resize(want);
            const size_t got = fread(raw.data(), 1, want, fp);
            fclose(fp);
            const long frames = (long)(got / (2u * (unsigned)n_chan));
            short *out = (short *)malloc(sizeof(short) * (size_t)(frames > 0 ? frames : 1));
            if (!out) return -5;
            for (long i = 0; i < frames; i++) out[i] = (short)le16(&raw[(size_t)(i * n_chan + channel) * 2]);
            *data = out;
            *n_samples = frames;
            if (fs) *fs = rate;
            return 0;
        } else {
            if (fseek(fp, (long)(sz + (sz & 1)), SEEK_CUR)) { fclose(fp); return -3; }
        }
    }
}

SE_EXPORT int se_wav_write(const char *path, const short *data, long n_samples, int fs)
{
    if (!path || (!data && n_samples > 0) || n_samples < 0) return -1;
    FILE *fp = fopen(path, "wb");
    if (!fp) return -2;
    unsigned char h[44];
    const uint32_t bytes = (uint32_t)n_samples * 2u;
    memcpy(h, "RIFF", 4); put32(h + 4, 36 + bytes); memcpy(h + 8, "WAVEfmt ", 8); put32(h + 16, 16);
    put16(h + 20, 1); put16(h + 22, 1); put32(h + 24, (uint32_t)fs); put32(h + 28, (uint32_t)fs * 2u);
    put16(h + 32, 2); put16(h + 34, 16); memcpy(h + 36, "data", 4); put32(h + 40, bytes);
    bool ok = fwrite(h, 1, 44, fp) == 44;
    std::vector<unsigned char> raw((size_t)bytes);
    for (long i = 0; i < n_samples; i++) put16(&raw[(size_t)i * 2], (uint16_t)data[i]);
    ok = ok && fwrite(raw.data(), 1, raw.size(), fp) == raw.size();
    ok = (fclose(fp) == 0) && ok;
    return ok ? 0 : -6;
}

SE_EXPORT void se_io_free(void *p) { free(p); }

// ---- Kaldi-style text matrices ---------------------------------------------------------------------
struct se_mask_reader {
    FILE *fp;
    char buf[64 * 15];          // the reference's line buffer (resyth_64sub_IBM/cpp/main.cpp:84)
    bool pending_header;        // buf holds a header line that rows() ran into
};

SE_EXPORT se_mask_reader *se_mask_reader_open(const char *path)
{
    if (!path) return nullptr;
    FILE *fp = fopen(path, "r");
    if (!fp) return nullptr;
    se_mask_reader *r = new se_mask_reader();
    r->fp = fp;
    r->pending_header = false;
    r->buf[0] = 0;
    return r;
}

SE_EXPORT int se_mask_reader_next(se_mask_reader *r, char *name, int name_cap)
{
    if (!r) return 0;
    for (;;) {
        if (!r->pending_header && !fgets(r->buf, sizeof r->buf, r->fp)) return 0;
        r->pending_header = false;
        const char *br = strstr(r->buf, "[");
        if (!br) continue;                      // a data line outside any matrix: the reference would misbehave; skip
        if (name && name_cap > 0) {
            std::string s(r->buf, (size_t)(br - r->buf));
            const size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
            s = (a == std::string::npos) ? std::string() : s.substr(a, b - a + 1);
            strncpy(name, s.c_str(), (size_t)name_cap - 1);
            name[name_cap - 1] = 0;
        }
        return 1;
    }
}

SE_EXPORT int se_mask_reader_rows(se_mask_reader *r, int num_frame, float *mask)
{
    if (!r || !mask || num_frame <= 0) return 0;
    int rows = 0;
    while (rows < num_frame && fgets(r->buf, sizeof r->buf, r->fp)) {
        if (strstr(r->buf, "[")) { r->pending_header = true; break; }
        std::stringstream ss{std::string(r->buf)};
        float *row = mask + (size_t)rows * 64;
        for (int j = 0; j < 64; j++) {
            float v = 0.0f;
            if (!(ss >> v)) v = 0.0f;           // operator>> on a failed stream leaves the zero-initialised row value
            row[j] = v;
        }
        rows++;
    }
    return rows;
}

SE_EXPORT void se_mask_reader_close(se_mask_reader *r)
{
    if (!r) return;
    if (r->fp) fclose(r->fp);
    delete r;
}

SE_EXPORT int se_write_kaldi_matrix(FILE *fp, const char *id, const float *m, int rows, int cols)
{
    if (!fp || !id || !m || rows < 1 || cols < 1) return -1;
    fprintf(fp, "%s_noisy [\n", id);
    for (int f = 0; f < rows - 1; f++) {
        for (int c = 0; c < cols; c++) fprintf(fp, "%.7f ", m[(size_t)f * cols + c]);
        fprintf(fp, "\n ");
    }
    for (int c = 0; c < cols; c++) fprintf(fp, "%.7f ", m[(size_t)(rows - 1) * cols + c]);
    fprintf(fp, "]\n");
    return ferror(fp) ? -6 : 0;
}

SE_EXPORT int se_write_wiener(FILE *fp, const float *gains, int rows)
{
    if (!fp || (!gains && rows > 0)) return -1;
    for (int r = 0; r < rows; r++) {
        for (int i = 0; i < 25; i++) fprintf(fp, "%f ", gains[(size_t)r * 25 + i]);
        fprintf(fp, "\n");
    }
    return ferror(fp) ? -6 : 0;
}

SE_EXPORT int se_write_sibm(FILE *fp, const char *id, int subband, const unsigned char *ibm, int frames)
{
    if (!fp || !id || !ibm || subband < 0 || subband >= 64) return -1;
    fprintf(fp, "%s_noisy_%d ", id, subband);
    for (int f = 0; f < frames; f++) fprintf(fp, "%d ", (int)ibm[(size_t)f * 64 + subband]);
    fprintf(fp, "\n");
    return ferror(fp) ? -6 : 0;
}

SE_EXPORT int se_write_frames_f(FILE *fp, const float *m, int frames, int n_chan)
{
    if (!fp || (!m && frames > 0) || n_chan < 1) return -1;
    for (int f = 0; f < frames; f++) {
        for (int c = 0; c < n_chan; c++) fprintf(fp, "%f ", m[(size_t)f * n_chan + c]);
        fprintf(fp, "\n");
    }
    return ferror(fp) ? -6 : 0;
}
