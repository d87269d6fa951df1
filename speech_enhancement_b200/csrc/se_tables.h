// Host-side constant tables of the hot path (built once per context, uploaded to the GPU).
// Each builder restates the reference's init code with the same float/double promotions so the
// tables are bit-identical to the reference's (checked by tests/test_abi_host.py::test_tables_bit_identical_to_oracle against the oracle).
#pragma once
#include <vector>

namespace se {

constexpr int kMel = 25;          // WF_MEL_ORDER, etsi/cpp/MelProcExports.h:19
constexpr int kFir = 17;          // NS_FILTER_LENGTH, etsi/cpp/NoiseSup.h:28
constexpr int kGtChannels = 64;   // NUMBER_CHANNEL, resyth_64sub_IBM/cpp/HuWang.h:13

struct NsTables {                 // one per NoiseSup profile
    int shift, frame, nfft, spec, ana, fs;
    std::vector<float> sig_window;            // Hanning, frame
    std::vector<float> ir_window;             // Hanning-17
    int bank_start[kMel], bank_len[kMel];
    int bank_w;                               // row width of bank (65 for P8 mel, 128 for P16 gamma)
    std::vector<float> bank;                  // [25][bank_w], zero padded
    std::vector<float> idct;                  // [25][25]
    float eps;                                // NS_EPS = (float)exp(-10.0)
};

struct CcTables {                 // CompCeps, 8 kHz
    float hamming[100];
    float dct[12 * 23];
    int start[23], len[23];
    float tri[23][129];
};

struct Hp16Tables {               // ETSI 16 kHz mode (etsi/cpp/16kHzProcessing.c)
    float dp[118];                            // half-band low-pass prototype; high-pass = reversed, alternating signs
    int fb_start[5], fb_len[5];
    float fb[5][65];                          // 5 un-normalised mel triangles from 80 Hz over the 65-bin PSD
    float dct26[12 * 26];                     // DCT over 23 + 3 log bands
};

struct GtTables {                 // 64-channel gammatone
    float cf[kGtChannels], bw[kGtChannels], mid[kGtChannels];
    float f1[kGtChannels], f2[kGtChannels], gain[kGtChannels];
    float hc[16];                             // hair-cell constants, see build_gt_tables
    double fall_d[160], rise_d[160];          // raised-cosine halves, double as in the reference
};

// SE_FLAG-less "exact" arithmetic mode (se_ctx_set_exact): the butterfly schedule of rfft(x, n, m)
// (etsi/cpp/rfft.c:45-180, split-radix after Sorensen et al. 1987) flattened into per-level lists of INDEPENDENT
// butterflies, so a warp can run one level in parallel and still produce every value with the reference's operations
// in the reference's order.  Level 0 = the length-two butterflies (op = first index); level k >= 1 = the L-shaped
// butterflies of size n2 = 2 << k (op = block base i | j << 16; j = 0 is the twiddle-free butterfly incl. its n2/8
// part, j >= 1 uses twiddle set j - 1 of that level: cc1, ss1, cc3, ss3 exactly as rfft.c:133-139 evaluates them --
// double cos/sin of a FLOAT angle in the C file, the float overloads in the plugin's C++ copy).
struct RfftPlan {
    int n, m;
    std::vector<unsigned> ops;
    int level_begin[16];          // first op of level k (k = 0..m-1), level_begin[m] = ops.size()
    int tw_begin[16];             // first twiddle set (4 floats each) of level k
    std::vector<float> tw;
};
void build_rfft_plan(RfftPlan &p, int n, int m, bool float_trig);

void build_ns_tables_p8(NsTables &t);
void build_ns_tables_p16(NsTables &t);
void build_cc_tables(CcTables &t);
void build_gt_tables(GtTables &t);
void build_hp16_tables(Hp16Tables &t);

}  // namespace se
