/* Test-infrastructure shim (ours, not reference code).
 * asdk::CWave is un-vendored.  The reference's sub-band code only uses it as a
 * (short*, length, fs) holder plus Write(path) (resyth_64sub_IBM/cpp/extractwav.cpp:28-29,123-125;
 * enhancement_extract_subband_linux/cpp/extractwav.cpp:88-90).  Write() hands the samples to a
 * sink callback instead of the file system so the oracle drivers can capture them in memory.   */
#ifndef SE_ORACLE_WAVE_SHIM_H
#define SE_ORACLE_WAVE_SHIM_H
#include <vector>
#include <string>
#include <cstring>
#include <iostream>
using namespace std;
/* enhancement_extract_subband_linux/cpp/HuWang.h:7 has its own PI commented out and relies on the
 * un-vendored asdk headers for it; the other variants define exactly this token sequence. */
#define PI (3.1415926535897932384626433832795)
extern "C" void se_shim_wave_sink(const char *path, const short *data, long n, int fs);
namespace asdk {
class CWave {
public:
    enum Channel { LEFT_CHANNEL = 0, RIGHT_CHANNEL = 1 };
    enum Type { PCM = 0 };
    CWave() : m_p(0), m_n(0), m_fs(16000) {}
    CWave(short *p, long n, int fs) : m_p(p), m_n(n), m_fs(fs) {}
    short *GetDataPtr() { return m_p; }
    long GetSampleNum() { return m_n; }
    int GetFs() { return m_fs; }
    int Read(const char *, Channel, bool, Type, int, int, int) { return -1; }
    int Write(const char *path) { se_shim_wave_sink(path, m_p, m_n, m_fs); return 0; }
private:
    short *m_p; long m_n; int m_fs;
};
}
#endif
