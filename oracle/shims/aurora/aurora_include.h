/* Test-infrastructure shim (ours, not reference code).
 * The reference's plugin-API NoiseSup (function/20141106_speech_enhancement/aurora_etsi/NoiseSupExports.h:5)
 * includes "aurora/aurora_include.h", an un-vendored header.  Only a handful of typedefs are used;
 * this shim supplies them so the reference sources compile in place from /root/reference.      */
#ifndef SE_ORACLE_AURORA_SHIM_H
#define SE_ORACLE_AURORA_SHIM_H
#include <stdint.h>
#include <stdio.h>
#include <math.h>
namespace aurora {
typedef void  *INSTANCE;
typedef void **PINSTANCE;
typedef int32_t int32s;
}
#ifndef X_FLOAT32
#define X_CHAR    char
#define X_INT16   short int
#define X_INT32   int
#define X_FLOAT32 float
#define X_FLOAT64 double
#endif
#ifndef PIx2
#define PIx2 6.28318530717958647692
#endif
#endif
