/* TEST / BENCH INFRASTRUCTURE — not part of the product.
 *
 * Times the reference's OWN CPU implementation of the hot path (oracle/_ref/libref_*_fast.so = the reference's sources
 * compiled in place, or — where that is absent — the plain-C port oracle/libse_oracle_fast.so) over a batch of
 * utterances that already sits in memory: no Python, no pickling, no file I/O inside the timed region.
 *
 * Worker model = the reference's own (one work item = one utterance, N workers pulling indices,
 * function/.../aurora_speech_enhancement/aurora_speech_enhancement.cpp:110-120), as N forked PROCESSES rather than
 * threads: the oracle/_ref drivers capture the "wav files" the reference writes in a global sink and silence its
 * stdout chatter by redirecting fd 1 (oracle/ref_g64.cpp, ref_p16.cpp), neither of which is thread-safe.  The library
 * is dlopen()ed in the CALLING process (so it shows up among the process's loaded objects) and inherited by the
 * workers; input arrays are shared copy-on-write; the work counter and the per-worker results live in a MAP_SHARED page.
 *
 *   workload 0  P8   etsi_denoise + the CompCeps step   ref_etsi_trace / so_etsi_denoise
 *                    (etsi/cpp/AdvFrontEnd.c:125-210, ParmInterface.c:275-293)
 *   workload 1  P16  etsi_denoise_mapping_func_Wiener on a fresh thread instance per utterance
 *                    (aurora_etsi/NoiseSup.cpp:913-1407; ref_p16_* / so_ns_* + so_p16_func_wiener)
 *   workload 2  G64  resynth, IBM variant (resyth_64sub_IBM/cpp/extractwav.cpp:9-132; ref_resynth / so_resynth)
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

typedef struct {
    volatile int next;            /* work counter */
    int pad[15];
    double checksum[256];         /* per worker */
    long units[256];              /* per worker: utterances done */
} shared_t;

typedef long (*p8_ref_fn)(const short *, long, short *, float *, long *, void *);
typedef long (*p8_port_fn)(const int16_t *, long, int16_t *, float *, float *, long *, void *);
typedef void *(*p16_create_fn)(void);
typedef void (*p16_destroy_fn)(void *);
typedef int (*p16_ref_fn)(void *, float *, int, float *, int *, int *, int *, int *, int *, float *, int *, char *, long);
typedef void *(*ns_create_fn)(int);
typedef int (*p16_port_fn)(void *, const float *, int, float *, int *, int *, int *, int *, int *, float *);
typedef int (*rs_ref_fn)(short *, long, float *, int, short *);
typedef void (*rs_port_fn)(const int16_t *, long, float *, int, int, int16_t *, float *);

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static void *g_lib = NULL;
static char g_lib_path[1024] = "";

/* dlopen in the caller's process; returns 0 on success. */
int ref_bench_open(const char *path)
{
    if (g_lib && strcmp(path, g_lib_path) == 0) return 0;
    if (g_lib) dlclose(g_lib);
    g_lib = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!g_lib) { fprintf(stderr, "ref_bench_open: %s\n", dlerror()); return -1; }
    snprintf(g_lib_path, sizeof g_lib_path, "%s", path);
    return 0;
}

static double work_p8(int port, const int16_t *x, long n)
{
    short *den = (short *)calloc((size_t)(n > 0 ? n : 1), sizeof(short));
    float *ceps = (float *)calloc((size_t)(n / 80 + 1) * 14, sizeof(float));
    long nc = 0, i;
    double cs = 0.0;
    if (port) ((p8_port_fn)dlsym(g_lib, "so_etsi_denoise"))(x, n, den, NULL, ceps, &nc, NULL);
    else ((p8_ref_fn)dlsym(g_lib, "ref_etsi_trace"))(x, n, den, ceps, &nc, NULL);
    for (i = 0; i < n; i++) cs += den[i];
    for (i = 0; i < nc * 14; i++) cs += ceps[i];
    free(den);
    free(ceps);
    return cs;
}

static double work_p16(int port, const int16_t *x, long n)
{
    const long body = n - n % 160, nb = body / 160;
    float *in = (float *)malloc(sizeof(float) * (size_t)(body > 0 ? body : 1));
    float *out = (float *)calloc((size_t)(body > 0 ? body : 1), sizeof(float));
    int *fl = (int *)calloc((size_t)(nb > 0 ? nb : 1) * 5, sizeof(int));
    float *gains = (float *)calloc((size_t)(nb > 0 ? nb : 1) * 25, sizeof(float));
    long i;
    int rows = 0;
    double cs = 0.0;
    for (i = 0; i < body; i++) in[i] = (float)x[i];          /* aurora_etsi_test.cpp:100-107 */
    if (port) {
        void *h = ((ns_create_fn)dlsym(g_lib, "so_ns_create"))(16);
        rows = ((p16_port_fn)dlsym(g_lib, "so_p16_func_wiener"))(h, in, (int)body, out, fl, fl + nb, fl + 2 * nb, fl + 3 * nb,
                                                               fl + 4 * nb, gains);
        ((p16_destroy_fn)dlsym(g_lib, "so_ns_destroy"))(h);
    } else {
        void *h = ((p16_create_fn)dlsym(g_lib, "ref_p16_create"))();
        ((p16_ref_fn)dlsym(g_lib, "ref_p16_func_wiener"))(h, in, (int)body, out, fl, fl + nb, fl + 2 * nb, fl + 3 * nb, fl + 4 * nb,
                                                         gains, &rows, NULL, 0);
        ((p16_destroy_fn)dlsym(g_lib, "ref_p16_destroy"))(h);
    }
    for (i = 0; i < body; i++) cs += out[i];
    for (i = 0; i < (long)rows * 25; i++) cs += gains[i];
    free(in); free(out); free(fl); free(gains);
    return cs;
}

static double work_resynth(int port, const int16_t *x, long n, const float *mask, long frames)
{
    short *out = (short *)calloc((size_t)(n > 0 ? n : 1), sizeof(short));
    float *m = (float *)malloc(sizeof(float) * (size_t)(frames > 0 ? frames : 1) * 64);
    short *wav = (short *)malloc(sizeof(short) * (size_t)(n > 0 ? n : 1));
    long i;
    double cs = 0.0;
    memcpy(m, mask, sizeof(float) * (size_t)frames * 64);   /* the IBM variant overwrites gated entries with 1.0 */
    memcpy(wav, x, sizeof(short) * (size_t)n);
    if (port) ((rs_port_fn)dlsym(g_lib, "so_resynth"))(wav, n, m, (int)frames, 0, out, NULL);
    else ((rs_ref_fn)dlsym(g_lib, "ref_resynth"))(wav, n, m, (int)frames, out);
    for (i = 0; i < n; i++) cs += out[i];
    free(out); free(m); free(wav);
    return cs;
}

/* Runs utterances [0, n_utts) of the batch on `workers` forked processes (work queue in shared memory).
 * Returns 0 and *seconds (wall clock fork -> last worker reaped), *checksum (sum of every output value, independent of
 * the worker count), or a negative value on failure. */
int ref_bench_run(int workload, int port, const int16_t *samples, const int64_t *offsets, int n_utts, const float *mask,
                  const int64_t *frame_offsets, int workers, double *seconds, double *checksum)
{
    shared_t *sh;
    pid_t pids[256];
    int w, rc = 0;
    double t0, cs = 0.0;
    if (!g_lib || workers < 1 || workers > 256 || workload < 0 || workload > 2) return -1;
    sh = (shared_t *)mmap(NULL, sizeof(shared_t), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (sh == MAP_FAILED) return -2;
    memset(sh, 0, sizeof *sh);
    fflush(stdout);
    fflush(stderr);
    t0 = now_s();
    for (w = 0; w < workers; w++) {
        pids[w] = fork();
        if (pids[w] < 0) { rc = -3; workers = w; break; }
        if (pids[w] == 0) {
            double sum = 0.0;
            long done = 0;
            for (;;) {
                const int u = __sync_fetch_and_add(&sh->next, 1);
                if (u >= n_utts) break;
                const int16_t *x = samples + offsets[u];
                const long n = (long)(offsets[u + 1] - offsets[u]);
                if (workload == 0) sum += work_p8(port, x, n);
                else if (workload == 1) sum += work_p16(port, x, n);
                else sum += work_resynth(port, x, n, mask + frame_offsets[u] * 64, (long)(frame_offsets[u + 1] - frame_offsets[u]));
                done++;
            }
            sh->checksum[w] = sum;
            sh->units[w] = done;
            _exit(0);
        }
    }
    for (w = 0; w < workers; w++) {
        int st = 0;
        if (waitpid(pids[w], &st, 0) < 0 || !WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = -4;
    }
    *seconds = now_s() - t0;
    {
        long total = 0;
        for (w = 0; w < workers; w++) { cs += sh->checksum[w]; total += sh->units[w]; }
        if (rc == 0 && total != n_utts) rc = -5;
    }
    *checksum = cs;
    munmap(sh, sizeof *sh);
    return rc;
}
