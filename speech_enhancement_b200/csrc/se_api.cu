// C ABI of libse_b200.so: context, tables upload, staging and the *_batch entry points (se_b200.h).
#include "se_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace se {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line)
{
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) return SE_ENODEVICE;
    if (e == cudaErrorMemoryAllocation) return SE_ENOMEM;
    return SE_ECUDA;
}

template <typename T> static int upload(se_ctx *ctx, const T *h, size_t count, const T **d)
{
    void *p = nullptr;
    SE_CUDA(cudaMalloc(&p, sizeof(T) * count));
    ctx->table_allocs.push_back(p);
    SE_CUDA(cudaMemcpy(p, h, sizeof(T) * count, cudaMemcpyHostToDevice));
    *d = reinterpret_cast<const T *>(p);
    return SE_OK;
}

static int upload_ns(se_ctx *ctx, const NsTables &h, NsDevTables &d, int bank_rows)
{
    int rc;
    if ((rc = upload(ctx, h.sig_window.data(), h.sig_window.size(), &d.sig_window))) return rc;
    if ((rc = upload(ctx, h.ir_window.data() + 8, 9, &d.fir_window))) return rc;
    std::vector<float> bankT((size_t)bank_rows * 32, 0.0f);
    int st[32] = {0}, ln[32] = {0};
    for (int j = 0; j < kMel; j++) {
        st[j] = h.bank_start[j];
        ln[j] = h.bank_len[j];
        if (h.bank_len[j] > bank_rows || (bank_rows != 128 && h.bank_len[j] > 10)) {   // kMelTaps (se_ns_stage.cuh)
            set_error("filter bank row %d too long (%d)", j, h.bank_len[j]);
            return SE_EINVAL;
        }
        if (bank_rows == 128) {
            // P16 dense Gamma bank: float4 per (band, lane) = the lane's 4 bins {lane, lane+32, lane+64, lane+96}
            for (int l = 0; l < 32; l++)
                for (int r = 0; r < 4; r++) bankT[((size_t)j * 32 + l) * 4 + r] = h.bank[(size_t)j * h.bank_w + l + 32 * r];
        } else {
            for (int i = 0; i < h.bank_len[j]; i++) bankT[(size_t)i * 32 + j] = h.bank[(size_t)j * h.bank_w + i];
        }
    }
    if ((rc = upload(ctx, bankT.data(), bankT.size(), &d.bankT))) return rc;
    if ((rc = upload(ctx, st, 32, &d.bank_start))) return rc;
    if ((rc = upload(ctx, ln, 32, &d.bank_len))) return rc;
    std::vector<float> idctT(25 * 16, 0.0f);
    for (int f = 0; f < kMel; f++)
        for (int t = 0; t < 9; t++) idctT[f * 16 + t] = h.idct[t * kMel + f];
    if ((rc = upload(ctx, idctT.data(), idctT.size(), &d.idctT))) return rc;
    std::vector<float> tw(2 * 132, 0.0f);
    for (int k = 0; k <= 128; k++) {
        tw[2 * k] = (float)std::cos(6.283185307179586476925287 * k / 256.0);
        tw[2 * k + 1] = (float)std::sin(6.283185307179586476925287 * k / 256.0);
    }
    if ((rc = upload(ctx, tw.data(), tw.size(), &d.tw_half))) return rc;
    d.eps = h.eps;
    // exact mode: the reference's rfft schedule (P8: rfft(x, 256, 8) of the C file; P16: the plugin's rfft(x, 512, 8))
    const RfftPlan &pl = (bank_rows == 128) ? ctx->h_rf16 : ctx->h_rf8;
    int lvl[32];
    for (int k = 0; k < 16; k++) { lvl[k] = k <= pl.m ? pl.level_begin[k] : 0; lvl[16 + k] = k <= pl.m ? pl.tw_begin[k] : 0; }
    const int *d_lvl = nullptr;
    if ((rc = upload(ctx, pl.ops.data(), pl.ops.size(), &d.rf_ops))) return rc;
    if ((rc = upload(ctx, lvl, 32, &d_lvl))) return rc;
    if ((rc = upload(ctx, pl.tw.data(), pl.tw.size(), &d.rf_tw))) return rc;
    d.rf_lvl = d_lvl;
    d.rf_m = pl.m;
    return SE_OK;
}

static int upload_cc(se_ctx *ctx)
{
    const CcTables &h = ctx->h_cc;
    CcDevTables &d = ctx->d_cc;
    int rc;
    float ham[200];
    for (int i = 0; i < 100; i++) { ham[i] = h.hamming[i]; ham[199 - i] = h.hamming[i]; }   // Window(), CompCeps.c:115-125
    if ((rc = upload(ctx, ham, 200, &d.hamming))) return rc;
    std::vector<float> triT(32 * 32, 0.0f);
    int st[32] = {0}, ln[32] = {0};
    for (int j = 0; j < 23; j++) {
        st[j] = h.start[j];
        ln[j] = h.len[j];
        if (h.len[j] > 32 || h.len[j] > (j >= 14 ? 24 : 12)) {   // 2 x kTriTaps for the bands that are split over two lanes
            set_error("CompCeps triangle %d too long (%d)", j, h.len[j]);
            return SE_EINVAL;
        }
        for (int i = 0; i < h.len[j]; i++) triT[i * 32 + j] = h.tri[j][i];
    }
    if ((rc = upload(ctx, triT.data(), triT.size(), &d.triT))) return rc;
    if ((rc = upload(ctx, st, 32, &d.tri_start))) return rc;
    if ((rc = upload(ctx, ln, 32, &d.tri_len))) return rc;
    std::vector<float> dctT(23 * 16, 0.0f);
    for (int j = 0; j < 23; j++)
        for (int t = 0; t < 12; t++) dctT[j * 16 + t] = h.dct[t * 23 + j];
    if ((rc = upload(ctx, dctT.data(), dctT.size(), &d.dctT))) return rc;
    // 16 kHz mode: bands 1..3 of the 5-band HP bank as dense rows (band 3 ends at bin 63), DCT over 26 bands
    const Hp16Tables &hp = ctx->h_hp16;
    std::vector<float> hpfb(3 * 64, 0.0f);
    for (int b = 0; b < 3; b++) {
        if (hp.fb_start[b + 1] + hp.fb_len[b + 1] > 64) { set_error("HP mel band %d leaves the 64-bin table", b + 1); return SE_EINVAL; }
        for (int i = 0; i < hp.fb_len[b + 1]; i++) hpfb[b * 64 + hp.fb_start[b + 1] + i] = hp.fb[b + 1][i];
    }
    if ((rc = upload(ctx, hpfb.data(), hpfb.size(), &d.hpfb))) return rc;
    std::vector<float> dct26T(26 * 16, 0.0f);
    for (int j = 0; j < 26; j++)
        for (int t = 0; t < 12; t++) dct26T[j * 16 + t] = hp.dct26[t * 26 + j];
    if ((rc = upload(ctx, dct26T.data(), dct26T.size(), &d.dct26T))) return rc;
    return SE_OK;
}

// Device -> host copy on the current lane's stream.  A cudaMemcpyAsync into pageable memory is synchronous: the driver
// waits inside the call for the kernel in front of it, holding a lock that stalls every other host thread's CUDA calls
// (measured: 8 threads calling the per-utterance adapters ran 1.6x faster than one thread; 7.6x with this).  So small
// copies into pageable memory land in a pinned arena first and are moved to the caller's array by
// finish_host_copies() right after the stream synchronisation that ends the API call.  Pinned / registered
// destinations (bench.py, se_host_alloc) and copies beyond the arena limit go direct.
constexpr uint64_t kBounceMax = 64ull << 20;

int d2h_async(se_ctx *ctx, void *host, const void *dev, uint64_t bytes)
{
    if (!host || bytes == 0) return SE_OK;
    bool direct = bytes > kBounceMax - ctx->bounce_used;
    if (!direct) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); direct = true; }
        else direct = at.type != cudaMemoryTypeUnregistered;
    }
    const uint64_t need = (bytes + 255) & ~255ull;
    if (!direct && ctx->bounce_used + need > ctx->bounce_cap) {
        if (!ctx->bounces.empty()) direct = true;               // arena in use: cannot move it
        else {
            if (ctx->bounce) cudaFreeHost(ctx->bounce);
            ctx->bounce = nullptr;
            ctx->bounce_cap = 0;
            uint64_t want = need + need / 4;
            if (want < (1ull << 20)) want = 1ull << 20;
            if (want > kBounceMax) want = kBounceMax;
            if (cudaHostAlloc(&ctx->bounce, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); ctx->bounce = nullptr; direct = true; }
            else ctx->bounce_cap = want;
        }
    }
    if (direct) {
        SE_CUDA(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
        return SE_OK;
    }
    void *pin = (char *)ctx->bounce + ctx->bounce_used;
    SE_CUDA(cudaMemcpyAsync(pin, dev, bytes, cudaMemcpyDeviceToHost, ctx->cur->stream));
    ctx->bounces.push_back({host, pin, bytes});
    ctx->bounce_used += need;
    return SE_OK;
}

// Call after the stream(s) carrying d2h_async copies have been synchronised.
void finish_host_copies(se_ctx *ctx)
{
    for (const auto &b : ctx->bounces) std::memcpy(b.dst, b.src, b.bytes);
    ctx->bounces.clear();
    ctx->bounce_used = 0;
}

int ensure_stage(se_ctx *ctx, int slot, uint64_t bytes, void **out)
{
    if (bytes == 0) bytes = 16;
    if (ctx->cur->stage_bytes[slot] < bytes) {
        if (ctx->cur->d_stage[slot]) {
            SE_CUDA(cudaStreamSynchronize(ctx->cur->stream));
            SE_CUDA(cudaFree(ctx->cur->d_stage[slot]));
            ctx->cur->d_stage[slot] = nullptr;
            ctx->cur->stage_bytes[slot] = 0;
        }
        const uint64_t want = bytes + bytes / 8;
        SE_CUDA(cudaMalloc(&ctx->cur->d_stage[slot], want));
        ctx->cur->stage_bytes[slot] = want;
    }
    *out = ctx->cur->d_stage[slot];
    return SE_OK;
}

int ensure_scratch(se_ctx *ctx, uint64_t bytes)
{
    if (ctx->scratch_bytes < bytes) {
        if (ctx->d_scratch) {
            SE_CUDA(cudaStreamSynchronize(ctx->cur->stream));
            SE_CUDA(cudaFree(ctx->d_scratch));
            ctx->d_scratch = nullptr;
            ctx->scratch_bytes = 0;
        }
        SE_CUDA(cudaMalloc(&ctx->d_scratch, bytes));
        ctx->scratch_bytes = bytes;
    }
    return SE_OK;
}

void prof_begin(se_ctx *ctx, int k)
{
    if (!ctx->profiling) return;
    se_ctx::Ev e;
    e.k = k;
    cudaEventCreate(&e.a);
    cudaEventCreate(&e.b);
    cudaEventRecord(e.a, ctx->cur->stream);
    ctx->pending.push_back(e);
}

void prof_end(se_ctx *ctx, int k)
{
    ctx->launches++;
    (void)k;
    if (!ctx->profiling) return;
    cudaEventRecord(ctx->pending.back().b, ctx->cur->stream);
}

static int drain_events(se_ctx *ctx)
{
    for (auto &e : ctx->pending) {
        float ms = 0.0f;
        SE_CUDA(cudaEventSynchronize(e.b));
        SE_CUDA(cudaEventElapsedTime(&ms, e.a, e.b));
        ctx->k_ms[e.k] += ms;
        ctx->k_launches[e.k]++;
        cudaEventDestroy(e.a);
        cudaEventDestroy(e.b);
    }
    ctx->pending.clear();
    return SE_OK;
}

}  // namespace se

using namespace se;

extern "C" {

const char *se_last_error(void) { return g_err; }

int se_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// Device-side part of se_ctx_create; any failure leaves a partially built context for se_ctx_destroy.
static int ctx_init_device(se_ctx *ctx)
{
    const int device = ctx->device;
    cudaDeviceProp prop;
    SE_CUDA(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    SE_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    SE_CUDA(cudaStreamCreateWithFlags(&ctx->up_stream, cudaStreamNonBlocking));
    ctx->lanes[0].stream = ctx->own_stream;
    for (int l = 0; l < kLanes; l++) {
        if (l > 0) SE_CUDA(cudaStreamCreateWithFlags(&ctx->lanes[l].stream, cudaStreamNonBlocking));
        SE_CUDA(cudaMalloc((void **)&ctx->lanes[l].d_work_counter, sizeof(int)));
        SE_CUDA(cudaEventCreateWithFlags(&ctx->lanes[l].done, cudaEventDisableTiming));
    }
    build_rfft_plan(ctx->h_rf8, 256, 8, false);
    build_rfft_plan(ctx->h_rf16, 512, 8, true);
    build_ns_tables_p8(ctx->h_p8);
    build_ns_tables_p16(ctx->h_p16);
    build_cc_tables(ctx->h_cc);
    build_gt_tables(ctx->h_gt);
    build_hp16_tables(ctx->h_hp16);
    int rc;
    if ((rc = upload_ns(ctx, ctx->h_p8, ctx->d_p8, 16)) || (rc = upload_ns(ctx, ctx->h_p16, ctx->d_p16, 128)) ||
        (rc = upload_cc(ctx)) || (rc = upload_g64_tables(ctx)))
        return rc;
    return SE_OK;
}

int se_ctx_create(int device, se_ctx **out)
{
    if (!out) return SE_EINVAL;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        set_error("no CUDA device available (libse_b200 has no CPU path): %s", cudaGetErrorString(e));
        return SE_ENODEVICE;
    }
    if (device < 0 || device >= n) { set_error("device %d out of range (0..%d)", device, n - 1); return SE_EINVAL; }
    SE_CUDA(cudaSetDevice(device));
    se_ctx *ctx = new se_ctx();
    ctx->device = device;
    ctx->exact = 0;
    ctx->own_stream = nullptr;
    std::memset(ctx->lanes, 0, sizeof ctx->lanes);
    ctx->cur = &ctx->lanes[0];
    ctx->d_scratch = nullptr;
    ctx->scratch_bytes = 0;
    ctx->scratch_limit = 48ull << 30;
    ctx->chunk_bytes = 512ull << 20;
    ctx->chunk_min_utts = 1024;
    ctx->chunk_lead = 3;
    ctx->up_stream = nullptr;
    ctx->feature_flush = 0;
    ctx->ns_pipeline = 0;
    ctx->bounce = nullptr;
    ctx->bounce_cap = ctx->bounce_used = 0;
    ctx->profiling = 0;
    std::memset(ctx->k_ms, 0, sizeof ctx->k_ms);
    std::memset(ctx->k_launches, 0, sizeof ctx->k_launches);
    ctx->launches = 0;
    const int rc = ctx_init_device(ctx);
    if (rc) {                       // partially built context: release what exists
        se_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return SE_OK;
}

void se_ctx_destroy(se_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (void *p : ctx->table_allocs) cudaFree(p);
    for (int l = 0; l < kLanes; l++) {
        se_lane &ln = ctx->lanes[l];
        for (int i = 0; i < 16; i++)
            if (ln.d_stage[i]) cudaFree(ln.d_stage[i]);
        if (ln.d_work_counter) cudaFree(ln.d_work_counter);
        if (ln.done) cudaEventDestroy(ln.done);
        if (l > 0 && ln.stream) cudaStreamDestroy(ln.stream);
    }
    if (ctx->d_scratch) cudaFree(ctx->d_scratch);
    if (ctx->bounce) cudaFreeHost(ctx->bounce);
    for (auto &e : ctx->pending) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    if (ctx->up_stream) cudaStreamDestroy(ctx->up_stream);
    for (cudaEvent_t e : ctx->up_events) cudaEventDestroy(e);
    delete ctx;
}

int se_ctx_set_stream(se_ctx *ctx, void *cuda_stream)
{
    if (!ctx) return SE_EINVAL;
    ctx->lanes[0].stream = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
    return SE_OK;
}

int se_ctx_sync(se_ctx *ctx)
{
    if (!ctx) return SE_EINVAL;
    std::lock_guard<std::recursive_mutex> guard(ctx->mu);
    SE_CUDA(cudaSetDevice(ctx->device));
    SE_CUDA(cudaStreamSynchronize(ctx->cur->stream));
    return SE_OK;
}

int se_ctx_set_scratch_limit(se_ctx *ctx, uint64_t bytes)
{
    if (!ctx || bytes < (1ull << 20)) return SE_EINVAL;
    ctx->scratch_limit = bytes;
    return SE_OK;
}

int se_ctx_set_pipeline(se_ctx *ctx, uint64_t chunk_bytes, int min_utts)
{
    if (!ctx || chunk_bytes == 0 || min_utts < 1) return SE_EINVAL;
    ctx->chunk_bytes = chunk_bytes;
    ctx->chunk_min_utts = min_utts;
    return SE_OK;
}

int se_ctx_set_pipeline_lead(se_ctx *ctx, int enable)
{
    if (!ctx || enable < 0 || enable > 3) return SE_EINVAL;
    ctx->chunk_lead = enable;
    return SE_OK;
}

int se_ctx_set_exact(se_ctx *ctx, int enable)
{
    if (!ctx) return SE_EINVAL;
    ctx->exact = enable ? 1 : 0;
    return SE_OK;
}

int se_ctx_set_ns_pipeline(se_ctx *ctx, int mode)
{
    if (!ctx || mode < 0 || mode > 2) return SE_EINVAL;
    ctx->ns_pipeline = mode;
    return SE_OK;
}

int se_ctx_set_feature_flush(se_ctx *ctx, int enable)
{
    if (!ctx) return SE_EINVAL;
    ctx->feature_flush = enable ? 1 : 0;
    return SE_OK;
}

int se_ctx_profiling(se_ctx *ctx, int enable)
{
    if (!ctx) return SE_EINVAL;
    ctx->profiling = enable ? 1 : 0;
    return SE_OK;
}

int se_ctx_kernel_time(se_ctx *ctx, int kernel_id, double *total_ms, int64_t *launches)
{
    if (!ctx || kernel_id < 0 || kernel_id >= SE_K_COUNT) return SE_EINVAL;
    std::lock_guard<std::recursive_mutex> guard(ctx->mu);
    int rc = drain_events(ctx);
    if (rc) return rc;
    if (total_ms) *total_ms = ctx->k_ms[kernel_id];
    if (launches) *launches = ctx->k_launches[kernel_id];
    return SE_OK;
}

int se_ctx_reset_kernel_times(se_ctx *ctx)
{
    if (!ctx) return SE_EINVAL;
    std::lock_guard<std::recursive_mutex> guard(ctx->mu);
    int rc = drain_events(ctx);
    if (rc) return rc;
    std::memset(ctx->k_ms, 0, sizeof ctx->k_ms);
    std::memset(ctx->k_launches, 0, sizeof ctx->k_launches);
    return SE_OK;
}

int64_t se_ctx_launch_count(se_ctx *ctx) { return ctx ? ctx->launches : 0; }

int se_shard_plan(const int64_t *offsets, int n_utts, int n_shards, int32_t *first_utt)
{
    if (!offsets || !first_utt || n_utts < 0 || n_shards < 1) { set_error("se_shard_plan: bad arguments"); return SE_EINVAL; }
    const int64_t base = offsets[0], total = offsets[n_utts] - base;
    first_utt[0] = 0;
    int u = 0;
    for (int k = 1; k < n_shards; k++) {
        const int64_t target = base + (int64_t)((__int128)total * k / n_shards);
        while (u < n_utts && offsets[u] < target) u++;           // first boundary at or after the target
        if (u > first_utt[k - 1] && offsets[u] - target > target - offsets[u - 1]) u--;   // nearest boundary
        first_utt[k] = u;
    }
    first_utt[n_shards] = n_utts;
    return SE_OK;
}

int64_t se_etsi_num_ceps_frames(int64_t n) { const int64_t c = n / 80 - 6; return c > 0 ? c : 0; }
int64_t se_etsi_num_feature_rows(int64_t n) { return n > 0 ? n / 80 : 0; }
int64_t se_etsi16k_num_feature_rows(int64_t n) { return n > 0 ? n / 160 : 0; }
int64_t se_ibm_num_frames(int64_t n) { return n >= 320 ? (n - 320) / 160 + 1 : 0; }
int64_t se_resynth_num_frames(int64_t n, int mode)
{
    if (mode == SE_RESYNTH_1DNN) return n / 160;
    const int64_t f = (n - 320) / 160 + 1;   // C integer division truncates toward zero (extractwav.cpp:67)
    return f > 0 ? f : 0;
}

}  // extern "C"

#include "se_api_batch.cuh"
