// se_resynth_noalloc(): resynth() for ONE utterance with NO device or host allocation in the call -- for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It has no
// counterpart in the reference.
//
// The two resynthesis passes need scratch memory (the [sample][64] float intermediate between the forward and the
// time-reversed gammatone pass, csrc/se_g64_resynth.cuh) and a list of work items.  se_resynth_batch takes both from
// cudaMalloc'ed context buffers; here they live in ONE file-scope __device__ array of fixed size that every GPU which
// loads the library holds until the process exits.  The work items are planned ON THE DEVICE by a one-thread kernel
// (same plan as launch_resynth, so the output equals se_resynth_batch's bit for bit), which also avoids any
// host-to-device copy.  Included at the end of se_g64.cu: same translation unit as the array and the kernels.
#pragma once
#include <mutex>

namespace se {

constexpr size_t kNaScratchBytes = 64ull << 20;               // 64 MiB per GPU, resident for the life of the process
constexpr size_t kNaItemBytes = 64ull << 10;                  // head of the array: the work items
constexpr int64_t kNaMaxRows = (int64_t)((kNaScratchBytes - kNaItemBytes) / (64 * sizeof(float)));   // 261 888 rows
constexpr int kNaMaxItems = (int)(kNaItemBytes / sizeof(GtItem));

static __device__ __align__(256) unsigned char g_na_scratch[kNaScratchBytes];

// The plan of launch_resynth (se_g64_host.cuh) for one utterance, written by one thread.
__global__ void na_plan_kernel(GtItem *items, int n, int frames, int step, int exact, int count)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int64_t rows = 0;
    for (int i = 0; i < count; i++) {
        GtItem it;
        const int a = i * step;
        const int b = (a + step < n) ? a + step : n;
        it.sample_off = 0;
        it.frame_off = 0;
        it.frames = frames;
        it.a = a;
        it.b = b;
        it.fs = exact ? 0 : ((a - kGtWarm > 0) ? a - kGtWarm : 0);
        it.fe = exact ? n : ((b + kGtWarm < n) ? b + kGtWarm : n);
        it.y_off = rows;
        rows += it.fe - it.a;
        items[i] = it;
    }
}

// Host-side plan: chunk length, item count and scratch rows for an utterance of n samples (mirrors launch_resynth).
static void na_plan_host(int64_t n, bool exact, int flags, int64_t *step, int *count, int64_t *rows)
{
    *step = exact ? n : gt_chunk_len(n, flags);
    *count = (int)((n + *step - 1) / *step);
    int64_t r = 0;
    for (int64_t a = 0; a < n; a += *step) {
        const int64_t b = (a + *step < n) ? a + *step : n;
        const int64_t fe = exact ? n : ((b + kGtWarm < n) ? b + kGtWarm : n);
        r += fe - a;
    }
    *rows = r;
}

// Device address of g_na_scratch, looked up once per device with that device current.  In host code the array's name
// is only a host-side stand-in for the symbol: it must never be handed to a kernel.
static unsigned char *na_scratch_ptr(int device, int *rc)
{
    static std::mutex mu;
    static unsigned char *ptr[64] = {nullptr};
    *rc = SE_OK;
    if (device < 0 || device >= 64) { set_error("se_resynth_noalloc: device index %d out of range", device); *rc = SE_EINVAL; return nullptr; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_na_scratch);
        if (e != cudaSuccess || !p) { *rc = cuda_fail(e, "cudaGetSymbolAddress(g_na_scratch)", __FILE__, __LINE__); return nullptr; }
        ptr[device] = reinterpret_cast<unsigned char *>(p);
    }
    return ptr[device];
}

int launch_resynth_noalloc(se_ctx *ctx, const int16_t *d_samples, int64_t n, const float *d_mask, int mode, int flags,
                           int16_t *d_out)
{
    const bool exact = (flags & SE_FLAG_EXACT) != 0;
    int64_t step = 0, rows = 0;
    int count = 0;
    na_plan_host(n, exact, flags, &step, &count, &rows);
    if (n > SE_RESYNTH_NOALLOC_MAX_SAMPLES || rows > kNaMaxRows || count > kNaMaxItems) {
        set_error("se_resynth_noalloc: %lld samples exceed the fixed scratch (maximum %d samples per call; "
                  "use se_resynth_batch for longer utterances)", (long long)n, SE_RESYNTH_NOALLOC_MAX_SAMPLES);
        return SE_EINVAL;
    }
    int rc = SE_OK;
    unsigned char *base = na_scratch_ptr(ctx->device, &rc);
    if (!base) return rc;
    GtItem *items = reinterpret_cast<GtItem *>(base);
    float *Y = reinterpret_cast<float *>(base + kNaItemBytes);
    float *out_f = (flags & SE_FLAG_OUT_F32) ? reinterpret_cast<float *>(d_out) : nullptr;
    const int frames = (int)se_resynth_num_frames(n, mode);
    cudaStream_t st = ctx->cur->stream;
    // every row the reversed pass reads is written by the forward pass of the same call and every item by the plan
    // kernel, so nothing in the shared array has to start at zero; stale contents of earlier calls are never read
    na_plan_kernel<<<1, 32, 0, st>>>(items, (int)n, frames, (int)step, exact ? 1 : 0, count);
    if (exact) gt_fwd_kernel<true><<<count, 64, 0, st>>>(d_samples, items, Y);
    else gt_fwd_fast_kernel<<<count, 64, 0, st>>>(d_samples, items, Y);
    if (exact) launch_bwd<true>(ctx, mode, count, Y, items, d_mask, d_out, out_f);
    else if (mode == SE_RESYNTH_IBM) gt_bwd_fast_kernel<SE_RESYNTH_IBM><<<count, 64, 0, st>>>(Y, items, d_mask, d_out, out_f);
    else gt_bwd_fast_kernel<SE_RESYNTH_ORI><<<count, 64, 0, st>>>(Y, items, d_mask, d_out, out_f);
    ctx->launches += 3;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

}  // namespace se
