// Persistent NoiseSup streams: the state of one DoNoiseSup / func_Wiener instance lives in device
// memory between pushes (se_b200.h "persistent NoiseSup streams").  Included by se_api_batch.cuh.
#pragma once

struct se_ns_stream {
    se_ctx *ctx;
    int profile;        // 8 | 16
    int zero_gate;      // NsLaunch::zero_check
    int fresh;          // 1 until the first push after create/reset: the kernel then initialises the state itself
    se::NsState *d_state;
    void *d_in, *d_out, *d_produced, *d_flags, *d_gains, *d_rows, *d_offsets, *d_ceps;
    uint64_t cap_blocks;
};

namespace se {

static int stream_reserve(se_ns_stream *s, uint64_t n_blocks)
{
    if (n_blocks <= s->cap_blocks) return SE_OK;
    se_ctx *ctx = s->ctx;
    SE_CUDA(cudaStreamSynchronize(ctx->cur->stream));
    void **bufs[] = {&s->d_in, &s->d_out, &s->d_produced, &s->d_flags, &s->d_gains, &s->d_ceps};
    for (void **b : bufs) { if (*b) cudaFree(*b); *b = nullptr; }
    s->cap_blocks = 0;
    const uint64_t cap = n_blocks + n_blocks / 4 + 16;
    const uint64_t shift = s->profile == 8 ? 80 : 160;
    SE_CUDA(cudaMalloc(&s->d_in, sizeof(float) * shift * cap));
    SE_CUDA(cudaMalloc(&s->d_out, sizeof(float) * shift * cap));
    SE_CUDA(cudaMalloc(&s->d_produced, sizeof(int32_t) * cap));
    SE_CUDA(cudaMalloc(&s->d_flags, sizeof(int32_t) * 5 * cap));
    SE_CUDA(cudaMalloc(&s->d_gains, sizeof(float) * 25 * cap));
    SE_CUDA(cudaMalloc(&s->d_ceps, sizeof(float) * 14 * cap));
    s->cap_blocks = cap;
    return SE_OK;
}

}  // namespace se

extern "C" {

int se_ns_stream_create(se_ctx *ctx, int profile, se_ns_stream **out)
{
    if (!ctx || !out || (profile != 8 && profile != 16)) { set_error("se_ns_stream_create: bad arguments"); return SE_EINVAL; }
    *out = nullptr;
    SE_CUDA(cudaSetDevice(ctx->device));
    se_ns_stream *s = new se_ns_stream();
    std::memset(s, 0, sizeof *s);
    s->ctx = ctx;
    s->profile = profile;
    s->fresh = 1;
    s->zero_gate = profile == 8 ? 0 : 2;   // DoNoiseSup has no gate; func_Wiener skips zero blocks anywhere
    cudaError_t e = cudaMalloc((void **)&s->d_state, sizeof(NsState));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_rows, sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_offsets, 2 * sizeof(int64_t));
    if (e != cudaSuccess) {
        se_ns_stream_destroy(s);
        return cuda_fail(e, "cudaMalloc(stream state)", __FILE__, __LINE__);
    }
    *out = s;
    return SE_OK;
}

void se_ns_stream_destroy(se_ns_stream *s)
{
    if (!s) return;
    std::lock_guard<std::recursive_mutex> guard(s->ctx->mu);
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->cur->stream);
    void *bufs[] = {s->d_state, s->d_in, s->d_out, s->d_produced, s->d_flags, s->d_gains, s->d_rows, s->d_offsets, s->d_ceps};
    for (void *b : bufs)
        if (b) cudaFree(b);
    delete s;
}

int se_ns_stream_reset(se_ns_stream *s)
{
    if (!s) return SE_EINVAL;
    s->fresh = 1;
    return SE_OK;
}

int se_ns_stream_set_zero_gate(se_ns_stream *s, int mode)
{
    if (!s || mode < 0 || mode > 2) return SE_EINVAL;
    s->zero_gate = mode;
    return SE_OK;
}

int se_ns_stream_push(se_ns_stream *s, const float *in, int n_blocks, float *out, int32_t *produced,
                      int32_t *flags, float *gains, int32_t *n_rows, float *ceps, int32_t *n_ceps)
{
    if (!s || n_blocks < 0 || (n_blocks > 0 && !in)) { set_error("se_ns_stream_push: bad arguments"); return SE_EINVAL; }
    if (n_rows) *n_rows = 0;
    if (n_ceps) *n_ceps = 0;
    if (n_blocks == 0) return SE_OK;
    se_ctx *ctx = s->ctx;
    std::lock_guard<std::recursive_mutex> guard(ctx->mu);    // the handle may be used on any host thread
    SE_CUDA(cudaSetDevice(ctx->device));
    SE_TRY(stream_reserve(s, (uint64_t)n_blocks));
    const int shift = s->profile == 8 ? 80 : 160;
    const int64_t n = (int64_t)n_blocks * shift;
    const int64_t h_off[2] = {0, n};
    SE_CUDA(cudaMemcpyAsync(s->d_offsets, h_off, sizeof h_off, cudaMemcpyHostToDevice, ctx->cur->stream));
    SE_CUDA(cudaMemcpyAsync(s->d_in, in, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, ctx->cur->stream));
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    a.samples = s->d_in;
    a.is_f32 = 1;
    a.offsets = (const int64_t *)s->d_offsets;
    a.n_utts = 1;
    a.state = s->d_state;
    a.load_state = s->fresh ? 0 : 1;
    a.produced = (int32_t *)s->d_produced;
    a.work_counter = ctx->cur->d_work_counter;
    if (s->profile == 8) {
        a.den_f32 = (float *)s->d_out;
        a.zero_check = s->zero_gate;
        if (ceps) {
            a.ceps = (float *)s->d_ceps;
            a.ceps_offsets = (const int64_t *)s->d_offsets;   // {0, n}: row 0, capacity n >= n_blocks
            a.n_ceps = (int32_t *)s->d_rows;
        }
        SE_TRY(ctx->exact ? se_exact::launch_ns_p8(ctx, a) : launch_ns_p8(ctx, a));
    } else {
        a.out_f32 = (float *)s->d_out;
        a.flags = (int32_t *)s->d_flags;
        a.total_blocks = n_blocks;
        a.gains = (float *)s->d_gains;
        a.n_rows = (int32_t *)s->d_rows;
        a.block_offsets = (const int64_t *)s->d_offsets;   // {0, ...}: rows and flags start at 0
        a.zero_check = s->zero_gate;
        SE_CUDA(cudaMemsetAsync(s->d_gains, 0, sizeof(float) * 25 * (size_t)n_blocks, ctx->cur->stream));
        SE_TRY(ctx->exact ? se_exact::launch_ns_p16(ctx, a) : launch_ns_p16(ctx, a));
    }
    s->fresh = 0;
    SE_TRY(d2h_async(ctx, out, s->d_out, sizeof(float) * (size_t)n));
    SE_TRY(d2h_async(ctx, produced, s->d_produced, sizeof(int32_t) * (size_t)n_blocks));
    if (s->profile == 8 && ceps) {
        SE_TRY(d2h_async(ctx, ceps, s->d_ceps, sizeof(float) * 14 * (size_t)n_blocks));
        SE_TRY(d2h_async(ctx, n_ceps, s->d_rows, sizeof(int32_t)));
    }
    if (s->profile == 16) {
        SE_TRY(d2h_async(ctx, flags, s->d_flags, sizeof(int32_t) * 5 * (size_t)n_blocks));
        SE_TRY(d2h_async(ctx, gains, s->d_gains, sizeof(float) * 25 * (size_t)n_blocks));
        SE_TRY(d2h_async(ctx, n_rows, s->d_rows, sizeof(int32_t)));
    }
    {
        const cudaError_t e = cudaStreamSynchronize(ctx->cur->stream);
        if (e != cudaSuccess) { ctx->bounces.clear(); ctx->bounce_used = 0; }
        SE_CUDA(e);
        finish_host_copies(ctx);
    }
    return SE_OK;
}

}  // extern "C"
