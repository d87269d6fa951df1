// The *_batch entry points of se_b200.h: staging (for SE_MEM_HOST callers) + kernel launches.
// Included at the end of se_api.cu.
#pragma once

namespace se {

// Staging slots (ctx->d_stage).  One call uses each slot for at most one array.
enum {
    kSlotOffsets = 0, kSlotOffsets2 = 1, kSlotOffsets3 = 2, kSlotIn = 3, kSlotIn2 = 4,
    kSlotOut = 5, kSlotOut2 = 6, kSlotOut3 = 7, kSlotOut4 = 8, kSlotOut5 = 9, kSlotOut6 = 10,
    kSlotGain = 15
};

// offsets / frame-offset arrays are HOST arrays in both memory modes: upload a device copy.
static int upload_offsets(se_ctx *ctx, int slot, const int64_t *h, int count, const int64_t **d)
{
    *d = nullptr;
    if (!h) return SE_OK;
    void *p = nullptr;
    int rc = ensure_stage(ctx, slot, sizeof(int64_t) * (size_t)count, &p);
    if (rc) return rc;
    SE_CUDA(cudaMemcpyAsync(p, h, sizeof(int64_t) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
    *d = reinterpret_cast<const int64_t *>(p);
    return SE_OK;
}

static int check_offsets(const int64_t *o, int n, const char *what)
{
    if (!o) { set_error("%s: offsets is NULL", what); return SE_EINVAL; }
    for (int i = 0; i < n; i++)
        if (o[i + 1] < o[i]) { set_error("%s: offsets not monotone at %d", what, i); return SE_EINVAL; }
    return SE_OK;
}

// An input array: device pointer as is (SE_MEM_DEVICE) or an H2D copy into a staging slot.
static int stage_in(se_ctx *ctx, int slot, const void *p, uint64_t bytes, int mem, const void **d)
{
    *d = nullptr;
    if (!p) return SE_OK;
    if (mem == SE_MEM_DEVICE) { *d = p; return SE_OK; }
    void *s = nullptr;
    int rc = ensure_stage(ctx, slot, bytes, &s);
    if (rc) return rc;
    if (bytes) SE_CUDA(cudaMemcpyAsync(s, p, bytes, cudaMemcpyHostToDevice, ctx->stream));
    *d = s;
    return SE_OK;
}

// An output array: device pointer as is, or a staging slot that stage_out_copy() later reads back.
static int stage_out(se_ctx *ctx, int slot, void *p, uint64_t bytes, int mem, bool zero, void **d)
{
    *d = nullptr;
    if (!p) return SE_OK;
    if (mem == SE_MEM_DEVICE) {
        *d = p;
    } else {
        int rc = ensure_stage(ctx, slot, bytes, d);
        if (rc) return rc;
    }
    if (zero && bytes) SE_CUDA(cudaMemsetAsync(*d, 0, bytes, ctx->stream));
    return SE_OK;
}

static int stage_out_copy(se_ctx *ctx, void *host, const void *dev, uint64_t bytes, int mem)
{
    if (!host || mem == SE_MEM_DEVICE || bytes == 0) return SE_OK;
    SE_CUDA(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return SE_OK;
}

static int enter(se_ctx *ctx, int mem, const char *what)
{
    if (!ctx) { set_error("%s: NULL context", what); return SE_EINVAL; }
    if (mem != SE_MEM_HOST && mem != SE_MEM_DEVICE) { set_error("%s: bad mem kind %d", what, mem); return SE_EINVAL; }
    SE_CUDA(cudaSetDevice(ctx->device));
    return SE_OK;
}

static int leave(se_ctx *ctx, int mem)
{
    if (mem == SE_MEM_HOST) SE_CUDA(cudaStreamSynchronize(ctx->stream));
    return SE_OK;
}

#define SE_TRY(expr) do { int rc__ = (expr); if (rc__) return rc__; } while (0)

}  // namespace se

extern "C" {

void *se_host_alloc(uint64_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        se::set_error("se_host_alloc: cudaHostAlloc(%llu) failed", (unsigned long long)bytes);
        return nullptr;
    }
    return p;
}

void se_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

int se_etsi_denoise_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                          int16_t *denoised, float *ceps, const int64_t *ceps_offsets, int32_t *n_ceps,
                          float *den_f32, uint8_t *vad_trace, const int64_t *trace_offsets, int mem)
{
    SE_TRY(enter(ctx, mem, "se_etsi_denoise_batch"));
    if (n_utts < 0 || (n_utts > 0 && !samples)) { set_error("se_etsi_denoise_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_etsi_denoise_batch"));
    if (ceps && !ceps_offsets) { set_error("se_etsi_denoise_batch: ceps needs ceps_offsets"); return SE_EINVAL; }
    if (vad_trace && !trace_offsets) { set_error("se_etsi_denoise_batch: vad_trace needs trace_offsets"); return SE_EINVAL; }
    const int64_t base = offsets[0], total = offsets[n_utts] - base;
    if (base != 0) { set_error("se_etsi_denoise_batch: offsets[0] must be 0"); return SE_EINVAL; }
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    const uint64_t ceps_bytes = ceps ? sizeof(float) * 14 * (uint64_t)ceps_offsets[n_utts] : 0;
    const uint64_t trace_bytes = vad_trace ? (uint64_t)trace_offsets[n_utts] : 0;
    void *d_den = nullptr, *d_ceps = nullptr, *d_nceps = nullptr, *d_denf = nullptr, *d_trace = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &a.offsets));
    SE_TRY(upload_offsets(ctx, kSlotOffsets2, ceps ? ceps_offsets : nullptr, n_utts + 1, &a.ceps_offsets));
    SE_TRY(upload_offsets(ctx, kSlotOffsets3, vad_trace ? trace_offsets : nullptr, n_utts + 1, &a.trace_offsets));
    SE_TRY(stage_in(ctx, kSlotIn, samples, sizeof(int16_t) * (uint64_t)total, mem, &a.samples));
    SE_TRY(stage_out(ctx, kSlotOut, denoised, sizeof(int16_t) * (uint64_t)total, mem, false, &d_den));
    SE_TRY(stage_out(ctx, kSlotOut2, ceps, ceps_bytes, mem, true, &d_ceps));
    SE_TRY(stage_out(ctx, kSlotOut3, n_ceps, sizeof(int32_t) * (uint64_t)n_utts, mem, false, &d_nceps));
    SE_TRY(stage_out(ctx, kSlotOut4, den_f32, sizeof(float) * (uint64_t)total, mem, false, &d_denf));
    SE_TRY(stage_out(ctx, kSlotOut5, vad_trace, trace_bytes, mem, true, &d_trace));
    a.is_f32 = 0;
    a.n_utts = n_utts;
    a.den_i16 = (int16_t *)d_den;
    a.ceps = (float *)d_ceps;
    a.n_ceps = (int32_t *)d_nceps;
    a.den_f32 = (float *)d_denf;
    a.vad_trace = (uint8_t *)d_trace;
    a.zero_check = 1;
    a.work_counter = ctx->d_work_counter;
    SE_TRY(launch_ns_p8(ctx, a));
    SE_TRY(stage_out_copy(ctx, denoised, d_den, sizeof(int16_t) * (uint64_t)total, mem));
    SE_TRY(stage_out_copy(ctx, ceps, d_ceps, ceps_bytes, mem));
    SE_TRY(stage_out_copy(ctx, n_ceps, d_nceps, sizeof(int32_t) * (uint64_t)n_utts, mem));
    SE_TRY(stage_out_copy(ctx, den_f32, d_denf, sizeof(float) * (uint64_t)total, mem));
    SE_TRY(stage_out_copy(ctx, vad_trace, d_trace, trace_bytes, mem));
    return leave(ctx, mem);
}

int se_p16_wiener_batch(se_ctx *ctx, const void *samples, int is_f32, const int64_t *offsets, int n_utts,
                        float *out_f32, int32_t *flags, float *gains, int32_t *n_rows,
                        const int64_t *block_offsets, int mem)
{
    SE_TRY(enter(ctx, mem, "se_p16_wiener_batch"));
    if (n_utts < 0 || (n_utts > 0 && !samples)) { set_error("se_p16_wiener_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_p16_wiener_batch"));
    if ((flags || gains) && !block_offsets) { set_error("se_p16_wiener_batch: flags/gains need block_offsets"); return SE_EINVAL; }
    if (offsets[0] != 0) { set_error("se_p16_wiener_batch: offsets[0] must be 0"); return SE_EINVAL; }
    if (block_offsets)
        for (int u = 0; u < n_utts; u++)
            if (block_offsets[u + 1] - block_offsets[u] < (offsets[u + 1] - offsets[u]) / 160) {
                set_error("se_p16_wiener_batch: block_offsets too small for utterance %d", u);
                return SE_EINVAL;
            }
    const int64_t total = offsets[n_utts];
    const int64_t total_blocks = block_offsets ? block_offsets[n_utts] : 0;
    const uint64_t in_bytes = (is_f32 ? sizeof(float) : sizeof(int16_t)) * (uint64_t)total;
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    void *d_out = nullptr, *d_flags = nullptr, *d_gains = nullptr, *d_rows = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &a.offsets));
    SE_TRY(upload_offsets(ctx, kSlotOffsets2, block_offsets, n_utts + 1, &a.block_offsets));
    SE_TRY(stage_in(ctx, kSlotIn, samples, in_bytes, mem, &a.samples));
    SE_TRY(stage_out(ctx, kSlotOut, out_f32, sizeof(float) * (uint64_t)total, mem, false, &d_out));
    SE_TRY(stage_out(ctx, kSlotOut2, flags, sizeof(int32_t) * 5 * (uint64_t)total_blocks, mem, true, &d_flags));
    SE_TRY(stage_out(ctx, kSlotOut3, gains, sizeof(float) * 25 * (uint64_t)total_blocks, mem, true, &d_gains));
    SE_TRY(stage_out(ctx, kSlotOut4, n_rows, sizeof(int32_t) * (uint64_t)n_utts, mem, false, &d_rows));
    a.is_f32 = is_f32 ? 1 : 0;
    a.n_utts = n_utts;
    a.out_f32 = (float *)d_out;
    a.flags = (int32_t *)d_flags;
    a.total_blocks = total_blocks;
    a.gains = (float *)d_gains;
    a.n_rows = (int32_t *)d_rows;
    a.zero_check = 2;
    a.work_counter = ctx->d_work_counter;
    SE_TRY(launch_ns_p16(ctx, a));
    SE_TRY(stage_out_copy(ctx, out_f32, d_out, sizeof(float) * (uint64_t)total, mem));
    SE_TRY(stage_out_copy(ctx, flags, d_flags, sizeof(int32_t) * 5 * (uint64_t)total_blocks, mem));
    SE_TRY(stage_out_copy(ctx, gains, d_gains, sizeof(float) * 25 * (uint64_t)total_blocks, mem));
    SE_TRY(stage_out_copy(ctx, n_rows, d_rows, sizeof(int32_t) * (uint64_t)n_utts, mem));
    return leave(ctx, mem);
}

}  // extern "C"

#include "se_api_stream.cuh"
#include "se_api_g64.cuh"
