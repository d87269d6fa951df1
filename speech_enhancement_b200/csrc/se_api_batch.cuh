// The *_batch entry points of se_b200.h: staging (for SE_MEM_HOST callers) + kernel launches.
// Included at the end of se_api.cu.
#pragma once

namespace se {

// Staging slots (ctx->cur->d_stage).  One call uses each slot for at most one array.
enum {
    kSlotOffsets = 0, kSlotOffsets2 = 1, kSlotOffsets3 = 2, kSlotIn = 3, kSlotIn2 = 4,
    kSlotOut = 5, kSlotOut2 = 6, kSlotOut3 = 7, kSlotOut4 = 8, kSlotOut5 = 9, kSlotOut6 = 10,
    kSlotInAll = 13, kSlotGain = 15
};

// offsets / frame-offset arrays are HOST arrays in both memory modes: upload a device copy.
static int upload_offsets(se_ctx *ctx, int slot, const int64_t *h, int count, const int64_t **d)
{
    *d = nullptr;
    if (!h) return SE_OK;
    void *p = nullptr;
    int rc = ensure_stage(ctx, slot, sizeof(int64_t) * (size_t)count, &p);
    if (rc) return rc;
    SE_CUDA(cudaMemcpyAsync(p, h, sizeof(int64_t) * (size_t)count, cudaMemcpyHostToDevice, ctx->cur->stream));
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
    if (bytes) SE_CUDA(cudaMemcpyAsync(s, p, bytes, cudaMemcpyHostToDevice, ctx->cur->stream));
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
    if (zero && bytes) SE_CUDA(cudaMemsetAsync(*d, 0, bytes, ctx->cur->stream));
    return SE_OK;
}

static int stage_out_copy(se_ctx *ctx, void *host, const void *dev, uint64_t bytes, int mem)
{
    if (!host || mem == SE_MEM_DEVICE || bytes == 0) return SE_OK;
    return d2h_async(ctx, host, dev, bytes);
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
    if (mem == SE_MEM_HOST) {
        const cudaError_t e = cudaStreamSynchronize(ctx->cur->stream);
        if (e != cudaSuccess) { ctx->bounces.clear(); ctx->bounce_used = 0; }
        SE_CUDA(e);
        finish_host_copies(ctx);
    }
    return SE_OK;
}

#define SE_TRY(expr) do { int rc__ = (expr); if (rc__) return rc__; } while (0)
// argument checks + cudaSetDevice, then the context lock for the rest of the call (se_ctx::mu)
#define SE_ENTER(ctx, mem, what)                                   \
    SE_TRY(enter(ctx, mem, what));                                 \
    std::lock_guard<std::recursive_mutex> se_ctx_guard__((ctx)->mu)


// ------------------------------------------------------------------------------------------------
// Shared driver of the two NoiseSup batch calls.
//
// SE_MEM_DEVICE: one launch on the context stream, pointers used as given.
// SE_MEM_HOST  : the batch is cut into contiguous utterance chunks (balanced by samples); chunk c runs
//                H2D -> kernel -> D2H in order on lane c % kLanes, so the copies of one chunk overlap the
//                kernels of its neighbours (PCIe both directions + SMs busy at once).
// In both modes utterances are queued longest-first inside a launch so the persistent warps finish
// together (the reference's one-file-at-a-time loop has no such concern).
// ------------------------------------------------------------------------------------------------
enum { kIdxSample = 0, kIdxAux1 = 1, kIdxAux2 = 2 };
struct NsOut {
    void *host;            // caller's array (host or device pointer), may be NULL
    uint64_t elem;         // bytes per indexed element
    int index;             // which offsets array indexes it
    bool zero;             // clear before the launch (rows the kernel may leave unwritten)
    bool planes5;          // int32 [5][total(aux1)] layout (P16 flags)
    int slot;
    size_t field;          // offsetof(NsLaunch, <pointer field>)
};
constexpr int kSlotOrder = 11, kSlotCounts = 12;

static void set_field(NsLaunch &a, size_t field, void *p) { *reinterpret_cast<void **>(reinterpret_cast<char *>(&a) + field) = p; }

static int ns_batch_run(se_ctx *ctx, int profile, NsLaunch a, const void *samples, uint64_t in_elem,
                        const int64_t *offsets, const int64_t *aux1, const int64_t *aux2, int n_utts,
                        NsOut *outs, int n_outs, int32_t *counts, size_t counts_field, int mem)
{
    const int64_t *idx[3] = {offsets, aux1, aux2};
    se_lane *lane0 = &ctx->lanes[0];
    ctx->cur = lane0;
    // queue order: longest utterance first within each chunk
    uint64_t io_bytes = in_elem * (uint64_t)offsets[n_utts];
    for (int k = 0; k < n_outs; k++)
        if (outs[k].host) io_bytes += outs[k].elem * (uint64_t)idx[outs[k].index][n_utts] * (outs[k].planes5 ? 5 : 1);
    int n_chunks = 1;
    if (mem == SE_MEM_HOST) {
        const uint64_t by_bytes = io_bytes / ctx->chunk_bytes, by_utts = (uint64_t)n_utts / (uint64_t)ctx->chunk_min_utts;
        n_chunks = (int)std::max<uint64_t>(1, std::min<uint64_t>(std::min(by_bytes, by_utts), 32));
    }
    // Pipelined batches open with a SMALL chunk: as many utterances as the single-utterance kernel (se_ns_pipe.cuh) runs at
    // once.  It finishes a 10 s utterance in ~1.6 ms where the batch kernel's one warp needs 5-6.5 ms, so the D2H engine --
    // the bottleneck of the whole call -- starts ~10 ms earlier.  Results do not depend on the chunking.
    int lead = 0;
    if (n_chunks > 1 && !ctx->exact && (ctx->chunk_lead & 1)) {
        const int cap = ns_pipe_capacity(ctx, profile, a);
        if (cap > 0 && n_utts >= 4 * cap) lead = cap;
    }
    std::vector<int32_t> first(n_chunks + 1 + (lead ? 1 : 0)), order(n_utts);
    if (lead) {
        first[0] = 0;
        SE_TRY(se_shard_plan(offsets + lead, n_utts - lead, n_chunks, first.data() + 1));
        for (int c = 1; c <= n_chunks + 1; c++) first[c] += lead;
        n_chunks++;
    } else {
        SE_TRY(se_shard_plan(offsets, n_utts, n_chunks, first.data()));
    }
    for (int u = 0; u < n_utts; u++) order[u] = u;
    for (int c = 0; c < n_chunks; c++)
        std::stable_sort(order.begin() + first[c], order.begin() + first[c + 1], [&](int32_t x, int32_t y) {
            return offsets[x + 1] - offsets[x] > offsets[y + 1] - offsets[y];
        });
    // whole-batch index arrays + per-utterance counters live on lane 0
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &a.offsets));
    const int64_t *d_aux1 = nullptr, *d_aux2 = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets2, aux1, n_utts + 1, &d_aux1));
    SE_TRY(upload_offsets(ctx, kSlotOffsets3, aux2, n_utts + 1, &d_aux2));
    if (profile == 8) { a.ceps_offsets = d_aux1; a.trace_offsets = d_aux2; } else a.block_offsets = d_aux1;
    void *d_order = nullptr, *d_counts = nullptr;
    SE_TRY(ensure_stage(ctx, kSlotOrder, sizeof(int32_t) * (size_t)n_utts, &d_order));
    SE_CUDA(cudaMemcpyAsync(d_order, order.data(), sizeof(int32_t) * (size_t)n_utts, cudaMemcpyHostToDevice, lane0->stream));
    a.order = (const int32_t *)d_order;
    if (counts) {
        if (mem == SE_MEM_DEVICE) d_counts = counts;
        else SE_TRY(ensure_stage(ctx, kSlotCounts, sizeof(int32_t) * (size_t)n_utts, &d_counts));
        set_field(a, counts_field, d_counts);
    }
    if (n_chunks > 1) {
        SE_CUDA(cudaEventRecord(lane0->done, lane0->stream));
        for (int l = 1; l < kLanes; l++) SE_CUDA(cudaStreamWaitEvent(ctx->lanes[l].stream, lane0->done, 0));
    }
    // SE_B200_TRACE=1: print the H2D / kernel / D2H completion times of every chunk (debug aid)
    struct TraceEv { cudaEvent_t e; int chunk; const char *what; };
    std::vector<TraceEv> trace;
    const bool tracing = mem == SE_MEM_HOST && std::getenv("SE_B200_TRACE") != nullptr;
    auto mark = [&](int c, const char *what) {
        if (!tracing) return;
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        cudaEventRecord(e, ctx->cur->stream);
        trace.push_back({e, c, what});
    };
    mark(-1, "start");
    // upload-ahead: every chunk's input goes up back to back on ctx->up_stream into ONE staging array that mirrors the
    // caller's; chunk c's lane waits for event c.  The upload engine is done early, and the download engine -- the
    // bottleneck of the call -- has the host link to itself for the rest of it.
    const bool ahead = mem == SE_MEM_HOST && n_chunks > 1 && (ctx->chunk_lead & 2);
    void *d_in_all = nullptr;
    if (ahead) {
        SE_TRY(ensure_stage(ctx, kSlotInAll, in_elem * (uint64_t)offsets[n_utts], &d_in_all));
        while ((int)ctx->up_events.size() < n_chunks) {
            cudaEvent_t e;
            SE_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ctx->up_events.push_back(e);
        }
        for (int c = 0; c < n_chunks; c++) {
            const uint64_t b0 = in_elem * (uint64_t)offsets[first[c]], b1 = in_elem * (uint64_t)offsets[first[c + 1]];
            if (b1 > b0)
                SE_CUDA(cudaMemcpyAsync((char *)d_in_all + b0, (const char *)samples + b0, b1 - b0, cudaMemcpyHostToDevice, ctx->up_stream));
            SE_CUDA(cudaEventRecord(ctx->up_events[c], ctx->up_stream));
            if (tracing) {
                cudaEvent_t e;
                if (cudaEventCreate(&e) == cudaSuccess) { cudaEventRecord(e, ctx->up_stream); trace.push_back({e, c, "h2d"}); }
            }
        }
    }
    int rc = SE_OK;
    for (int c = 0; c < n_chunks && rc == SE_OK; c++) {
        const int u0 = first[c], u1 = first[c + 1];
        if (u1 == u0) continue;
        ctx->cur = &ctx->lanes[n_chunks > 1 ? c % kLanes : 0];
        NsLaunch b = a;
        b.utt_begin = u0;
        b.n_utts = u1 - u0;
        b.work_counter = ctx->cur->d_work_counter;
        const int64_t lo[3] = {offsets[u0], aux1 ? aux1[u0] : 0, aux2 ? aux2[u0] : 0};
        const int64_t hi[3] = {offsets[u1], aux1 ? aux1[u1] : 0, aux2 ? aux2[u1] : 0};
        void *d_out[8] = {nullptr};
        auto body = [&]() -> int {
            if (mem == SE_MEM_HOST) {
                b.sample_base = lo[0];
                if (profile == 8) { b.ceps_base = lo[1]; b.trace_base = lo[2]; } else b.block_base = lo[1];
                b.total_blocks = hi[1] - lo[1];
                if (ahead) {
                    SE_CUDA(cudaStreamWaitEvent(ctx->cur->stream, ctx->up_events[c], 0));
                    b.samples = d_in_all;
                    b.in_shift = lo[0];
                } else {
                    SE_TRY(stage_in(ctx, kSlotIn, (const char *)samples + in_elem * (uint64_t)lo[0],
                                    in_elem * (uint64_t)(hi[0] - lo[0]), mem, &b.samples));
                }
            } else {
                b.samples = samples;
            }
            if (in_elem == sizeof(int16_t)) {      // int16 input: the byte range the kernel's bulk-async prefetch may touch
                b.bulk_lo = (const char *)b.samples;
                b.bulk_hi = b.bulk_lo + in_elem * (uint64_t)(mem == SE_MEM_HOST && !ahead ? hi[0] - lo[0] : offsets[n_utts]);
            }
            for (int k = 0; k < n_outs; k++) {
                const NsOut &o = outs[k];
                if (!o.host) continue;
                const uint64_t cnt = (uint64_t)(hi[o.index] - lo[o.index]) * (o.planes5 ? 5 : 1);
                if (mem == SE_MEM_HOST) SE_TRY(stage_out(ctx, o.slot, o.host, o.elem * cnt, mem, o.zero, &d_out[k]));
                else d_out[k] = o.host;          // device arrays are written in place (caller zeroes nothing)
                set_field(b, o.field, d_out[k]);
            }
            if (!ahead) mark(c, "h2d");
            SE_TRY(ctx->exact ? (profile == 8 ? se_exact::launch_ns_p8(ctx, b) : se_exact::launch_ns_p16(ctx, b))
                               : (profile == 8 ? launch_ns_p8(ctx, b) : launch_ns_p16(ctx, b)));
            mark(c, "kernel");
            if (mem != SE_MEM_HOST) return SE_OK;
            for (int k = 0; k < n_outs; k++) {
                const NsOut &o = outs[k];
                if (!o.host) continue;
                const uint64_t cnt = (uint64_t)(hi[o.index] - lo[o.index]);
                if (cnt == 0) continue;
                if (o.planes5 && n_chunks > 1) {
                    const uint64_t T = (uint64_t)idx[o.index][n_utts];
                    SE_CUDA(cudaMemcpy2DAsync((char *)o.host + o.elem * (uint64_t)lo[o.index], o.elem * T, d_out[k],
                                              o.elem * cnt, o.elem * cnt, 5, cudaMemcpyDeviceToHost, ctx->cur->stream));
                } else if (o.planes5) {      // one chunk: the 5 planes are contiguous on both sides
                    SE_TRY(d2h_async(ctx, o.host, d_out[k], o.elem * cnt * 5));
                } else if (n_chunks > 1) {   // pipelined big batches: straight to the caller's (pinned) array
                    SE_CUDA(cudaMemcpyAsync((char *)o.host + o.elem * (uint64_t)lo[o.index], d_out[k], o.elem * cnt,
                                            cudaMemcpyDeviceToHost, ctx->cur->stream));
                } else {
                    SE_TRY(d2h_async(ctx, (char *)o.host + o.elem * (uint64_t)lo[o.index], d_out[k], o.elem * cnt));
                }
            }
            mark(c, "d2h");
            return SE_OK;
        };
        rc = body();
    }
    ctx->cur = lane0;
    if (n_chunks > 1) {
        for (int l = 1; l < kLanes; l++) {
            cudaEventRecord(ctx->lanes[l].done, ctx->lanes[l].stream);
            cudaStreamWaitEvent(lane0->stream, ctx->lanes[l].done, 0);
        }
    }
    if (rc) {
        cudaStreamSynchronize(lane0->stream);
        ctx->bounces.clear();
        ctx->bounce_used = 0;
        for (auto &t : trace) cudaEventDestroy(t.e);
        return rc;
    }
    if (counts && mem == SE_MEM_HOST) SE_TRY(d2h_async(ctx, counts, d_counts, sizeof(int32_t) * (size_t)n_utts));
    const int rc_leave = leave(ctx, mem);
    if (tracing && !trace.empty()) {
        std::fprintf(stderr, "[se_b200 trace] %d chunks\n", n_chunks);
        for (size_t i = 1; i < trace.size(); i++) {
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, trace[0].e, trace[i].e);
            std::fprintf(stderr, "[se_b200 trace] chunk %2d %-6s done at %8.3f ms\n", trace[i].chunk, trace[i].what, ms);
        }
        for (auto &t : trace) cudaEventDestroy(t.e);
    }
    return rc_leave;
}

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
    SE_ENTER(ctx, mem, "se_etsi_denoise_batch");
    if (n_utts < 0 || (n_utts > 0 && !samples)) { set_error("se_etsi_denoise_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_etsi_denoise_batch"));
    if (ceps && !ceps_offsets) { set_error("se_etsi_denoise_batch: ceps needs ceps_offsets"); return SE_EINVAL; }
    if (vad_trace && !trace_offsets) { set_error("se_etsi_denoise_batch: vad_trace needs trace_offsets"); return SE_EINVAL; }
    if (offsets[0] != 0) { set_error("se_etsi_denoise_batch: offsets[0] must be 0"); return SE_EINVAL; }
    if (ceps) SE_TRY(check_offsets(ceps_offsets, n_utts, "se_etsi_denoise_batch(ceps_offsets)"));
    if ((ceps && ceps_offsets[0] != 0) || (vad_trace && trace_offsets[0] != 0)) {
        set_error("se_etsi_denoise_batch: ceps_offsets[0] / trace_offsets[0] must be 0");
        return SE_EINVAL;
    }
    if (vad_trace) {
        SE_TRY(check_offsets(trace_offsets, n_utts, "se_etsi_denoise_batch(trace_offsets)"));
        for (int u = 0; u < n_utts; u++)
            if (trace_offsets[u + 1] - trace_offsets[u] < (offsets[u + 1] - offsets[u]) / 80) {
                set_error("se_etsi_denoise_batch: trace_offsets too small for utterance %d", u);
                return SE_EINVAL;
            }
    }
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    a.zero_check = 1;
    NsOut outs[] = {
        {denoised, sizeof(int16_t), kIdxSample, false, false, kSlotOut, offsetof(NsLaunch, den_i16)},
        {ceps, sizeof(float) * 14, kIdxAux1, true, false, kSlotOut2, offsetof(NsLaunch, ceps)},
        {den_f32, sizeof(float), kIdxSample, false, false, kSlotOut4, offsetof(NsLaunch, den_f32)},
        {vad_trace, 1, kIdxAux2, true, false, kSlotOut5, offsetof(NsLaunch, vad_trace)},
    };
    return ns_batch_run(ctx, 8, a, samples, sizeof(int16_t), offsets, ceps ? ceps_offsets : nullptr,
                        vad_trace ? trace_offsets : nullptr, n_utts, outs, 4, n_ceps, offsetof(NsLaunch, n_ceps), mem);
}

int se_etsi_features_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                           float *features, const int64_t *feat_offsets, int32_t *n_rows, int16_t *denoised, int mem)
{
    SE_ENTER(ctx, mem, "se_etsi_features_batch");
    if (n_utts < 0 || (n_utts > 0 && (!samples || !features || !feat_offsets))) {
        set_error("se_etsi_features_batch: bad arguments");
        return SE_EINVAL;
    }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_etsi_features_batch"));
    SE_TRY(check_offsets(feat_offsets, n_utts, "se_etsi_features_batch(feat_offsets)"));
    if (offsets[0] != 0 || feat_offsets[0] != 0) { set_error("se_etsi_features_batch: offsets[0] / feat_offsets[0] must be 0"); return SE_EINVAL; }
    for (int u = 0; u < n_utts; u++)
        if (feat_offsets[u + 1] - feat_offsets[u] <
            se_etsi_num_feature_rows(offsets[u + 1] - offsets[u]) + (ctx->feature_flush ? SE_FEATURE_FLUSH_ROWS : 0)) {
            set_error("se_etsi_features_batch: feat_offsets too small for utterance %d", u);
            return SE_EINVAL;
        }
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    a.zero_check = 1;
    a.features = 1;
    a.flush = ctx->feature_flush;
    NsOut outs[] = {
        {denoised, sizeof(int16_t), kIdxSample, false, false, kSlotOut, offsetof(NsLaunch, den_i16)},
        {features, sizeof(float) * 15, kIdxAux1, true, false, kSlotOut2, offsetof(NsLaunch, ceps)},
    };
    return ns_batch_run(ctx, 8, a, samples, sizeof(int16_t), offsets, feat_offsets, nullptr, n_utts, outs, 2, n_rows,
                        offsetof(NsLaunch, n_ceps), mem);
}

int se_etsi16k_features_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                              float *features, const int64_t *feat_offsets, int32_t *n_rows, int16_t *denoised_lp,
                              const int64_t *den_offsets, int mem)
{
    SE_ENTER(ctx, mem, "se_etsi16k_features_batch");
    if (n_utts < 0 || (n_utts > 0 && (!samples || !features || !feat_offsets)) || (denoised_lp && !den_offsets)) {
        set_error("se_etsi16k_features_batch: bad arguments");
        return SE_EINVAL;
    }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_etsi16k_features_batch"));
    SE_TRY(check_offsets(feat_offsets, n_utts, "se_etsi16k_features_batch(feat_offsets)"));
    if (offsets[0] != 0 || feat_offsets[0] != 0) { set_error("se_etsi16k_features_batch: offsets[0] / feat_offsets[0] must be 0"); return SE_EINVAL; }
    if (denoised_lp) SE_TRY(check_offsets(den_offsets, n_utts, "se_etsi16k_features_batch(den_offsets)"));
    if (denoised_lp && den_offsets[0] != 0) { set_error("se_etsi16k_features_batch: den_offsets[0] must be 0"); return SE_EINVAL; }
    for (int u = 0; u < n_utts; u++) {
        const int64_t calls = se_etsi16k_num_feature_rows(offsets[u + 1] - offsets[u]);
        if (feat_offsets[u + 1] - feat_offsets[u] < calls + (ctx->feature_flush ? SE_FEATURE_FLUSH_ROWS : 0)) {
            set_error("se_etsi16k_features_batch: feat_offsets too small for utterance %d", u);
            return SE_EINVAL;
        }
        if (denoised_lp && den_offsets[u + 1] - den_offsets[u] < calls * 80) {
            set_error("se_etsi16k_features_batch: den_offsets too small for utterance %d", u);
            return SE_EINVAL;
        }
    }
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    a.zero_check = 1;
    a.features = 2;
    a.flush = ctx->feature_flush;
    NsOut outs[] = {
        {denoised_lp, sizeof(int16_t), kIdxAux2, false, false, kSlotOut, offsetof(NsLaunch, den_i16)},
        {features, sizeof(float) * 15, kIdxAux1, true, false, kSlotOut2, offsetof(NsLaunch, ceps)},
    };
    return ns_batch_run(ctx, 8, a, samples, sizeof(int16_t), offsets, feat_offsets, denoised_lp ? den_offsets : nullptr,
                        n_utts, outs, 2, n_rows, offsetof(NsLaunch, n_ceps), mem);
}

int se_p16_wiener_batch(se_ctx *ctx, const void *samples, int is_f32, const int64_t *offsets, int n_utts,
                        float *out_f32, int32_t *flags, float *gains, int32_t *n_rows,
                        const int64_t *block_offsets, int mem)
{
    SE_ENTER(ctx, mem, "se_p16_wiener_batch");
    if (n_utts < 0 || (n_utts > 0 && !samples)) { set_error("se_p16_wiener_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_p16_wiener_batch"));
    if ((flags || gains) && !block_offsets) { set_error("se_p16_wiener_batch: flags/gains need block_offsets"); return SE_EINVAL; }
    if (offsets[0] != 0) { set_error("se_p16_wiener_batch: offsets[0] must be 0"); return SE_EINVAL; }
    if (block_offsets) {
        SE_TRY(check_offsets(block_offsets, n_utts, "se_p16_wiener_batch(block_offsets)"));
        if (block_offsets[0] != 0) { set_error("se_p16_wiener_batch: block_offsets[0] must be 0"); return SE_EINVAL; }
        for (int u = 0; u < n_utts; u++)
            if (block_offsets[u + 1] - block_offsets[u] < (offsets[u + 1] - offsets[u]) / 160) {
                set_error("se_p16_wiener_batch: block_offsets too small for utterance %d", u);
                return SE_EINVAL;
            }
    }
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    a.is_f32 = is_f32 ? 1 : 0;
    a.zero_check = 2;
    a.total_blocks = block_offsets ? block_offsets[n_utts] : 0;
    NsOut outs[] = {
        {out_f32, sizeof(float), kIdxSample, false, false, kSlotOut, offsetof(NsLaunch, out_f32)},
        {flags, sizeof(int32_t), kIdxAux1, true, true, kSlotOut2, offsetof(NsLaunch, flags)},
        {gains, sizeof(float) * 25, kIdxAux1, true, false, kSlotOut3, offsetof(NsLaunch, gains)},
    };
    return ns_batch_run(ctx, 16, a, samples, is_f32 ? sizeof(float) : sizeof(int16_t), offsets, block_offsets, nullptr,
                        n_utts, outs, 3, n_rows, offsetof(NsLaunch, n_rows), mem);
}

int se_ns_spectra_batch(se_ctx *ctx, int profile, const void *samples, int is_f32, const int64_t *offsets, int n_utts,
                        float *spectra, int32_t *produced, const int64_t *block_offsets, int mem)
{
    SE_ENTER(ctx, mem, "se_ns_spectra_batch");
    if ((profile != 8 && profile != 16) || n_utts < 0 || (n_utts > 0 && (!samples || !spectra || !block_offsets)) ||
        (profile == 8 && is_f32)) {
        set_error("se_ns_spectra_batch: bad arguments");
        return SE_EINVAL;
    }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_ns_spectra_batch"));
    SE_TRY(check_offsets(block_offsets, n_utts, "se_ns_spectra_batch(block_offsets)"));
    if (offsets[0] != 0 || block_offsets[0] != 0) { set_error("se_ns_spectra_batch: offsets[0] must be 0"); return SE_EINVAL; }
    const int shift = profile == 8 ? 80 : 160, spec = profile == 8 ? 65 : 129;
    for (int u = 0; u < n_utts; u++)
        if (block_offsets[u + 1] - block_offsets[u] < (offsets[u + 1] - offsets[u]) / shift) {
            set_error("se_ns_spectra_batch: block_offsets too small for utterance %d", u);
            return SE_EINVAL;
        }
    NsLaunch a;
    std::memset(&a, 0, sizeof a);
    a.is_f32 = is_f32 ? 1 : 0;
    a.zero_check = profile == 8 ? 1 : 2;
    a.total_blocks = block_offsets[n_utts];
    // P8 indexes per-call outputs through its trace offsets (aux 2), P16 through block_offsets (aux 1)
    const int idx = profile == 8 ? kIdxAux2 : kIdxAux1;
    NsOut outs[] = {
        {spectra, sizeof(float) * 8 * (uint64_t)spec, idx, true, false, kSlotOut, offsetof(NsLaunch, spec_trace)},
        {produced, sizeof(int32_t), idx, true, false, kSlotOut2, offsetof(NsLaunch, produced)},
    };
    return ns_batch_run(ctx, profile, a, samples, is_f32 ? sizeof(float) : sizeof(int16_t), offsets,
                        profile == 8 ? nullptr : block_offsets, profile == 8 ? block_offsets : nullptr, n_utts, outs, 2,
                        nullptr, 0, mem);
}

}  // extern "C"

#include "se_api_stream.cuh"
#include "se_api_g64.cuh"
