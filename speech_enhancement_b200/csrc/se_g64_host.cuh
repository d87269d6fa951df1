// Host launchers of the G64 kernels.  Included by se_g64.cu (via se_g64_resynth.cuh).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

namespace se {

int upload_g64_tables(se_ctx *ctx)
{
    const GtTables &h = ctx->h_gt;
    GtDevTables d;
    for (int c = 0; c < kGtChannels; c++) { d.f1[c] = h.f1[c]; d.f2[c] = h.f2[c]; d.gain[c] = h.gain[c]; d.mid[c] = h.mid[c]; }
    for (int i = 0; i < 16; i++) d.hc[i] = h.hc[i];
    for (int c = 0; c < kGtChannels; c++) {
        const double re = (double)h.f1[c], im = (double)h.f2[c];
        const double r = std::hypot(re, im), th = std::atan2(im, re);
        d.rho[c] = (float)r;
        d.rot_c[c] = (float)(re / r);
        d.rot_s[c] = (float)(im / r);
        d.turns[c] = th / 6.283185307179586476925287;
    }
    SE_CUDA(cudaMemcpyToSymbol(c_gt, &d, sizeof d));
    float win[320], wf[160], wr[160], wb[160];
    // Hamming-320 of the SpecInfo restatement: 0.54 - 0.46 cos(2 pi k / 319)
    for (int k = 0; k < 320; k++) win[k] = (float)(0.54 - 0.46 * std::cos(6.28318530717958647692 * k / 319.0));
    for (int k = 0; k < 160; k++) {
        wf[k] = (float)h.fall_d[k];                          // first "+=" on a zero weight (extractwav.cpp:105)
        wr[k] = (float)h.rise_d[k];
        wb[k] = (float)((double)wf[k] + h.rise_d[k]);        // falling half already stored as float, then += rising
    }
    SE_CUDA(cudaMemcpyToSymbol(c_ibm_win, win, sizeof win));
    SE_CUDA(cudaMemcpyToSymbol(c_w_fall, wf, sizeof wf));
    SE_CUDA(cudaMemcpyToSymbol(c_w_rise, wr, sizeof wr));
    SE_CUDA(cudaMemcpyToSymbol(c_w_both, wb, sizeof wb));
    SE_CUDA(cudaMemcpyToSymbol(c_w_fall_d, h.fall_d, sizeof h.fall_d));
    SE_CUDA(cudaMemcpyToSymbol(c_w_rise_d, h.rise_d, sizeof h.rise_d));
    return SE_OK;
}

int launch_subband(se_ctx *ctx, const G64Launch &a, int16_t *out)
{
    if (a.utt_count <= 0) return SE_OK;
    prof_begin(ctx, SE_K_SUBBAND);
    if (a.flags & SE_FLAG_EXACT) subband_kernel<true><<<a.utt_count, 64, 0, ctx->cur->stream>>>(a.samples, a.offsets, a.utt_begin, out);
    else subband_kernel<false><<<a.utt_count, 64, 0, ctx->cur->stream>>>(a.samples, a.offsets, a.utt_begin, out);
    prof_end(ctx, SE_K_SUBBAND);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_ibm(se_ctx *ctx, const G64Launch &a, float *irm, uint64_t *bits, const int64_t *frame_offsets)
{
    if (a.utt_count <= 0) return SE_OK;
    prof_begin(ctx, SE_K_IBM);
    ibm_kernel<<<a.utt_count, 128, 0, ctx->cur->stream>>>(a.samples, a.samples2, a.offsets, a.utt_begin, irm,
                                                     reinterpret_cast<uint32_t *>(bits), frame_offsets);
    prof_end(ctx, SE_K_IBM);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

template <bool EXACT>
static void launch_bwd(se_ctx *ctx, int mode, int count, const float *Y, const GtItem *items, const float *mask, int16_t *out,
                       float *out_f)
{
    if (mode == SE_RESYNTH_IBM)
        gt_bwd_kernel<EXACT, SE_RESYNTH_IBM><<<count, 64, 0, ctx->cur->stream>>>(Y, items, mask, out, out_f);
    else
        gt_bwd_kernel<EXACT, SE_RESYNTH_ORI><<<count, 64, 0, ctx->cur->stream>>>(Y, items, mask, out, out_f);
}

// h_offsets / h_frame_offsets: host copies (work items are planned on the host and grouped against the
// scratch limit: one group = one forward + one reversed launch).
int launch_resynth(se_ctx *ctx, const G64Launch &a, const int64_t *h_offsets, const float *mask,
                   const int64_t *h_frame_offsets, int mode, int16_t *out, cudaEvent_t mask_ready)
{
    float *out_f = (a.flags & SE_FLAG_OUT_F32) ? reinterpret_cast<float *>(out) : nullptr;
    const bool exact = (a.flags & SE_FLAG_EXACT) != 0;
    const int end = a.utt_begin + a.utt_count;
    const int64_t max_rows = (int64_t)(ctx->scratch_limit / (64 * sizeof(float)));
    std::vector<GtItem> items;
    int64_t rows = 0;
    auto flush = [&]() -> int {
        if (items.empty()) return SE_OK;
        int rc = ensure_scratch(ctx, (uint64_t)rows * 64 * sizeof(float));
        if (rc) return rc;
        void *d_items = nullptr;
        rc = ensure_stage(ctx, 14, sizeof(GtItem) * items.size(), &d_items);
        if (rc) return rc;
        SE_CUDA(cudaMemcpyAsync(d_items, items.data(), sizeof(GtItem) * items.size(), cudaMemcpyHostToDevice, ctx->cur->stream));
        SE_CUDA(cudaStreamSynchronize(ctx->cur->stream));      // items is a host vector about to be reused
        float *Y = reinterpret_cast<float *>(ctx->d_scratch);
        const GtItem *it = reinterpret_cast<const GtItem *>(d_items);
        const int count = (int)items.size();
        prof_begin(ctx, SE_K_GT_FWD);
        if (exact) gt_fwd_kernel<true><<<count, 64, 0, ctx->cur->stream>>>(a.samples, it, Y);
        else gt_fwd_fast_kernel<<<count, 64, 0, ctx->cur->stream>>>(a.samples, it, Y);
        prof_end(ctx, SE_K_GT_FWD);
        if (mask_ready) SE_CUDA(cudaStreamWaitEvent(ctx->cur->stream, mask_ready, 0));   // the mask upload ran beside the forward pass
        prof_begin(ctx, SE_K_GT_BWD);
        if (exact) launch_bwd<true>(ctx, mode, count, Y, it, mask, out, out_f);
        else if (mode == SE_RESYNTH_IBM) gt_bwd_fast_kernel<SE_RESYNTH_IBM><<<count, 64, 0, ctx->cur->stream>>>(Y, it, mask, out, out_f);
        else gt_bwd_fast_kernel<SE_RESYNTH_ORI><<<count, 64, 0, ctx->cur->stream>>>(Y, it, mask, out, out_f);
        prof_end(ctx, SE_K_GT_BWD);
        SE_CUDA(cudaGetLastError());
        items.clear();
        rows = 0;
        return SE_OK;
    };
    for (int u = a.utt_begin; u < end; u++) {
        const int64_t n = h_offsets[u + 1] - h_offsets[u];
        if (n <= 0) continue;
        if (n > 0x7fffffff) { set_error("se_resynth_batch: utterance %d longer than 2^31 samples", u); return SE_EINVAL; }
        const int64_t step = exact ? n : gt_chunk_len(n, a.flags);
        for (int64_t ca = 0; ca < n; ca += step) {
            GtItem it;
            it.sample_off = h_offsets[u];
            it.frame_off = h_frame_offsets[u];
            it.frames = (int32_t)(h_frame_offsets[u + 1] - h_frame_offsets[u]);
            it.a = (int32_t)ca;
            it.b = (int32_t)std::min<int64_t>(n, ca + step);
            it.fs = exact ? 0 : (int32_t)std::max<int64_t>(0, ca - kGtWarm);
            it.fe = exact ? (int32_t)n : (int32_t)std::min<int64_t>(n, (int64_t)it.b + kGtWarm);
            const int64_t need = it.fe - it.a;       // the forward pass stores [a, fe) only
            if (rows + need > max_rows && !items.empty()) { int rc = flush(); if (rc) return rc; }
            it.y_off = rows;
            rows += need;
            items.push_back(it);
        }
    }
    return flush();
}

int launch_gammatone(se_ctx *ctx, const float *in, float *out, const int32_t *channel, int64_t n, int count, int flags)
{
    if (count <= 0) return SE_OK;
    if (flags & SE_FLAG_EXACT) gammatone_kernel<true><<<(count + 63) / 64, 64, 0, ctx->cur->stream>>>(in, out, channel, n, count);
    else gammatone_kernel<false><<<(count + 63) / 64, 64, 0, ctx->cur->stream>>>(in, out, channel, n, count);
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_gammatone_coef(se_ctx *ctx, const float *in, float *out, const float *coef, int64_t n, int count, int flags)
{
    if (count <= 0) return SE_OK;
    if (flags & SE_FLAG_EXACT) gammatone_coef_kernel<true><<<(count + 63) / 64, 64, 0, ctx->cur->stream>>>(in, out, coef, n, count);
    else gammatone_coef_kernel<false><<<(count + 63) / 64, 64, 0, ctx->cur->stream>>>(in, out, coef, n, count);
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_ibm_rows(se_ctx *ctx, const int16_t *pure64, const int16_t *noise64, const int64_t *offsets, int n_utts,
                    float *irm, uint64_t *bits, const int64_t *frame_offsets, int64_t total_frames)
{
    if (n_utts <= 0 || total_frames <= 0) return SE_OK;
    if (total_frames > 0x7fffffff) { set_error("se_ibm_rows_batch: more than 2^31 frames in one call"); return SE_EINVAL; }
    prof_begin(ctx, SE_K_IBM);
    ibm_rows_kernel<<<(unsigned)total_frames, 128, 0, ctx->cur->stream>>>(pure64, noise64, offsets, n_utts, irm,
                                                                        reinterpret_cast<uint32_t *>(bits), frame_offsets);
    prof_end(ctx, SE_K_IBM);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_haircell(se_ctx *ctx, const float *in, float *out, int64_t n, int count)
{
    if (count <= 0) return SE_OK;
    haircell_kernel<<<(count + 63) / 64, 64, 0, ctx->cur->stream>>>(in, out, n, count);
    ctx->launches++;
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

int launch_addnoise(se_ctx *ctx, const int16_t *pure, int16_t *noise, int16_t *noisy, const int64_t *offsets, int n_utts,
                    int db)
{
    if (n_utts <= 0) return SE_OK;
    void *g = nullptr;
    int rc = ensure_stage(ctx, 15, sizeof(float) * (size_t)n_utts, &g);
    if (rc) return rc;
    const float db_lin = (float)std::pow(10.0, db / 10.0);
    prof_begin(ctx, SE_K_ADDNOISE);
    addnoise_gain_kernel<<<(n_utts + 63) / 64, 64, 0, ctx->cur->stream>>>(pure, noise, offsets, n_utts, db_lin, (float *)g);
    if (n_utts > (0x7fffffff >> 6)) { set_error("se_addnoise_batch: more than 2^25 utterances in one call"); return SE_EINVAL; }
    addnoise_mix_kernel<<<(unsigned)n_utts * 64u, 256, 0, ctx->cur->stream>>>(pure, noise, noisy, offsets, (const float *)g);
    prof_end(ctx, SE_K_ADDNOISE);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

}  // namespace se
