// Host launchers of the G64 kernels.  Included by se_g64.cu (via se_g64_resynth.cuh).
#pragma once
#include <cmath>

namespace se {

int upload_g64_tables(se_ctx *ctx)
{
    const GtTables &h = ctx->h_gt;
    GtDevTables d;
    for (int c = 0; c < kGtChannels; c++) { d.f1[c] = h.f1[c]; d.f2[c] = h.f2[c]; d.gain[c] = h.gain[c]; d.mid[c] = h.mid[c]; }
    for (int i = 0; i < 16; i++) d.hc[i] = h.hc[i];
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
    ibm_kernel<<<a.utt_count, 64, 0, ctx->cur->stream>>>(a.samples, a.samples2, a.offsets, a.utt_begin, irm,
                                                     reinterpret_cast<uint32_t *>(bits), frame_offsets);
    prof_end(ctx, SE_K_IBM);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

template <bool EXACT>
static void launch_bwd(se_ctx *ctx, int mode, int count, const float *Y, const int64_t *offsets, int utt_begin,
                       int64_t base, const float *mask, const int64_t *fo, int16_t *out)
{
    if (mode == SE_RESYNTH_IBM)
        gt_bwd_kernel<EXACT, SE_RESYNTH_IBM><<<count, 64, 0, ctx->cur->stream>>>(Y, offsets, utt_begin, base, mask, fo, out);
    else
        gt_bwd_kernel<EXACT, SE_RESYNTH_ORI><<<count, 64, 0, ctx->cur->stream>>>(Y, offsets, utt_begin, base, mask, fo, out);
}

// h_offsets: host copy of the sample offsets (to chunk the batch against the scratch limit).
int launch_resynth(se_ctx *ctx, const G64Launch &a, const int64_t *h_offsets, const float *mask,
                   const int64_t *frame_offsets, int mode, int16_t *out)
{
    const bool exact = (a.flags & SE_FLAG_EXACT) != 0;
    int u = a.utt_begin;
    const int end = a.utt_begin + a.utt_count;
    const uint64_t max_samples = ctx->scratch_limit / (64 * sizeof(float));
    while (u < end) {
        int v = u;
        uint64_t samples = 0;
        while (v < end && (v == u || samples + (uint64_t)(h_offsets[v + 1] - h_offsets[v]) <= max_samples)) {
            samples += (uint64_t)(h_offsets[v + 1] - h_offsets[v]);
            v++;
        }
        if (samples > 0) {
            int rc = ensure_scratch(ctx, samples * 64 * sizeof(float));
            if (rc) return rc;
            float *Y = reinterpret_cast<float *>(ctx->d_scratch);
            const int64_t base = h_offsets[u];
            prof_begin(ctx, SE_K_GT_FWD);
            if (exact) gt_fwd_kernel<true><<<v - u, 64, 0, ctx->cur->stream>>>(a.samples, a.offsets, u, base, Y);
            else gt_fwd_kernel<false><<<v - u, 64, 0, ctx->cur->stream>>>(a.samples, a.offsets, u, base, Y);
            prof_end(ctx, SE_K_GT_FWD);
            prof_begin(ctx, SE_K_GT_BWD);
            if (exact) launch_bwd<true>(ctx, mode, v - u, Y, a.offsets, u, base, mask, frame_offsets, out);
            else launch_bwd<false>(ctx, mode, v - u, Y, a.offsets, u, base, mask, frame_offsets, out);
            prof_end(ctx, SE_K_GT_BWD);
            SE_CUDA(cudaGetLastError());
        }
        u = v;
    }
    return SE_OK;
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
    addnoise_mix_kernel<<<dim3(64, n_utts), 256, 0, ctx->cur->stream>>>(pure, noise, noisy, offsets, (const float *)g);
    prof_end(ctx, SE_K_ADDNOISE);
    SE_CUDA(cudaGetLastError());
    return SE_OK;
}

}  // namespace se
