// G64 batch entry points + known-answer primitives of se_b200.h.  Included by se_api_batch.cuh.
#pragma once

extern "C" {

int se_subband_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                     int16_t *out, int flags, int mem)
{
    SE_ENTER(ctx, mem, "se_subband_batch");
    if (n_utts < 0 || (n_utts > 0 && (!samples || !out))) { set_error("se_subband_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_subband_batch"));
    if (offsets[0] != 0) { set_error("se_subband_batch: offsets[0] must be 0"); return SE_EINVAL; }
    const uint64_t total = (uint64_t)offsets[n_utts];
    G64Launch a;
    std::memset(&a, 0, sizeof a);
    const void *d_in = nullptr;
    void *d_out = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &a.offsets));
    SE_TRY(stage_in(ctx, kSlotIn, samples, sizeof(int16_t) * total, mem, &d_in));
    SE_TRY(stage_out(ctx, kSlotOut, out, sizeof(int16_t) * 64 * total, mem, false, &d_out));
    a.samples = (const int16_t *)d_in;
    a.utt_begin = 0;
    a.utt_count = n_utts;
    a.flags = flags;
    SE_TRY(launch_subband(ctx, a, (int16_t *)d_out));
    SE_TRY(stage_out_copy(ctx, out, d_out, sizeof(int16_t) * 64 * total, mem));
    return leave(ctx, mem);
}

int se_ibm_batch(se_ctx *ctx, const int16_t *pure, const int16_t *noise, const int64_t *offsets, int n_utts,
                 float *irm, uint64_t *ibm_bits, const int64_t *frame_offsets, int mem)
{
    SE_ENTER(ctx, mem, "se_ibm_batch");
    if (n_utts < 0 || (n_utts > 0 && (!pure || !noise || !frame_offsets))) { set_error("se_ibm_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_ibm_batch"));
    if (offsets[0] != 0) { set_error("se_ibm_batch: offsets[0] must be 0"); return SE_EINVAL; }
    for (int u = 0; u < n_utts; u++)
        if (frame_offsets[u + 1] - frame_offsets[u] < se_ibm_num_frames(offsets[u + 1] - offsets[u])) {
            set_error("se_ibm_batch: frame_offsets too small for utterance %d", u);
            return SE_EINVAL;
        }
    const uint64_t total = (uint64_t)offsets[n_utts], frames = (uint64_t)frame_offsets[n_utts];
    G64Launch a;
    std::memset(&a, 0, sizeof a);
    const void *d_p = nullptr, *d_n = nullptr;
    const int64_t *d_fo = nullptr;
    void *d_irm = nullptr, *d_bits = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &a.offsets));
    SE_TRY(upload_offsets(ctx, kSlotOffsets2, frame_offsets, n_utts + 1, &d_fo));
    SE_TRY(stage_in(ctx, kSlotIn, pure, sizeof(int16_t) * total, mem, &d_p));
    SE_TRY(stage_in(ctx, kSlotIn2, noise, sizeof(int16_t) * total, mem, &d_n));
    SE_TRY(stage_out(ctx, kSlotOut, irm, sizeof(float) * 64 * frames, mem, true, &d_irm));
    SE_TRY(stage_out(ctx, kSlotOut2, ibm_bits, sizeof(uint64_t) * frames, mem, true, &d_bits));
    a.samples = (const int16_t *)d_p;
    a.samples2 = (const int16_t *)d_n;
    a.utt_begin = 0;
    a.utt_count = n_utts;
    a.flags = SE_FLAG_EXACT;
    SE_TRY(launch_ibm(ctx, a, (float *)d_irm, (uint64_t *)d_bits, d_fo));
    SE_TRY(stage_out_copy(ctx, irm, d_irm, sizeof(float) * 64 * frames, mem));
    SE_TRY(stage_out_copy(ctx, ibm_bits, d_bits, sizeof(uint64_t) * frames, mem));
    return leave(ctx, mem);
}

int se_resynth_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                     const float *mask, const int64_t *frame_offsets, int mode, int flags,
                     int16_t *out, int mem)
{
    SE_ENTER(ctx, mem, "se_resynth_batch");
    if (n_utts < 0 || (n_utts > 0 && (!samples || !mask || !frame_offsets || !out)) || mode < 0 || mode > 2) {
        set_error("se_resynth_batch: bad arguments");
        return SE_EINVAL;
    }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_resynth_batch"));
    SE_TRY(check_offsets(frame_offsets, n_utts, "se_resynth_batch(frame_offsets)"));
    if (offsets[0] != 0) { set_error("se_resynth_batch: offsets[0] must be 0"); return SE_EINVAL; }
    for (int u = 0; u < n_utts; u++)
        if (frame_offsets[u + 1] - frame_offsets[u] != se_resynth_num_frames(offsets[u + 1] - offsets[u], mode)) {
            set_error("se_resynth_batch: utterance %d has %lld mask rows, expected %lld", u,
                      (long long)(frame_offsets[u + 1] - frame_offsets[u]),
                      (long long)se_resynth_num_frames(offsets[u + 1] - offsets[u], mode));
            return SE_EINVAL;
        }
    const uint64_t total = (uint64_t)offsets[n_utts], frames = (uint64_t)frame_offsets[n_utts];
    G64Launch a;
    std::memset(&a, 0, sizeof a);
    const void *d_in = nullptr, *d_mask = nullptr;
    const int64_t *d_fo = nullptr;
    void *d_out = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &a.offsets));
    SE_TRY(stage_in(ctx, kSlotIn, samples, sizeof(int16_t) * total, mem, &d_in));
    // The mask (1.6x the bytes of the samples) is only read by the reversed pass: host callers get it uploaded on the
    // context's upload stream while the forward pass runs.
    cudaEvent_t mask_ready = nullptr;
    if (mem == SE_MEM_HOST && frames > 0) {
        void *dm = nullptr;
        SE_TRY(ensure_stage(ctx, kSlotIn2, sizeof(float) * 64 * frames, &dm));
        if (ctx->up_events.empty()) {
            cudaEvent_t e;
            SE_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ctx->up_events.push_back(e);
        }
        SE_CUDA(cudaEventRecord(ctx->cur->done, ctx->cur->stream));         // the samples go up first, undisturbed
        SE_CUDA(cudaStreamWaitEvent(ctx->up_stream, ctx->cur->done, 0));
        SE_CUDA(cudaMemcpyAsync(dm, mask, sizeof(float) * 64 * frames, cudaMemcpyHostToDevice, ctx->up_stream));
        SE_CUDA(cudaEventRecord(ctx->up_events[0], ctx->up_stream));
        mask_ready = ctx->up_events[0];
        d_mask = dm;
    } else {
        SE_TRY(stage_in(ctx, kSlotIn2, mask, sizeof(float) * 64 * frames, mem, &d_mask));
    }
    const uint64_t out_elem = (flags & SE_FLAG_OUT_F32) ? sizeof(float) : sizeof(int16_t);
    SE_TRY(stage_out(ctx, kSlotOut, out, out_elem * total, mem, false, &d_out));
    a.samples = (const int16_t *)d_in;
    a.utt_begin = 0;
    a.utt_count = n_utts;
    a.flags = flags;
    SE_TRY(launch_resynth(ctx, a, offsets, (const float *)d_mask, frame_offsets, mode, (int16_t *)d_out, mask_ready));
    SE_TRY(stage_out_copy(ctx, out, d_out, out_elem * total, mem));
    return leave(ctx, mem);
}

// No counterpart in the reference: added (se_b200.h).  Allocates nothing:
// no cudaMalloc / cudaFree / staging / pinned arena -- device pointers in, device pointer out, fixed static scratch.
int se_resynth_noalloc(se_ctx *ctx, const int16_t *d_samples, int64_t n_samples, const float *d_mask, int mode, int flags,
                       int16_t *d_out)
{
    if (!ctx) { set_error("se_resynth_noalloc: NULL context"); return SE_EINVAL; }
    std::lock_guard<std::recursive_mutex> guard(ctx->mu);
    SE_CUDA(cudaSetDevice(ctx->device));
    if (n_samples < 0 || mode < 0 || mode > 2 || (n_samples > 0 && (!d_samples || !d_mask || !d_out))) {
        set_error("se_resynth_noalloc: bad arguments");
        return SE_EINVAL;
    }
    if (n_samples == 0) return SE_OK;
    return launch_resynth_noalloc(ctx, d_samples, n_samples, d_mask, mode, flags, d_out);
}

int se_addnoise_batch(se_ctx *ctx, const int16_t *pure, int16_t *noise, int16_t *noisy,
                      const int64_t *offsets, int n_utts, int db, int mem)
{
    SE_ENTER(ctx, mem, "se_addnoise_batch");
    if (n_utts < 0 || (n_utts > 0 && (!pure || !noise || !noisy))) { set_error("se_addnoise_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_addnoise_batch"));
    if (offsets[0] != 0) { set_error("se_addnoise_batch: offsets[0] must be 0"); return SE_EINVAL; }
    const uint64_t bytes = sizeof(int16_t) * (uint64_t)offsets[n_utts];
    const int64_t *d_off = nullptr;
    const void *d_p = nullptr, *d_nz_in = nullptr;
    void *d_noisy = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &d_off));
    SE_TRY(stage_in(ctx, kSlotIn, pure, bytes, mem, &d_p));
    SE_TRY(stage_in(ctx, kSlotIn2, noise, bytes, mem, &d_nz_in));
    SE_TRY(stage_out(ctx, kSlotOut, noisy, bytes, mem, false, &d_noisy));
    int16_t *d_nz = const_cast<int16_t *>((const int16_t *)d_nz_in);   // scaled in place
    SE_TRY(launch_addnoise(ctx, (const int16_t *)d_p, d_nz, (int16_t *)d_noisy, d_off, n_utts, db));
    SE_TRY(stage_out_copy(ctx, noise, d_nz, bytes, mem));
    SE_TRY(stage_out_copy(ctx, noisy, d_noisy, bytes, mem));
    return leave(ctx, mem);
}

int se_rfft_batch(se_ctx *ctx, float *x, int n, int m, int count, int mem)
{
    SE_ENTER(ctx, mem, "se_rfft_batch");
    if (count < 0 || (count > 0 && !x)) { set_error("se_rfft_batch: bad arguments"); return SE_EINVAL; }
    if (!((n == 256 && m == 8) || (n == 512 && m == 8))) {
        set_error("se_rfft_batch: only (256,8) and (512,8) are on the hot path");
        return SE_EUNSUPPORTED;
    }
    if (count == 0) return SE_OK;
    const uint64_t bytes = sizeof(float) * (uint64_t)n * (uint64_t)count;
    const void *d_in = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, x, bytes, mem, &d_in));
    float *d = const_cast<float *>((const float *)d_in);
    SE_TRY(ctx->exact ? se_exact::launch_rfft(ctx, d, n, m, count) : launch_rfft(ctx, d, n, m, count));
    SE_TRY(stage_out_copy(ctx, x, d, bytes, mem));
    return leave(ctx, mem);
}

int se_wave_proc_batch(se_ctx *ctx, float *frames, int n_frames, int mem)
{
    SE_ENTER(ctx, mem, "se_wave_proc_batch");
    if (n_frames < 0 || (n_frames > 0 && !frames)) { set_error("se_wave_proc_batch: bad arguments"); return SE_EINVAL; }
    if (n_frames == 0) return SE_OK;
    const uint64_t bytes = sizeof(float) * 200u * (uint64_t)n_frames;
    const void *d_in = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, frames, bytes, mem, &d_in));
    float *d = const_cast<float *>((const float *)d_in);
    SE_TRY(ctx->exact ? se_exact::launch_wave_proc(ctx, d, n_frames) : launch_wave_proc(ctx, d, n_frames));
    SE_TRY(stage_out_copy(ctx, frames, d, bytes, mem));
    return leave(ctx, mem);
}

int se_comp_ceps_batch(se_ctx *ctx, const float *frames201, int n_frames, float *coef14, int mem)
{
    SE_ENTER(ctx, mem, "se_comp_ceps_batch");
    if (n_frames < 0 || (n_frames > 0 && (!frames201 || !coef14))) { set_error("se_comp_ceps_batch: bad arguments"); return SE_EINVAL; }
    if (n_frames == 0) return SE_OK;
    const void *d_in = nullptr;
    void *d_out = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, frames201, sizeof(float) * 201u * (uint64_t)n_frames, mem, &d_in));
    SE_TRY(stage_out(ctx, kSlotOut, coef14, sizeof(float) * 14u * (uint64_t)n_frames, mem, false, &d_out));
    SE_TRY(ctx->exact ? se_exact::launch_comp_ceps(ctx, (const float *)d_in, (float *)d_out, n_frames)
                      : launch_comp_ceps(ctx, (const float *)d_in, (float *)d_out, n_frames));
    SE_TRY(stage_out_copy(ctx, coef14, d_out, sizeof(float) * 14u * (uint64_t)n_frames, mem));
    return leave(ctx, mem);
}

int se_qmf_split_batch(se_ctx *ctx, const int16_t *samples, int64_t n, int count, float *lp, float *hp, int mem)
{
    SE_ENTER(ctx, mem, "se_qmf_split_batch");
    if (count < 0 || n < 0 || n % 160 != 0 || (count > 0 && n > 0 && (!samples || !lp || !hp))) {
        set_error("se_qmf_split_batch: bad arguments (n must be a multiple of 160)");
        return SE_EINVAL;
    }
    if (count == 0 || n == 0) return SE_OK;
    const uint64_t total = (uint64_t)n * (uint64_t)count;
    const void *d_in = nullptr;
    void *d_lp = nullptr, *d_hp = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, samples, sizeof(int16_t) * total, mem, &d_in));
    SE_TRY(stage_out(ctx, kSlotOut, lp, sizeof(float) * (total / 2), mem, false, &d_lp));
    SE_TRY(stage_out(ctx, kSlotOut2, hp, sizeof(float) * (total / 2), mem, false, &d_hp));
    SE_TRY(ctx->exact ? se_exact::launch_qmf_split(ctx, (const int16_t *)d_in, n, count, (float *)d_lp, (float *)d_hp)
                      : launch_qmf_split(ctx, (const int16_t *)d_in, n, count, (float *)d_lp, (float *)d_hp));
    SE_TRY(stage_out_copy(ctx, lp, d_lp, sizeof(float) * (total / 2), mem));
    SE_TRY(stage_out_copy(ctx, hp, d_hp, sizeof(float) * (total / 2), mem));
    return leave(ctx, mem);
}

int se_gammatone_batch(se_ctx *ctx, const float *in, float *out, const int32_t *channel, int64_t n,
                       int count, int flags, int mem)
{
    SE_ENTER(ctx, mem, "se_gammatone_batch");
    if (count < 0 || n < 0 || (count > 0 && (!in || !out || !channel))) { set_error("se_gammatone_batch: bad arguments"); return SE_EINVAL; }
    if (count == 0 || n == 0) return SE_OK;
    if (mem == SE_MEM_HOST)
        for (int i = 0; i < count; i++)
            if (channel[i] < 0 || channel[i] >= kGtChannels) { set_error("se_gammatone_batch: channel[%d] out of range", i); return SE_EINVAL; }
    const uint64_t bytes = sizeof(float) * (uint64_t)n * (uint64_t)count;
    const void *d_in = nullptr, *d_ch = nullptr;
    void *d_out = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, in, bytes, mem, &d_in));
    SE_TRY(stage_in(ctx, kSlotIn2, channel, sizeof(int32_t) * (uint64_t)count, mem, &d_ch));
    SE_TRY(stage_out(ctx, kSlotOut, out, bytes, mem, false, &d_out));
    SE_TRY(launch_gammatone(ctx, (const float *)d_in, (float *)d_out, (const int32_t *)d_ch, n, count, flags));
    SE_TRY(stage_out_copy(ctx, out, d_out, bytes, mem));
    return leave(ctx, mem);
}

int se_gammatone_coef_batch(se_ctx *ctx, const float *in, float *out, const float *coef, int64_t n, int count, int flags,
                            int mem)
{
    SE_ENTER(ctx, mem, "se_gammatone_coef_batch");
    if (count < 0 || n < 0 || (count > 0 && (!in || !out || !coef))) { set_error("se_gammatone_coef_batch: bad arguments"); return SE_EINVAL; }
    if (count == 0 || n == 0) return SE_OK;
    const uint64_t bytes = sizeof(float) * (uint64_t)n * (uint64_t)count;
    const void *d_in = nullptr, *d_coef = nullptr;
    void *d_out = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, in, bytes, mem, &d_in));
    SE_TRY(stage_in(ctx, kSlotIn2, coef, sizeof(float) * 3 * (uint64_t)count, SE_MEM_HOST, &d_coef));   // coef: host array
    SE_TRY(stage_out(ctx, kSlotOut, out, bytes, mem, false, &d_out));
    SE_TRY(launch_gammatone_coef(ctx, (const float *)d_in, (float *)d_out, (const float *)d_coef, n, count, flags));
    SE_TRY(stage_out_copy(ctx, out, d_out, bytes, mem));
    return leave(ctx, mem);
}

int se_ibm_rows_batch(se_ctx *ctx, const int16_t *pure64, const int16_t *noise64, const int64_t *offsets, int n_utts,
                      float *irm, uint64_t *ibm_bits, const int64_t *frame_offsets, int mem)
{
    SE_ENTER(ctx, mem, "se_ibm_rows_batch");
    if (n_utts < 0 || (n_utts > 0 && (!pure64 || !noise64 || !frame_offsets))) { set_error("se_ibm_rows_batch: bad arguments"); return SE_EINVAL; }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_ibm_rows_batch"));
    SE_TRY(check_offsets(frame_offsets, n_utts, "se_ibm_rows_batch(frame_offsets)"));
    if (offsets[0] != 0 || frame_offsets[0] != 0) { set_error("se_ibm_rows_batch: offsets[0] must be 0"); return SE_EINVAL; }
    for (int u = 0; u < n_utts; u++)
        if (frame_offsets[u + 1] - frame_offsets[u] < se_ibm_num_frames(offsets[u + 1] - offsets[u])) {
            set_error("se_ibm_rows_batch: frame_offsets too small for utterance %d", u);
            return SE_EINVAL;
        }
    const uint64_t total = (uint64_t)offsets[n_utts], frames = (uint64_t)frame_offsets[n_utts];
    const int64_t *d_off = nullptr, *d_fo = nullptr;
    const void *d_p = nullptr, *d_n = nullptr;
    void *d_irm = nullptr, *d_bits = nullptr;
    SE_TRY(upload_offsets(ctx, kSlotOffsets, offsets, n_utts + 1, &d_off));
    SE_TRY(upload_offsets(ctx, kSlotOffsets2, frame_offsets, n_utts + 1, &d_fo));
    SE_TRY(stage_in(ctx, kSlotIn, pure64, sizeof(int16_t) * 64 * total, mem, &d_p));
    SE_TRY(stage_in(ctx, kSlotIn2, noise64, sizeof(int16_t) * 64 * total, mem, &d_n));
    SE_TRY(stage_out(ctx, kSlotOut, irm, sizeof(float) * 64 * frames, mem, true, &d_irm));
    SE_TRY(stage_out(ctx, kSlotOut2, ibm_bits, sizeof(uint64_t) * frames, mem, true, &d_bits));
    SE_TRY(launch_ibm_rows(ctx, (const int16_t *)d_p, (const int16_t *)d_n, d_off, n_utts, (float *)d_irm, (uint64_t *)d_bits,
                           d_fo, (int64_t)frames));
    SE_TRY(stage_out_copy(ctx, irm, d_irm, sizeof(float) * 64 * frames, mem));
    SE_TRY(stage_out_copy(ctx, ibm_bits, d_bits, sizeof(uint64_t) * frames, mem));
    return leave(ctx, mem);
}

int se_haircell_batch(se_ctx *ctx, const float *in, float *out, int64_t n, int count, int mem)
{
    SE_ENTER(ctx, mem, "se_haircell_batch");
    if (count < 0 || n < 0 || (count > 0 && (!in || !out))) { set_error("se_haircell_batch: bad arguments"); return SE_EINVAL; }
    if (count == 0 || n == 0) return SE_OK;
    const uint64_t bytes = sizeof(float) * (uint64_t)n * (uint64_t)count;
    const void *d_in = nullptr;
    void *d_out = nullptr;
    SE_TRY(stage_in(ctx, kSlotIn, in, bytes, mem, &d_in));
    SE_TRY(stage_out(ctx, kSlotOut, out, bytes, mem, false, &d_out));
    SE_TRY(launch_haircell(ctx, (const float *)d_in, (float *)d_out, n, count));
    SE_TRY(stage_out_copy(ctx, out, d_out, bytes, mem));
    return leave(ctx, mem);
}

// Host copies of the init-time tables (built on the host, identical to what the kernels read).
// ctx may be NULL: the tables are then built on the fly (no GPU needed — used by the CPU table tests).
int se_get_table(se_ctx *ctx, const char *name, float *out, int capacity, int *count)
{
    if (!name || !count) { set_error("se_get_table: bad arguments"); return SE_EINVAL; }
    std::vector<float> v;
    const std::string key(name);
    NsTables p8l, p16l; CcTables ccl; GtTables gtl;
    const NsTables *ns = nullptr;
    std::string field;
    if (key.rfind("p8.", 0) == 0) {
        if (ctx) ns = &ctx->h_p8; else { build_ns_tables_p8(p8l); ns = &p8l; }
        field = key.substr(3);
    } else if (key.rfind("p16.", 0) == 0) {
        if (ctx) ns = &ctx->h_p16; else { build_ns_tables_p16(p16l); ns = &p16l; }
        field = key.substr(4);
    }
    if (ns) {
        if (field == "sig_window") v = ns->sig_window;
        else if (field == "ir_window") v = ns->ir_window;
        else if (field == "bank") v = ns->bank;
        else if (field == "idct") v = ns->idct;
        else if (field == "bank_start") v.assign(ns->bank_start, ns->bank_start + kMel);
        else if (field == "bank_len") v.assign(ns->bank_len, ns->bank_len + kMel);
        else if (field == "eps") v.assign(1, ns->eps);
        else { set_error("se_get_table: unknown table %s", name); return SE_EINVAL; }
    } else if (key.rfind("cc.", 0) == 0) {
        const CcTables *cc = &ccl;
        if (ctx) cc = &ctx->h_cc; else build_cc_tables(ccl);
        field = key.substr(3);
        if (field == "hamming") v.assign(cc->hamming, cc->hamming + 100);
        else if (field == "dct") v.assign(cc->dct, cc->dct + 12 * 23);
        else if (field == "start") v.assign(cc->start, cc->start + 23);
        else if (field == "len") v.assign(cc->len, cc->len + 23);
        else if (field == "tri") v.assign(&cc->tri[0][0], &cc->tri[0][0] + 23 * 129);
        else { set_error("se_get_table: unknown table %s", name); return SE_EINVAL; }
    } else if (key.rfind("hp16.", 0) == 0) {
        Hp16Tables hpl;
        const Hp16Tables *hp = &hpl;
        if (ctx) hp = &ctx->h_hp16; else build_hp16_tables(hpl);
        field = key.substr(5);
        if (field == "dp") v.assign(hp->dp, hp->dp + 118);
        else if (field == "fb") v.assign(&hp->fb[0][0], &hp->fb[0][0] + 5 * 65);
        else if (field == "fb_start") v.assign(hp->fb_start, hp->fb_start + 5);
        else if (field == "fb_len") v.assign(hp->fb_len, hp->fb_len + 5);
        else if (field == "dct26") v.assign(hp->dct26, hp->dct26 + 12 * 26);
        else { set_error("se_get_table: unknown table %s", name); return SE_EINVAL; }
    } else if (key.rfind("gt.", 0) == 0) {
        const GtTables *gt = &gtl;
        if (ctx) gt = &ctx->h_gt; else build_gt_tables(gtl);
        field = key.substr(3);
        const float *src = nullptr;
        if (field == "cf") src = gt->cf; else if (field == "bw") src = gt->bw; else if (field == "mid") src = gt->mid;
        else if (field == "f1") src = gt->f1; else if (field == "f2") src = gt->f2; else if (field == "gain") src = gt->gain;
        if (src) v.assign(src, src + kGtChannels);
        else if (field == "hc") v.assign(gt->hc, gt->hc + 16);
        else { set_error("se_get_table: unknown table %s", name); return SE_EINVAL; }
    } else {
        set_error("se_get_table: unknown table %s", name);
        return SE_EINVAL;
    }
    *count = (int)v.size();
    if (out) {
        if (capacity < (int)v.size()) { set_error("se_get_table: capacity %d < %d", capacity, (int)v.size()); return SE_EINVAL; }
        std::memcpy(out, v.data(), sizeof(float) * v.size());
    }
    return SE_OK;
}


}  // extern "C"

// ---- extract_Wiener as one device-resident chain --------------------------------------------------------------
namespace se {
// dst[d_off[u] + i] = src[s_off[u] + i] for i < d_off[u+1] - d_off[u]: drops each utterance's len % 160 tail
__global__ void __launch_bounds__(256) compact_kernel(const int16_t *src, const int64_t *s_off, const int64_t *d_off, int16_t *dst)
{
    const int u = blockIdx.x;
    const int64_t n = d_off[u + 1] - d_off[u];
    const int16_t *s = src + s_off[u];
    int16_t *d = dst + d_off[u];
    for (int64_t i = (int64_t)blockIdx.y * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.y * blockDim.x) d[i] = s[i];
}
}  // namespace se

extern "C" {

int64_t se_extract_wiener_num_blocks(int64_t n_samples) { return n_samples > 0 ? 64 * (n_samples / 160) : 0; }

int se_extract_wiener_batch(se_ctx *ctx, const int16_t *samples, const int64_t *offsets, int n_utts,
                            float *gains, int32_t *flags, int32_t *n_rows, const int64_t *block_offsets, int mem)
{
    SE_ENTER(ctx, mem, "se_extract_wiener_batch");
    if (n_utts < 0 || (n_utts > 0 && (!samples || !block_offsets || (!gains && !flags)))) {
        set_error("se_extract_wiener_batch: bad arguments");
        return SE_EINVAL;
    }
    if (n_utts == 0) return SE_OK;
    SE_TRY(check_offsets(offsets, n_utts, "se_extract_wiener_batch"));
    if (offsets[0] != 0 || block_offsets[0] != 0) { set_error("se_extract_wiener_batch: offsets[0] / block_offsets[0] must be 0"); return SE_EINVAL; }
    for (int u = 0; u < n_utts; u++)
        if (block_offsets[u + 1] - block_offsets[u] != se_extract_wiener_num_blocks(offsets[u + 1] - offsets[u])) {
            set_error("se_extract_wiener_batch: block_offsets[%d+1] - block_offsets[%d] must be 64 * (len / 160)", u, u);
            return SE_EINVAL;
        }
    const int64_t T = block_offsets[n_utts];        // rows of the whole batch: the pitch of the caller's flags planes
    // per truncated sample: 2 B compact + 128 B sub-band rows + (100 B gains + 20 B flags) * 64 / 160
    const uint64_t kPerSample = 2 + 128 + 48 + 1;
    int u0 = 0;
    while (u0 < n_utts) {
        // group of utterances whose intermediates fit the scratch limit (at least one utterance)
        int u1 = u0;
        uint64_t tot = 0;
        while (u1 < n_utts) {
            const uint64_t l = (uint64_t)((offsets[u1 + 1] - offsets[u1]) / 160 * 160);
            if (u1 > u0 && (tot + l) * kPerSample + 4096 > ctx->scratch_limit) break;
            tot += l;
            u1++;
        }
        const int cnt = u1 - u0;
        std::vector<int64_t> s_off(cnt + 1), t_off(cnt + 1), w_off(cnt + 1), b_off(cnt + 1);
        t_off[0] = 0;
        for (int k = 0; k <= cnt; k++) s_off[k] = offsets[u0 + k] - offsets[u0];
        for (int k = 0; k < cnt; k++) t_off[k + 1] = t_off[k] + (offsets[u0 + k + 1] - offsets[u0 + k]) / 160 * 160;
        for (int k = 0; k <= cnt; k++) { w_off[k] = 64 * t_off[k]; b_off[k] = w_off[k] / 160; }
        const uint64_t ts = (uint64_t)t_off[cnt], B = (uint64_t)b_off[cnt];
        if (ts > 0) {
            auto up256 = [](uint64_t x) { return (x + 255) & ~255ull; };
            const uint64_t o_rows = up256(2 * ts), o_gains = o_rows + up256(128 * ts), o_flags = o_gains + up256(100 * B),
                           o_cnt = o_flags + up256(20 * B), need = o_cnt + up256(4 * (uint64_t)cnt);
            SE_TRY(ensure_scratch(ctx, need));
            char *sc = (char *)ctx->d_scratch;
            int16_t *d_compact = (int16_t *)sc, *d_rows = (int16_t *)(sc + o_rows);
            float *d_gains = (float *)(sc + o_gains);
            int32_t *d_flags = (int32_t *)(sc + o_flags), *d_cnt = (int32_t *)(sc + o_cnt);
            const void *d_in = nullptr;
            SE_TRY(stage_in(ctx, kSlotIn, samples + offsets[u0], sizeof(int16_t) * (uint64_t)(offsets[u1] - offsets[u0]), mem, &d_in));
            const int64_t *d_soff = nullptr, *d_toff = nullptr;
            SE_TRY(upload_offsets(ctx, 13, s_off.data(), cnt + 1, &d_soff));
            SE_TRY(upload_offsets(ctx, 14, t_off.data(), cnt + 1, &d_toff));
            se::compact_kernel<<<dim3((unsigned)cnt, 8), 256, 0, ctx->cur->stream>>>((const int16_t *)d_in, d_soff, d_toff, d_compact);
            SE_CUDA(cudaGetLastError());
            ctx->launches++;
            // subbband() of every utterance: 64 hair-cell rows of len - len % 160 samples, back to back = the 64
            // channel files extract_Wiener walks (aurora_etsi_test.cpp:90-107, tail wiped at :100-101)
            SE_TRY(se_subband_batch(ctx, d_compact, t_off.data(), cnt, d_rows, SE_FLAG_EXACT, SE_MEM_DEVICE));
            // ... through func_Wiener on ONE thread instance without re-init (:21-22): state carries over from channel to
            // channel, i.e. the 64 rows are one stream of 64 * len samples
            SE_CUDA(cudaMemsetAsync(d_gains, 0, 100 * B, ctx->cur->stream));
            SE_CUDA(cudaMemsetAsync(d_flags, 0, 20 * B, ctx->cur->stream));
            SE_TRY(se_p16_wiener_batch(ctx, d_rows, 0, w_off.data(), cnt, nullptr, d_flags, d_gains, d_cnt, b_off.data(), SE_MEM_DEVICE));
            const int64_t r0 = block_offsets[u0];
            const cudaMemcpyKind kind = mem == SE_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
            if (gains && mem == SE_MEM_HOST) SE_TRY(d2h_async(ctx, gains + 25 * r0, d_gains, 100 * B));
            else if (gains) SE_CUDA(cudaMemcpyAsync(gains + 25 * r0, d_gains, 100 * B, kind, ctx->cur->stream));
            if (flags) SE_CUDA(cudaMemcpy2DAsync(flags + r0, sizeof(int32_t) * (size_t)T, d_flags, sizeof(int32_t) * B, sizeof(int32_t) * B, 5, kind, ctx->cur->stream));
            if (n_rows) SE_CUDA(cudaMemcpyAsync(n_rows + u0, d_cnt, sizeof(int32_t) * (size_t)cnt, kind, ctx->cur->stream));
            if (u1 < n_utts) {                                                    // the next group reuses the scratch
                SE_CUDA(cudaStreamSynchronize(ctx->cur->stream));
                finish_host_copies(ctx);
            }
        } else if (n_rows) {
            if (mem == SE_MEM_HOST) std::memset(n_rows + u0, 0, sizeof(int32_t) * (size_t)cnt);
            else SE_CUDA(cudaMemsetAsync(n_rows + u0, 0, sizeof(int32_t) * (size_t)cnt, ctx->cur->stream));
        }
        u0 = u1;
    }
    return leave(ctx, mem);
}

}  // extern "C"
