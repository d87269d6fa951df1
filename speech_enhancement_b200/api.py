"""ctypes view of libse_b200.so (include/se_b200.h) for tests and bench.py.

The product is the C library; this module only marshals numpy arrays / raw device pointers into the C
ABI.  It never computes anything itself and has no fallback: if the CUDA library is missing or no GPU
is usable, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# SE_B200_LIB: load another build of the same library (kernel A/B runs: tools/build_variants.sh)
LIB_PATH = os.environ.get("SE_B200_LIB") or os.path.join(_HERE, "lib", "libse_b200.so")
DROPIN_PATH = os.path.join(_HERE, "lib", "libse_b200_dropin.so")

SE_MEM_HOST, SE_MEM_DEVICE = 0, 1
SE_FLAG_EXACT = 1
SE_FLAG_OUT_F32 = 2
SE_FLAG_LATENCY = 4
RESYNTH_IBM, RESYNTH_ORI, RESYNTH_1DNN = 0, 1, 2
K_NS_P8, K_NS_P16, K_GT_FWD, K_GT_BWD, K_SUBBAND, K_IBM, K_ADDNOISE = range(7)
KERNEL_NAMES = {K_NS_P8: "ns_p8", K_NS_P16: "ns_p16", K_GT_FWD: "gt_fwd", K_GT_BWD: "gt_bwd",
                K_SUBBAND: "subband", K_IBM: "ibm", K_ADDNOISE: "addnoise"}

_vp = C.c_void_p
_lib = None


class SeError(RuntimeError):
    pass


def lib():
    """Load libse_b200.so; fail loudly if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SeError("libse_b200.so is not built (%s); run __graft_entry__.build() — there is no "
                          "CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.se_last_error.restype = C.c_char_p
        L.se_host_alloc.restype = _vp
        L.se_host_alloc.argtypes = [C.c_uint64]
        L.se_host_free.argtypes = [_vp]
        L.se_ctx_create.argtypes = [C.c_int, C.POINTER(_vp)]
        L.se_ctx_destroy.argtypes = [_vp]
        L.se_ctx_set_stream.argtypes = [_vp, _vp]
        L.se_ctx_sync.argtypes = [_vp]
        L.se_ctx_set_scratch_limit.argtypes = [_vp, C.c_uint64]
        L.se_ctx_profiling.argtypes = [_vp, C.c_int]
        L.se_ctx_set_pipeline.argtypes = [_vp, C.c_uint64, C.c_int]
        L.se_ctx_set_pipeline_lead.argtypes = [_vp, C.c_int]
        L.se_ctx_set_feature_flush.argtypes = [_vp, C.c_int]
        L.se_ctx_set_ns_pipeline.argtypes = [_vp, C.c_int]
        L.se_ctx_set_exact.argtypes = [_vp, C.c_int]
        L.se_extract_wiener_batch.argtypes = [_vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, C.c_int]
        L.se_ctx_kernel_time.argtypes = [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.se_ctx_reset_kernel_times.argtypes = [_vp]
        L.se_ctx_launch_count.restype = C.c_int64
        L.se_ctx_launch_count.argtypes = [_vp]
        for f in ("se_etsi_num_ceps_frames", "se_etsi_num_feature_rows", "se_etsi16k_num_feature_rows", "se_ibm_num_frames"):
            getattr(L, f).restype = C.c_int64
            getattr(L, f).argtypes = [C.c_int64]
        L.se_resynth_num_frames.restype = C.c_int64
        L.se_resynth_num_frames.argtypes = [C.c_int64, C.c_int]
        L.se_etsi_denoise_batch.argtypes = [_vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int]
        L.se_wave_proc_batch.argtypes = [_vp, _vp, C.c_int, C.c_int]
        L.se_haircell_batch.argtypes = [_vp, _vp, _vp, C.c_int64, C.c_int, C.c_int]
        L.se_comp_ceps_batch.argtypes = [_vp, _vp, C.c_int, _vp, C.c_int]
        L.se_etsi_features_batch.argtypes = [_vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, C.c_int]
        L.se_etsi16k_features_batch.argtypes = [_vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_int]
        L.se_qmf_split_batch.argtypes = [_vp, _vp, C.c_int64, C.c_int, _vp, _vp, C.c_int]
        L.se_p16_wiener_batch.argtypes = [_vp, _vp, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_int]
        L.se_ns_spectra_batch.argtypes = [_vp, C.c_int, _vp, C.c_int, _vp, C.c_int, _vp, _vp, _vp, C.c_int]
        L.se_ns_stream_create.argtypes = [_vp, C.c_int, C.POINTER(_vp)]
        L.se_ns_stream_destroy.argtypes = [_vp]
        L.se_ns_stream_reset.argtypes = [_vp]
        L.se_ns_stream_set_zero_gate.argtypes = [_vp, C.c_int]
        L.se_ns_stream_push.argtypes = [_vp, _vp, C.c_int] + [_vp] * 7
        L.se_subband_batch.argtypes = [_vp, _vp, _vp, C.c_int, _vp, C.c_int, C.c_int]
        L.se_ibm_batch.argtypes = [_vp, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, C.c_int]
        L.se_resynth_batch.argtypes = [_vp, _vp, _vp, C.c_int, _vp, _vp, C.c_int, C.c_int, _vp, C.c_int]
        L.se_resynth_noalloc.argtypes = [_vp, _vp, C.c_int64, _vp, C.c_int, C.c_int, _vp]
        L.se_addnoise_batch.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int]
        L.se_rfft_batch.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int]
        L.se_gammatone_batch.argtypes = [_vp, _vp, _vp, _vp, C.c_int64, C.c_int, C.c_int, C.c_int]
        L.se_shard_plan.argtypes = [_vp, C.c_int, C.c_int, _vp]
        L.se_get_table.argtypes = [_vp, C.c_char_p, _vp, C.c_int, C.POINTER(C.c_int)]
        _lib = L
    return _lib


def _check(rc, what):
    if rc != 0:
        raise SeError("%s failed (%d): %s" % (what, rc, lib().se_last_error().decode(errors="replace")))


def _ptr(a):
    """numpy array -> host pointer; int -> raw (device) pointer; None -> NULL."""
    if a is None:
        return None
    if isinstance(a, (int, np.integer)):
        return int(a)
    return a.ctypes.data


def _i64(a):
    return np.ascontiguousarray(a, np.int64)


def get_table(name, ctx=None):
    L = lib()
    cnt = C.c_int(0)
    h = ctx.h if ctx is not None else None
    _check(L.se_get_table(h, name.encode(), None, 0, C.byref(cnt)), "se_get_table")
    out = np.zeros(cnt.value, np.float32)
    _check(L.se_get_table(h, name.encode(), out.ctypes.data, cnt.value, C.byref(cnt)), "se_get_table")
    return out


def shard_plan(offsets, n_shards):
    """Contiguous utterance ranges balanced by sample count: int32[n_shards+1] (host-side only)."""
    offsets = _i64(offsets)
    out = np.zeros(n_shards + 1, np.int32)
    _check(lib().se_shard_plan(offsets.ctypes.data, len(offsets) - 1, int(n_shards), out.ctypes.data), "se_shard_plan")
    return out


def etsi_num_ceps_frames(n):
    return int(lib().se_etsi_num_ceps_frames(int(n)))


def ibm_num_frames(n):
    return int(lib().se_ibm_num_frames(int(n)))


def resynth_num_frames(n, mode):
    return int(lib().se_resynth_num_frames(int(n), int(mode)))


class PinnedArray:
    """numpy view over se_host_alloc() memory."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(np.atleast_1d(shape).tolist())
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self.ptr = lib().se_host_alloc(max(nbytes, 16))
        if not self.ptr:
            raise SeError("se_host_alloc(%d) failed: %s" % (nbytes, lib().se_last_error().decode()))
        buf = (C.c_char * max(nbytes, 16)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().se_host_free(self.ptr)
            self.ptr = None


class Context:
    """One GPU context (se_ctx)."""

    def __init__(self, device=0):
        self.L = lib()
        h = _vp()
        _check(self.L.se_ctx_create(int(device), C.byref(h)), "se_ctx_create")
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            self.L.se_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def sync(self):
        _check(self.L.se_ctx_sync(self.h), "se_ctx_sync")

    def set_stream(self, cuda_stream):
        _check(self.L.se_ctx_set_stream(self.h, cuda_stream), "se_ctx_set_stream")

    def set_scratch_limit(self, nbytes):
        _check(self.L.se_ctx_set_scratch_limit(self.h, int(nbytes)), "se_ctx_set_scratch_limit")

    def set_pipeline_lead(self, mode):
        """bits: 1 = small first chunk on the pipelined kernel, 2 = inputs of all chunks uploaded ahead (default 3)"""
        _check(self.L.se_ctx_set_pipeline_lead(self.h, int(mode)), "se_ctx_set_pipeline_lead")

    def set_pipeline(self, chunk_bytes, min_utts):
        _check(self.L.se_ctx_set_pipeline(self.h, int(chunk_bytes), int(min_utts)), "se_ctx_set_pipeline")

    def set_exact(self, enable):
        """NoiseSup family in the reference's own arithmetic (se_ctx_set_exact)."""
        _check(self.L.se_ctx_set_exact(self.h, int(bool(enable))), "se_ctx_set_exact")

    def profiling(self, on):
        _check(self.L.se_ctx_profiling(self.h, 1 if on else 0), "se_ctx_profiling")

    def reset_kernel_times(self):
        _check(self.L.se_ctx_reset_kernel_times(self.h), "se_ctx_reset_kernel_times")

    def kernel_time(self, kid):
        ms = C.c_double(0)
        n = C.c_int64(0)
        _check(self.L.se_ctx_kernel_time(self.h, kid, C.byref(ms), C.byref(n)), "se_ctx_kernel_time")
        return ms.value, n.value

    def launch_count(self):
        return int(self.L.se_ctx_launch_count(self.h))

    # ---- raw calls (pointers may be numpy arrays or device addresses) -------------------------
    def etsi_denoise_raw(self, samples, offsets, n_utts, denoised, ceps=None, ceps_offsets=None, n_ceps=None,
                         den_f32=None, vad_trace=None, trace_offsets=None, mem=SE_MEM_HOST):
        _check(self.L.se_etsi_denoise_batch(self.h, _ptr(samples), _ptr(offsets), n_utts, _ptr(denoised), _ptr(ceps),
                                            _ptr(ceps_offsets), _ptr(n_ceps), _ptr(den_f32), _ptr(vad_trace),
                                            _ptr(trace_offsets), mem), "se_etsi_denoise_batch")

    def etsi_features_raw(self, samples, offsets, n_utts, features, feat_offsets, n_rows=None, denoised=None,
                          mem=SE_MEM_HOST):
        _check(self.L.se_etsi_features_batch(self.h, _ptr(samples), _ptr(offsets), n_utts, _ptr(features),
                                             _ptr(feat_offsets), _ptr(n_rows), _ptr(denoised), mem),
               "se_etsi_features_batch")

    def etsi16k_features_raw(self, samples, offsets, n_utts, features, feat_offsets, n_rows=None, denoised_lp=None,
                             den_offsets=None, mem=SE_MEM_HOST):
        _check(self.L.se_etsi16k_features_batch(self.h, _ptr(samples), _ptr(offsets), n_utts, _ptr(features),
                                                _ptr(feat_offsets), _ptr(n_rows), _ptr(denoised_lp), _ptr(den_offsets),
                                                mem), "se_etsi16k_features_batch")

    def p16_wiener_raw(self, samples, is_f32, offsets, n_utts, out_f32, flags, gains, n_rows, block_offsets,
                       mem=SE_MEM_HOST):
        _check(self.L.se_p16_wiener_batch(self.h, _ptr(samples), int(is_f32), _ptr(offsets), n_utts, _ptr(out_f32),
                                          _ptr(flags), _ptr(gains), _ptr(n_rows), _ptr(block_offsets), mem),
               "se_p16_wiener_batch")

    def resynth_raw(self, samples, offsets, n_utts, mask, frame_offsets, mode, flags, out, mem=SE_MEM_HOST):
        _check(self.L.se_resynth_batch(self.h, _ptr(samples), _ptr(offsets), n_utts, _ptr(mask), _ptr(frame_offsets),
                                       mode, flags, _ptr(out), mem), "se_resynth_batch")

    def resynth_noalloc_raw(self, d_samples, n_samples, d_mask, mode, flags, d_out):
        """se_resynth_noalloc: device pointers only, nothing allocated, asynchronous on the context stream."""
        _check(self.L.se_resynth_noalloc(self.h, _ptr(d_samples), int(n_samples), _ptr(d_mask), int(mode), int(flags),
                                         _ptr(d_out)), "se_resynth_noalloc")

    def subband_raw(self, samples, offsets, n_utts, out, flags, mem=SE_MEM_HOST):
        _check(self.L.se_subband_batch(self.h, _ptr(samples), _ptr(offsets), n_utts, _ptr(out), flags, mem),
               "se_subband_batch")

    def ibm_raw(self, pure, noise, offsets, n_utts, irm, bits, frame_offsets, mem=SE_MEM_HOST):
        _check(self.L.se_ibm_batch(self.h, _ptr(pure), _ptr(noise), _ptr(offsets), n_utts, _ptr(irm), _ptr(bits),
                                   _ptr(frame_offsets), mem), "se_ibm_batch")

    # ---- numpy conveniences --------------------------------------------------------------------
    def etsi_denoise(self, samples, offsets, ceps=True, den_f32=False, trace=False):
        samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        lens = np.diff(offsets)
        out = {"den_i16": np.zeros(len(samples), np.int16)}
        co = to = None
        if ceps:
            co = np.zeros(nu + 1, np.int64)
            co[1:] = np.cumsum(np.maximum(lens // 80 - 6, 0))
            out["ceps"] = np.zeros((max(int(co[-1]), 1), 14), np.float32)
            out["ceps_offsets"] = co
            out["n_ceps"] = np.zeros(max(nu, 1), np.int32)
        if den_f32:
            out["den_f32"] = np.zeros(len(samples), np.float32)
        if trace:
            to = np.zeros(nu + 1, np.int64)
            to[1:] = np.cumsum(lens // 80)
            out["trace"] = np.zeros(max(int(to[-1]), 1), np.uint8)
            out["trace_offsets"] = to
        self.etsi_denoise_raw(samples, offsets, nu, out["den_i16"], out.get("ceps"), co, out.get("n_ceps"),
                              out.get("den_f32"), out.get("trace"), to)
        return out

    def haircell(self, x):
        x = np.ascontiguousarray(x, np.float32)
        x2 = x.reshape(-1, x.shape[-1])
        out = np.zeros_like(x2)
        _check(self.L.se_haircell_batch(self.h, _ptr(x2), _ptr(out), x2.shape[1], x2.shape[0], SE_MEM_HOST), "se_haircell_batch")
        return out.reshape(x.shape)

    def comp_ceps(self, frames201):
        f = np.ascontiguousarray(frames201, np.float32).reshape(-1, 201)
        out = np.zeros((len(f), 14), np.float32)
        _check(self.L.se_comp_ceps_batch(self.h, _ptr(f), len(f), _ptr(out), SE_MEM_HOST), "se_comp_ceps_batch")
        return out

    def wave_proc(self, frames):
        f = np.ascontiguousarray(frames, np.float32).reshape(-1, 200).copy()
        _check(self.L.se_wave_proc_batch(self.h, _ptr(f), len(f), SE_MEM_HOST), "se_wave_proc_batch")
        return f

    def etsi_features(self, samples, offsets, denoised=False):
        """Full ETSI feature chain: list of per-utterance [rows][15] arrays (+ int16 denoised stream)."""
        samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        fo = np.zeros(nu + 1, np.int64)
        fo[1:] = np.cumsum(np.diff(offsets) // 80 + getattr(self, "_flush_rows", 0))
        feat = np.zeros((max(int(fo[-1]), 1), 15), np.float32)
        rows = np.zeros(max(nu, 1), np.int32)
        den = np.zeros(len(samples), np.int16) if denoised else None
        self.etsi_features_raw(samples, offsets, nu, feat, fo, rows, den)
        per_utt = [feat[fo[u]:fo[u] + rows[u]] for u in range(nu)]
        return (per_utt, den) if denoised else per_utt

    def set_ns_pipeline(self, mode):
        """0: kernel chosen by batch size, 1: one warp per utterance, 2: one CTA per utterance (pipelined warps)."""
        _check(self.L.se_ctx_set_ns_pipeline(self.h, int(mode)), "se_ctx_set_ns_pipeline")

    def set_feature_flush(self, enable):
        """FlushAdvProcess at the end of every utterance of the feature-chain calls (6 more rows)."""
        _check(self.L.se_ctx_set_feature_flush(self.h, int(bool(enable))), "se_ctx_set_feature_flush")
        self._flush_rows = 6 if enable else 0

    def etsi16k_features(self, samples, offsets, denoised=False):
        """ETSI 16 kHz mode: per-utterance [rows][15] arrays (+ per-utterance int16 low-band denoised streams)."""
        samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        calls = np.diff(offsets) // 160
        fo = np.zeros(nu + 1, np.int64)
        fo[1:] = np.cumsum(calls + getattr(self, "_flush_rows", 0))
        do = np.zeros(nu + 1, np.int64)
        do[1:] = np.cumsum(calls) * 80
        feat = np.zeros((max(int(fo[-1]), 1), 15), np.float32)
        rows = np.zeros(max(nu, 1), np.int32)
        den = np.zeros(max(int(do[-1]), 1), np.int16) if denoised else None
        self.etsi16k_features_raw(samples, offsets, nu, feat, fo, rows, den, do if denoised else None)
        per_utt = [feat[fo[u]:fo[u] + rows[u]] for u in range(nu)]
        return (per_utt, [den[do[u]:do[u + 1]] for u in range(nu)]) if denoised else per_utt

    def qmf_split(self, x):
        """Do16kProcessing over [count][n] int16 signals (n % 160 == 0): (lp, hp) float32 [count][n/2]."""
        x = np.ascontiguousarray(x, np.int16)
        x2 = x.reshape(-1, x.shape[-1])
        lp = np.zeros((x2.shape[0], x2.shape[1] // 2), np.float32)
        hp = np.zeros_like(lp)
        _check(self.L.se_qmf_split_batch(self.h, _ptr(x2), x2.shape[1], x2.shape[0], _ptr(lp), _ptr(hp), SE_MEM_HOST),
               "se_qmf_split_batch")
        return lp, hp

    def p16_wiener(self, samples, offsets):
        samples = np.ascontiguousarray(samples)
        is_f32 = samples.dtype == np.float32
        if not is_f32:
            samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        bo = np.zeros(nu + 1, np.int64)
        bo[1:] = np.cumsum(np.diff(offsets) // 160)
        tb = max(int(bo[-1]), 1)
        out = dict(out=np.zeros(len(samples), np.float32), flags=np.zeros((5, tb), np.int32),
                   gains=np.zeros((tb, 25), np.float32), n_rows=np.zeros(max(nu, 1), np.int32), block_offsets=bo)
        if bo[-1] == 0:
            out["flags"] = np.zeros((5, 0), np.int32)
        self.p16_wiener_raw(samples, is_f32, offsets, nu, out["out"], out["flags"] if bo[-1] else None,
                            out["gains"] if bo[-1] else None, out["n_rows"], bo)
        return out

    def ns_spectra(self, profile, samples, offsets):
        """Per-block spectral state (se_ns_spectra_batch): spectra[block][8][spec], produced[block], block_offsets."""
        samples = np.ascontiguousarray(samples)
        is_f32 = samples.dtype == np.float32
        if not is_f32:
            samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        shift, spec = (80, 65) if profile == 8 else (160, 129)
        bo = np.zeros(nu + 1, np.int64)
        bo[1:] = np.cumsum(np.diff(offsets) // shift)
        tb = max(int(bo[-1]), 1)
        sp = np.zeros((tb, 8, spec), np.float32)
        produced = np.zeros(tb, np.int32)
        _check(self.L.se_ns_spectra_batch(self.h, profile, _ptr(samples), int(is_f32), _ptr(offsets), nu, _ptr(sp),
                                          _ptr(produced), _ptr(bo), SE_MEM_HOST), "se_ns_spectra_batch")
        return sp[:int(bo[-1])], produced[:int(bo[-1])], bo

    def subband(self, samples, offsets, exact=True):
        samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        out = np.zeros(64 * len(samples), np.int16)
        self.subband_raw(samples, offsets, len(offsets) - 1, out, SE_FLAG_EXACT if exact else 0)
        return out

    def ibm(self, pure, noise, offsets):
        pure = np.ascontiguousarray(pure, np.int16)
        noise = np.ascontiguousarray(noise, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        fo = np.zeros(nu + 1, np.int64)
        fo[1:] = np.cumsum([ibm_num_frames(n) for n in np.diff(offsets)])
        nf = max(int(fo[-1]), 1)
        irm = np.zeros((nf, 64), np.float32)
        bits = np.zeros(nf, np.uint64)
        self.ibm_raw(pure, noise, offsets, nu, irm, bits, fo)
        return irm[:int(fo[-1])], bits[:int(fo[-1])], fo

    def extract_wiener(self, samples, offsets):
        """extract_Wiener over a batch: per-utterance (gains[64*(len//160)][25], flags[5][64*(len//160)], n_rows)."""
        samples = np.ascontiguousarray(samples, np.int16)
        offsets = _i64(offsets)
        nu = len(offsets) - 1
        bo = np.zeros(nu + 1, np.int64)
        bo[1:] = np.cumsum(64 * (np.diff(offsets) // 160))
        tb = max(int(bo[-1]), 1)
        gains = np.zeros((tb, 25), np.float32)
        flags = np.zeros((5, tb), np.int32)
        rows = np.zeros(max(nu, 1), np.int32)
        _check(self.L.se_extract_wiener_batch(self.h, _ptr(samples), _ptr(offsets), nu, _ptr(gains), _ptr(flags),
                                              _ptr(rows), _ptr(bo), SE_MEM_HOST), "se_extract_wiener_batch")
        return [(gains[bo[u]:bo[u + 1]], flags[:, bo[u]:bo[u + 1]], int(rows[u])) for u in range(nu)]

    def resynth(self, samples, offsets, mask, frame_offsets, mode, exact=True, f32=False, latency=False):
        samples = np.ascontiguousarray(samples, np.int16)
        mask = np.ascontiguousarray(mask, np.float32)
        out = np.zeros(len(samples), np.float32 if f32 else np.int16)
        self.resynth_raw(samples, _i64(offsets), len(offsets) - 1, mask, _i64(frame_offsets), mode,
                         (SE_FLAG_EXACT if exact else 0) | (SE_FLAG_OUT_F32 if f32 else 0) |
                         (SE_FLAG_LATENCY if latency else 0), out)
        return out

    def addnoise(self, pure, noise, offsets, db):
        pure = np.ascontiguousarray(pure, np.int16)
        noise = np.ascontiguousarray(noise, np.int16).copy()
        noisy = np.zeros_like(pure)
        offsets = _i64(offsets)
        _check(self.L.se_addnoise_batch(self.h, pure.ctypes.data, noise.ctypes.data, noisy.ctypes.data,
                                        offsets.ctypes.data, len(offsets) - 1, int(db), SE_MEM_HOST), "se_addnoise_batch")
        return noise, noisy

    def rfft(self, x, m=8):
        x = np.ascontiguousarray(x, np.float32).copy()
        x2 = x.reshape(-1, x.shape[-1])
        _check(self.L.se_rfft_batch(self.h, x2.ctypes.data, x2.shape[1], m, x2.shape[0], SE_MEM_HOST), "se_rfft_batch")
        return x

    def gammatone(self, x, channels, exact=True):
        x = np.ascontiguousarray(x, np.float32)
        x2 = x.reshape(-1, x.shape[-1])
        ch = np.ascontiguousarray(channels, np.int32)
        out = np.zeros_like(x2)
        _check(self.L.se_gammatone_batch(self.h, x2.ctypes.data, out.ctypes.data, ch.ctypes.data, x2.shape[1],
                                         x2.shape[0], SE_FLAG_EXACT if exact else 0, SE_MEM_HOST), "se_gammatone_batch")
        return out.reshape(x.shape)

    def ns_stream(self, profile):
        return NsStream(self, profile)


class NsStream:
    """Persistent NoiseSup stream (se_ns_stream)."""

    def __init__(self, ctx, profile):
        self.ctx = ctx
        self.profile = profile
        self.shift = 80 if profile == 8 else 160
        h = _vp()
        _check(ctx.L.se_ns_stream_create(ctx.h, profile, C.byref(h)), "se_ns_stream_create")
        self.h = h

    def close(self):
        if self.h:
            self.ctx.L.se_ns_stream_destroy(self.h)
            self.h = None

    def reset(self):
        _check(self.ctx.L.se_ns_stream_reset(self.h), "se_ns_stream_reset")

    def set_zero_gate(self, mode):
        _check(self.ctx.L.se_ns_stream_set_zero_gate(self.h, mode), "se_ns_stream_set_zero_gate")

    def push(self, x):
        x = np.ascontiguousarray(x, np.float32)
        nb = len(x) // self.shift
        out = np.zeros(max(nb, 1) * self.shift, np.float32)
        produced = np.zeros(max(nb, 1), np.int32)
        flags = np.zeros((5, max(nb, 1)), np.int32)
        gains = np.zeros((max(nb, 1), 25), np.float32)
        ceps = np.zeros((max(nb, 1), 14), np.float32)
        n_rows = C.c_int32(0)
        n_ceps = C.c_int32(0)
        p16 = self.profile == 16
        _check(self.ctx.L.se_ns_stream_push(self.h, x.ctypes.data, nb, out.ctypes.data, produced.ctypes.data,
                                            flags.ctypes.data if p16 else None, gains.ctypes.data if p16 else None,
                                            C.addressof(n_rows), None if p16 else ceps.ctypes.data,
                                            C.addressof(n_ceps)), "se_ns_stream_push")
        return dict(out=out[:nb * self.shift], produced=produced[:nb], flags=flags[:, :nb],
                    gains=gains[:n_rows.value], ceps=ceps[:n_ceps.value])
