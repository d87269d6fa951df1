"""Smallest run of every kernel, for compute-sanitizer (one tool per gpurun call).  Both NoiseSup kernels are driven
explicitly: the batch kernel (with its bulk-async input staging) and the pipelined single-utterance kernel (mbarrier
FIFOs), and the chunked host-buffer pipeline with its lead chunk and upload stream.  compute-sanitizer is closed on the
shared pool these rounds were built on ("runs under it have left GPUs needing a reset"), so the new kernels are covered by
the bit-identity tests between the two kernels and by the parity tests only; run these where the tool is available:
    compute-sanitizer --tool memcheck  python tools/sanitize_small.py
    compute-sanitizer --tool racecheck python tools/sanitize_small.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speech_enhancement_b200 import api, synth  # noqa: E402

with api.Context(0) as ctx:
    s8, o8 = synth.corpus(0, 5, 8000, min_s=0.2, max_s=0.5)
    s16, o16 = synth.corpus(10, 5, 16000, min_s=0.2, max_s=0.5)
    for mode in (1, 2):                       # 1: one warp per utterance, 2: one CTA per utterance (pipelined warps)
        ctx.set_ns_pipeline(mode)
        r = ctx.etsi_denoise(s8, o8, ceps=True, den_f32=True, trace=True)
        r = ctx.p16_wiener(s16, o16)
        r = ctx.p16_wiener(s16.astype(np.float32), o16)
        r = ctx.p16_wiener(s16[3:], np.maximum(o16 - 3, 0))        # input not 16-byte aligned: bulk copies start below it
    ctx.set_ns_pipeline(0)
    if os.environ.get("SE_SANITIZE_CHUNKS", "1") == "1":            # host-buffer pipeline: lead chunk + upload stream
        sl, ol = synth.corpus(50, 1300, 8000, min_s=0.02, max_s=0.06)
        ctx.set_pipeline(64 << 10, 64)
        ctx.etsi_denoise(sl, ol, ceps=True)
        ctx.set_pipeline(512 << 20, 1024)
    st = ctx.ns_stream(16)
    st.push(s16[:1600].astype(np.float32)); st.push(s16[1600:3200].astype(np.float32)); st.close()
    st = ctx.ns_stream(8)
    st.push(s8[:800].astype(np.float32)); st.push(s8[800:1600].astype(np.float32)); st.close()
    ctx.etsi_features(s8, o8, denoised=True)
    ctx.etsi16k_features(s16, o16, denoised=True)
    ctx.set_feature_flush(True)
    ctx.etsi_features(s8, o8)
    ctx.etsi16k_features(s16, o16)
    ctx.set_feature_flush(False)
    ctx.extract_wiener(s16, o16)
    ctx.qmf_split(s16[:1600].reshape(2, 800)[:, :640])
    ctx.rfft(np.random.default_rng(0).standard_normal((3, 256)).astype(np.float32))
    ctx.rfft(np.random.default_rng(0).standard_normal((3, 512)).astype(np.float32))
    n = 20000
    x, clean, noise = synth.utterance(3, 16000, n, components=True)
    off = np.array([0, n, n + 700], np.int64)
    xx = np.concatenate([x, x[:700]])
    for mode in (0, 1, 2):
        fo = np.concatenate([[0], np.cumsum([api.resynth_num_frames(int(k), mode) for k in np.diff(off)])])
        m = np.random.default_rng(mode).random((int(fo[-1]), 64)).astype(np.float32)
        for exact in (True, False):
            ctx.resynth(xx, off, m, fo, mode, exact=exact)
        ctx.resynth(xx, off, m, fo, mode, exact=False, latency=True)
    ctx.subband(x[:3000], [0, 3000], exact=True)
    ctx.subband(x[:3000], [0, 3000], exact=False)
    ctx.ibm(clean[:3000], noise[:3000], [0, 3000])
    ctx.addnoise(clean[:3000], noise[:3000], [0, 3000], 5)
    ctx.gammatone(np.ones((4, 500), np.float32), np.arange(4, dtype=np.int32))
print("sanitize_small: done")
