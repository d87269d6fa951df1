"""The pipelined single-utterance NoiseSup kernel (csrc/se_ns_pipe.cuh: one CTA per utterance, a block's work spread over
4-5 warps) against the batch kernel (one warp per utterance) through the same C-ABI calls: every output bit-identical, on
ragged batches with empty, sub-block and zero-prefixed utterances, with more utterances than resident CTAs, and once
against the oracle directly (the default rule picks the pipelined kernel for a lone utterance)."""
import numpy as np
import pytest

from conftest import ceps_err
from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu


def _batch(fs, shift, lens, seed):
    xs = []
    for i, n in enumerate(lens):
        x = synth.utterance(seed + i, fs, n)
        if i % 3 == 1 and n > 4 * shift:
            x[:3 * shift + 7] = 0          # leading all-zero frames: the zero-frame gate
        if i % 4 == 2 and n > 12 * shift:
            x[8 * shift:10 * shift] = 0    # zero blocks in the middle (skipped by func_Wiener, processed by DoAdvProcess)
        xs.append(x)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    return (np.concatenate(xs) if sum(lens) else np.zeros(0, np.int16)), offsets


def _both(ctx, fn):
    ctx.set_ns_pipeline(1)
    a = fn()
    ctx.set_ns_pipeline(2)
    try:
        b = fn()
    finally:
        ctx.set_ns_pipeline(0)
    return a, b


def test_p8_pipelined_equals_batch_kernel(ctx):
    lens = [8000, 80 * 37 + 11, 0, 79, 80, 240, 560, 16000, 4000, 81, 12345]
    x, off = _batch(8000, 80, lens, 700)
    a, b = _both(ctx, lambda: ctx.etsi_denoise(x, off, ceps=True, den_f32=True, trace=True))
    for k in ("den_i16", "den_f32", "ceps", "n_ceps", "trace"):
        assert np.array_equal(a[k], b[k]), k
    assert int(a["n_ceps"].sum()) > 300


def test_p16_pipelined_equals_batch_kernel(ctx):
    lens = [16000, 160 * 21 + 5, 0, 159, 160, 480, 1120, 32000, 8000, 161, 23456]
    x, off = _batch(16000, 160, lens, 720)
    a, b = _both(ctx, lambda: ctx.p16_wiener(x, off))
    for k in ("out", "flags", "gains", "n_rows"):
        assert np.array_equal(a[k], b[k]), k
    af, bf = _both(ctx, lambda: ctx.p16_wiener(x.astype(np.float32), off))      # the plugin's float entry
    for k in ("out", "flags", "gains", "n_rows"):
        assert np.array_equal(af[k], bf[k]), k
        assert np.array_equal(af[k], a[k]), k
    assert int(a["n_rows"].sum()) > 400


def test_more_utterances_than_resident_ctas(ctx):
    """1 500 short utterances: every CTA of the pipelined kernel works through several of them in a row (the FIFO
    phases carry over from one utterance to the next)."""
    rng = np.random.default_rng(5)
    lens = [int(v) for v in rng.integers(0, 2400, 1500)]
    x, off = _batch(8000, 80, lens, 4000)
    a, b = _both(ctx, lambda: ctx.etsi_denoise(x, off, ceps=True))
    for k in ("den_i16", "ceps", "n_ceps"):
        assert np.array_equal(a[k], b[k]), k
    x16, off16 = _batch(16000, 160, [2 * v for v in lens[:700]], 9000)
    a, b = _both(ctx, lambda: ctx.p16_wiener(x16, off16))
    for k in ("out", "flags", "gains", "n_rows"):
        assert np.array_equal(a[k], b[k]), k


def test_lone_utterance_default_rule_against_the_oracle(ctx, oracle):
    """One utterance per call is what the reference's drivers do: the default rule sends it to the pipelined kernel."""
    x = synth.utterance(31, 8000, 24000)
    l0 = ctx.launch_count()
    got = ctx.etsi_denoise(x, [0, len(x)], ceps=True, den_f32=True)
    assert ctx.launch_count() > l0
    ref = oracle.etsi_denoise(x, with_ceps=True)
    assert int(got["n_ceps"][0]) == len(ref["ceps"])
    d = np.abs(got["den_i16"].astype(np.int32) - ref["den_i16"].astype(np.int32))
    assert d.max() <= 1
    assert ceps_err(ref["ceps"], got["ceps"][:len(ref["ceps"])]).max() <= 1e-4


def test_host_pipeline_lead_chunk_and_upload_ahead(ctx):
    """SE_MEM_HOST calls cut the batch into chunks; the first one is sized for the pipelined kernel and every chunk's
    input is uploaded ahead on its own stream (se_ctx_set_pipeline_lead bits 1 and 2).  None of that may change a bit."""
    rng = np.random.default_rng(11)
    lens = [int(v) for v in rng.integers(0, 1800, 1400)]
    x8, off8 = _batch(8000, 80, lens, 12000)
    x16, off16 = _batch(16000, 160, [2 * v for v in lens], 15000)
    ref8 = ctx.etsi_denoise(x8, off8, ceps=True, den_f32=True, trace=True)          # one chunk (default sizes)
    ref16 = ctx.p16_wiener(x16, off16)
    try:
        ctx.set_pipeline(256 << 10, 64)                                              # many small chunks
        for mode in (0, 1, 2, 3):
            ctx.set_pipeline_lead(mode)
            got8 = ctx.etsi_denoise(x8, off8, ceps=True, den_f32=True, trace=True)
            for k in ("den_i16", "den_f32", "ceps", "n_ceps", "trace"):
                assert np.array_equal(ref8[k], got8[k]), (mode, k)
            got16 = ctx.p16_wiener(x16, off16)
            for k in ("out", "flags", "gains", "n_rows"):
                assert np.array_equal(ref16[k], got16[k]), (mode, k)
    finally:
        ctx.set_pipeline(512 << 20, 1024)
        ctx.set_pipeline_lead(3)
