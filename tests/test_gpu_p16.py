"""P16 (plugin etsi_denoise_mapping_func_Wiener) parity: CUDA path vs the oracle.

Tolerances: block counts / row counts / FrameCounter exact; frame-dropping flags equal except at
threshold ties; 25-band gains rel 1e-4; denoised float stream >= 60 dB SNR.
"""
import numpy as np
import pytest

from conftest import rel_err, snr_db
from speech_enhancement_b200 import synth

pytestmark = pytest.mark.gpu

GAIN_REL = 1e-4
SNR_MIN_DB = 60.0
TIE_FRACTION = 1e-3


def _oracle_utt(oracle, x):
    s = oracle.p16_stream()
    r = s.func_wiener(np.asarray(x, np.float32))
    s.close()
    return r


def _compare(ref, r, lo, hi, b0, nrows):
    nb = (hi - lo) // 160
    assert nrows == len(ref["gains"])
    assert snr_db(ref["out"][:nb * 160], r["out"][lo:lo + nb * 160]) >= SNR_MIN_DB
    assert np.all(r["out"][lo + nb * 160:hi] == 0)
    if nrows:
        e = rel_err(ref["gains"], r["gains"][b0:b0 + nrows], 1e-3)
        assert e.max() <= GAIN_REL, "gains rel err %g" % e.max()
    fl = r["flags"][:, b0:b0 + nb]
    assert np.array_equal(fl[4], ref["frame_counter"])          # integer bookkeeping: exact
    mism = 0
    for i, k in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")):
        mism += int(np.sum(fl[i] != ref[k]))
    return mism, nb


def test_c2_style_ragged_batch(ctx, oracle):
    """BASELINE config 2 shape: ragged 16 kHz utterances (non-multiples of 160), int16 input."""
    samples, offsets = synth.corpus(200, 10, 16000, min_s=0.5, max_s=2.5)
    r = ctx.p16_wiener(samples, offsets)
    tot_m = tot_b = 0
    for u in range(len(offsets) - 1):
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        ref = _oracle_utt(oracle, samples[lo:hi])
        m, nb = _compare(ref, r, lo, hi, int(r["block_offsets"][u]), int(r["n_rows"][u]))
        tot_m += m
        tot_b += nb
    assert tot_m <= max(1, int(TIE_FRACTION * tot_b * 4))


def test_zero_blocks_anywhere_and_f32_input(ctx, oracle):
    """func_Wiener skips all-zero 160-sample blocks anywhere in the stream (NoiseSup.cpp:1159-1170)."""
    x = synth.utterance(4, 16000, 160 * 80).astype(np.float32)
    x[160 * 10:160 * 13] = 0.0
    x[160 * 40:160 * 41] = 0.0
    x[:160] = 0.0
    r = ctx.p16_wiener(x, [0, len(x)])
    ref = _oracle_utt(oracle, x)
    m, nb = _compare(ref, r, 0, len(x), 0, int(r["n_rows"][0]))
    assert int(r["n_rows"][0]) == 80 - 5 - 4
    assert m <= 1


def test_short_and_empty(ctx, oracle):
    lens = [0, 159, 160, 160 * 4, 160 * 5, 160 * 6 + 1]
    xs = [synth.utterance(30 + i, 16000, n) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    r = ctx.p16_wiener(np.concatenate(xs), offsets)
    assert list(r["n_rows"]) == [0, 0, 0, 0, 1, 2]
    for u, x in enumerate(xs):
        ref = _oracle_utt(oracle, x)
        _compare(ref, r, int(offsets[u]), int(offsets[u + 1]), int(r["block_offsets"][u]), int(r["n_rows"][u]))


def test_stream_carries_state_across_pushes(ctx, oracle):
    """extract_Wiener reuses one thread instance over 64 channel files without re-init
    (aurora_etsi_test.cpp:21-141): a stream pushed twice == the oracle stream called twice."""
    a = synth.utterance(50, 16000, 160 * 40).astype(np.float32)
    b = synth.utterance(51, 16000, 160 * 30).astype(np.float32)
    s = oracle.p16_stream()
    ra, rb = s.func_wiener(a), s.func_wiener(b)
    s.close()
    g = ctx.ns_stream(16)
    ga, gb = g.push(a), g.push(b)
    g.close()
    for ref, got in ((ra, ga), (rb, gb)):
        assert len(ref["gains"]) == len(got["gains"])
        assert snr_db(ref["out"], got["out"]) >= SNR_MIN_DB
        assert rel_err(ref["gains"], got["gains"], 1e-3).max() <= GAIN_REL
        assert np.array_equal(got["flags"][4], ref["frame_counter"])


def test_pipelined_chunks_do_not_change_results(ctx):
    """SE_MEM_HOST calls are cut into overlapping H2D/kernel/D2H chunks on several streams; forcing tiny
    chunks must give bit-identical outputs (P16 and P8)."""
    samples, offsets = synth.corpus(400, 37, 16000, min_s=0.3, max_s=1.5)
    a = ctx.p16_wiener(samples, offsets)
    s8, o8 = synth.corpus(500, 41, 8000, min_s=0.3, max_s=1.5)
    a8 = ctx.etsi_denoise(s8, o8, ceps=True, den_f32=True, trace=True)
    ctx.set_pipeline(64 << 10, 3)
    try:
        b = ctx.p16_wiener(samples, offsets)
        b8 = ctx.etsi_denoise(s8, o8, ceps=True, den_f32=True, trace=True)
    finally:
        ctx.set_pipeline(512 << 20, 1024)
    for k in ("out", "flags", "gains", "n_rows"):
        assert np.array_equal(a[k], b[k]), k
    for k in ("den_i16", "ceps", "n_ceps", "den_f32", "trace"):
        assert np.array_equal(a8[k], b8[k]), k
