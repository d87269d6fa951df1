"""P8 (ETSI NoiseSup + CompCeps) parity: CUDA path through the C ABI vs the oracle.

Tolerances (north_star): frame counts / indexing exact; VAD flags equal except at threshold ties
(bounded below); cepstra rel 1e-4 on the frame's scale (conftest.ceps_err); denoised waveform >= 60 dB SNR vs the oracle's float
stream and int16 within +-1 LSB.
"""
import numpy as np
import pytest

from conftest import ceps_err, snr_db
from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu

CEPS_REL = 1e-4      # north_star: spectra/cepstra within rel 1e-4
SNR_MIN_DB = 60.0
TIE_FRACTION = 1e-3  # "expect << 0.1 % of frames"


def _compare_utt(o, x, r, lo, hi, co, nc, to):
    ref = o.etsi_denoise(x, with_ceps=True, trace=True)
    calls = ref["calls"]
    assert calls == len(x) // 80
    n_body = calls * 80
    # frame counts: exact
    assert nc == len(ref["ceps"])
    den = r["den_i16"][lo:hi]
    assert np.all(den[n_body:] == 0)
    d = np.abs(den[:n_body].astype(np.int32) - ref["den_i16"][:n_body].astype(np.int32))
    assert d.max(initial=0) <= 1, "int16 denoised differs by more than 1 LSB"
    assert snr_db(ref["den_f32"][:n_body], r["den_f32"][lo:lo + n_body]) >= SNR_MIN_DB
    if nc:
        e = ceps_err(ref["ceps"], r["ceps"][co:co + nc])
        assert e.max() <= CEPS_REL, "cepstra rel err %g" % e.max()
    # VAD / frame-dropping flags
    tr = r["trace"][to:to + calls]
    info = ref["info"]
    skipped = info["produced"] < 0
    assert np.array_equal((tr & 128) != 0, skipped)
    assert np.array_equal((tr & 64) != 0, info["produced"] > 0)
    assert np.array_equal((tr & 32) != 0, info["stage0_ran"] > 0)
    ran = info["stage0_ran"] > 0
    mism = 0
    for bit, key in ((1, "flag_vad"), (2, "sp_var"), (4, "sp_spec"), (8, "sp_mel"), (16, "sp_vadns")):
        mism += int(np.sum(((tr[ran] & bit) != 0) != (info[key][ran] != 0)))
    return mism, int(ran.sum())


def test_c1_single_10s_utterance(ctx, oracle):
    """BASELINE config 1: one 10 s 8 kHz synthetic noisy utterance."""
    x = synth.utterance(0, 8000, 80000)
    r = ctx.etsi_denoise(x, [0, len(x)], ceps=True, den_f32=True, trace=True)
    assert int(r["n_ceps"][0]) == 80000 // 80 - 6
    mism, ran = _compare_utt(oracle, x, r, 0, len(x), 0, int(r["n_ceps"][0]), 0)
    assert mism <= max(1, int(TIE_FRACTION * ran * 5))


def test_ragged_batch(ctx, oracle):
    """Ragged lengths (tail drop), leading-zero utterance (u=7), very short and empty utterances."""
    lens = [synth.ragged_length(u, 8000, 0.5, 3.0) for u in range(12)]
    lens[3] = 0
    lens[4] = 79
    lens[5] = 80 * 5 + 3      # stage latency: nothing produced
    lens[6] = 80 * 7          # exactly one cepstral frame
    xs = [synth.utterance(u, 8000, n) for u, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    samples = np.concatenate(xs) if sum(lens) else np.zeros(0, np.int16)
    r = ctx.etsi_denoise(samples, offsets, ceps=True, den_f32=True, trace=True)
    tot_m = tot_r = 0
    for u, x in enumerate(xs):
        m, ran = _compare_utt(oracle, x, r, int(offsets[u]), int(offsets[u + 1]), int(r["ceps_offsets"][u]),
                              int(r["n_ceps"][u]), int(r["trace_offsets"][u]))
        tot_m += m
        tot_r += ran
    assert tot_m <= max(1, int(TIE_FRACTION * tot_r * 5))


def test_all_zero_and_zero_lead(ctx, oracle):
    z = np.zeros(80 * 20, np.int16)
    x = np.concatenate([np.zeros(80 * 3, np.int16), synth.utterance(2, 8000, 80 * 60)])
    samples = np.concatenate([z, x])
    offsets = np.array([0, len(z), len(z) + len(x)], np.int64)
    r = ctx.etsi_denoise(samples, offsets, ceps=True, den_f32=True, trace=True)
    assert int(r["n_ceps"][0]) == 0 and np.all(r["den_i16"][:len(z)] == 0)
    assert np.all((r["trace"][:20] & 128) != 0)
    _compare_utt(oracle, x, r, len(z), len(z) + len(x), int(r["ceps_offsets"][1]), int(r["n_ceps"][1]),
                 int(r["trace_offsets"][1]))
    assert int(r["n_ceps"][1]) == 60 - 6       # leading zero frames do not count (ParmInterface.c:244-250)


def test_shard_invariance(ctx):
    """Per-utterance outputs are bit-identical regardless of how the batch is split (multi-GPU sharding)."""
    samples, offsets = synth.corpus(100, 16, 8000, min_s=0.5, max_s=2.0)
    full = ctx.etsi_denoise(samples, offsets, ceps=True)
    for lo, hi in ((0, 5), (5, 16)):
        sub = ctx.etsi_denoise(samples[offsets[lo]:offsets[hi]], offsets[lo:hi + 1] - offsets[lo], ceps=True)
        assert np.array_equal(sub["den_i16"], full["den_i16"][offsets[lo]:offsets[hi]])
        c0, c1 = int(full["ceps_offsets"][lo]), int(full["ceps_offsets"][hi])
        assert np.array_equal(sub["ceps"][:c1 - c0], full["ceps"][c0:c1])


def test_rfft_known_answers(ctx, oracle):
    rng = np.random.default_rng(7)
    x = rng.standard_normal((64, 256)).astype(np.float32) * 1000
    y = ctx.rfft(x, 8)
    for i in range(len(x)):
        ref = oracle.rfft(x[i], 8)
        assert np.max(np.abs(ref - y[i])) <= 1e-4 * np.max(np.abs(ref))
    x = rng.standard_normal((64, 512)).astype(np.float32) * 1000
    y = ctx.rfft(x, 8)          # the plugin's truncated schedule
    for i in range(len(x)):
        ref = oracle.rfft(x[i], 8, cxx=True)
        assert np.max(np.abs(ref - y[i])) <= 1e-4 * np.max(np.abs(ref))


def test_stream_matches_batch(ctx):
    """DoNoiseSup-style persistent stream pushed in pieces == one batch call."""
    x = synth.utterance(11, 8000, 80 * 200)
    full = ctx.etsi_denoise(x, [0, len(x)], ceps=True, den_f32=True)
    s = ctx.ns_stream(8)
    s.set_zero_gate(1)
    outs, ceps = [], []
    for lo, hi in ((0, 1), (1, 2), (2, 7), (7, 50), (50, 200)):
        r = s.push(x[lo * 80:hi * 80].astype(np.float32))
        outs.append(r["out"])
        ceps.append(r["ceps"])
    s.close()
    assert np.array_equal(np.concatenate(outs), full["den_f32"])
    assert np.array_equal(np.concatenate(ceps), full["ceps"][:int(full["n_ceps"][0])])
