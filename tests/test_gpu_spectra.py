"""Spectrum-level parity (SURVEY.md §8a rows a6 FFTtoPSD, a7 PSDMean, a9 FilterCalc; north_star "spectra ... within
rel 1e-4"): se_ns_spectra_batch runs the product's fused NoiseSup kernel and returns, per block, the SPECTRUM state of
both Wiener stages.  Compared per block and bin with the oracle driven call by call (etsi/cpp/NoiseSup.c:249-303,
449-563; plugin twin aurora_etsi/NoiseSup.cpp), tolerance |a-b| <= 1e-4 max(|a|, |b|, e^-10) as SURVEY §8(c) fixes it.
"""
import numpy as np
import pytest

from speech_enhancement_b200 import synth

pytestmark = pytest.mark.gpu

SPEC_REL = 1e-4
SPEC_FLOOR = float(np.exp(-10.0))
PLANES = ("nSigSE1", "noiseSE1", "denSigSE1", "nSigSE2", "noiseSE2", "denSigSE2", "sqrtPSDMean1", "sqrtPSDMean2")


def spec_rel(ref, got):
    ref = np.asarray(ref, np.float64)
    got = np.asarray(got, np.float64)
    return np.abs(ref - got) / np.maximum(np.maximum(np.abs(ref), np.abs(got)), SPEC_FLOOR)


def _check(ctx, oracle, prof, xs):
    offsets = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
    batch = np.concatenate(xs) if xs else np.zeros(0, np.int16)
    sp, produced, bo = ctx.ns_spectra(prof, batch, offsets)
    shift = 80 if prof == 8 else 160
    worst = np.zeros(8)
    n_cmp = 0
    for u, x in enumerate(xs):
        rsp, ran = oracle.ns_spectra(prof, x)
        assert bo[u + 1] - bo[u] == len(x) // shift == len(ran)
        g = sp[bo[u]:bo[u + 1]]
        # which stage ran in which block, and the zero-gate skips: exact
        assert np.array_equal(produced[bo[u]:bo[u + 1]], ran), "utterance %d: stage bookkeeping differs" % u
        for st, bit in ((0, 2), (1, 1)):
            m = (ran & bit) != 0
            for w in (3 * st, 3 * st + 1, 3 * st + 2, 6 + st):
                if m.any():
                    e = spec_rel(rsp[m, w], g[m, w])
                    worst[w] = max(worst[w], e.max())
                    n_cmp += e.size
                # planes of a stage that did not run stay zero
                assert not g[~m, w].any()
    for w, name in enumerate(PLANES):
        assert worst[w] <= SPEC_REL, "%s: rel err %.3g" % (name, worst[w])
    return n_cmp, worst


def test_p8_spectra_ragged_batch(ctx, oracle):
    """P8: 10 s utterance + ragged ones, one with leading zero frames (u = 7), one too short for stage 2, one empty."""
    xs = [synth.utterance(0, 8000, 80000)]
    xs += [synth.utterance(u, 8000, synth.ragged_length(u, 8000, 0.5, 3.0)) for u in (3, 7, 9)]
    xs += [synth.utterance(20, 8000, 80 * 3 + 11), np.zeros(0, np.int16)]
    n, worst = _check(ctx, oracle, 8, xs)
    assert n > 8 * 65 * 1000
    print("P8 spectra: %d values compared, worst rel err per plane %s" % (n, np.array2string(worst, precision=2)))


def test_p16_spectra_ragged_batch(ctx, oracle):
    """P16 (plugin profile, truncated rfft(x,512,8) schedule, Gamma bank): 10 s + ragged + zero blocks mid-stream."""
    a = synth.utterance(5, 16000, 160000)
    b = synth.utterance(6, 16000, synth.ragged_length(6, 16000, 1.0, 3.0)).copy()
    b[160 * 30:160 * 33] = 0
    xs = [a, b, synth.utterance(8, 16000, 160 * 4 + 7)]
    n, worst = _check(ctx, oracle, 16, xs)
    assert n > 8 * 129 * 1000
    print("P16 spectra: %d values compared, worst rel err per plane %s" % (n, np.array2string(worst, precision=2)))


def test_spectra_float_input_p16(ctx, oracle):
    """The plugin's own input type is float (esti_denoise_in.inData): same state as the int16 path."""
    x = synth.utterance(12, 16000, 32000)
    a, pa, _ = ctx.ns_spectra(16, x, [0, len(x)])
    b, pb, _ = ctx.ns_spectra(16, x.astype(np.float32), [0, len(x)])
    assert np.array_equal(a, b) and np.array_equal(pa, pb)
