"""the ETSI 16 kHz mode (etsi/cpp/16kHzProcessing.c and its hooks in NoiseSup.c / CompCeps.c):
QMF split -> NoiseSup on the low band + spectral subtraction / coding of three 4-8 kHz mel bands -> WaveProc ->
26-band CompCeps -> PostProc -> VAD, against the oracle (pinned bit-for-bit on the reference build by
tests/test_oracle_golden.py::test_etsi16k_*).

Bars as for the 8 kHz chain (tests/test_gpu_features.py): row counts and VAD flags identical, low-band int16
stream within 1 LSB, cepstra within 1e-4 of the frame's scale except WaveProc peak flips, bounded by
conftest.check_feature_envelope; the exact arithmetic mode is bit-identical on every row (tests/test_gpu_exact.py).
"""
import numpy as np
import pytest

from conftest import FEAT_OUTLIER_TOL, ceps_err, check_feature_envelope
from speech_enhancement_b200 import synth

pytestmark = pytest.mark.gpu

OUTLIER_TOL = FEAT_OUTLIER_TOL


def _batch(lens, zeros, seed=700):
    xs = []
    for i, (n, z) in enumerate(zip(lens, zeros)):
        x = synth.utterance(seed + i, 16000, n)
        x[:160 * z] = 0
        xs.append(x)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    return xs, offsets


def test_qmf_split_known_answer(ctx, oracle):
    """Do16kProcessing alone: both half-bands within float rounding of the sequential 118-tap sums."""
    rng = np.random.default_rng(5)
    n = 160 * 40
    t = np.arange(n)
    sigs = [synth.utterance(11, 16000, n),
            (rng.standard_normal(n) * 9000).clip(-32768, 32767).astype(np.int16),
            (12000 * np.sin(2 * np.pi * 6000 * t / 16000.0)).astype(np.int16),       # lands in the high band
            (12000 * np.sin(2 * np.pi * 1000 * t / 16000.0)).astype(np.int16),       # lands in the low band
            rng.integers(-32768, 32768, n).astype(np.int16),
            np.zeros(n, np.int16)]
    lp, hp = ctx.qmf_split(np.stack(sigs))
    for k, x in enumerate(sigs):
        rl, rh = oracle.qmf_split(x)
        scale = max(float(np.abs(x).max()), 1.0)
        assert np.abs(lp[k] - rl).max() <= 2e-6 * scale, k
        assert np.abs(hp[k] - rh).max() <= 2e-6 * scale, k
    # a 6 kHz tone must come out in the high band, a 1 kHz tone in the low band
    assert np.abs(hp[2][200:]).max() > 100 * np.abs(lp[2][200:]).max()
    assert np.abs(lp[3][200:]).max() > 100 * np.abs(hp[3][200:]).max()


def test_features16k_vs_oracle(ctx, oracle):
    lens = [32000, 48000, 16000, 2000, 800, 24691, 1920, 0, 159, 160 * 13, 80000]
    zeros = [0, 3, 7, 0, 0, 1, 0, 0, 0, 0, 2]
    xs, offsets = _batch(lens, zeros)
    got, den = ctx.etsi16k_features(np.concatenate(xs), offsets, denoised=True)
    n_all = 0
    for u, x in enumerate(xs):
        ref, ref_den = oracle.etsi_features_16k(x)
        g = got[u]
        assert g.shape == ref.shape, "utterance %d: %s vs %s" % (u, g.shape, ref.shape)
        assert den[u].shape == ref_den.shape
        d = np.abs(den[u].astype(np.int32) - ref_den.astype(np.int32))
        assert d.size == 0 or d.max() <= 1, "utterance %d: low-band stream off by %d" % (u, d.max())
        if len(ref) == 0:
            continue
        assert np.array_equal(g[:, 14], ref[:, 14]), "utterance %d: VAD flags differ" % u
        assert not g[:zeros[u]].any()
        err = ceps_err(ref[:, :14], g[:, :14]).max(axis=1)
        n_all += len(err)
        assert err.max() <= OUTLIER_TOL, "utterance %d frame %d: %g" % (u, int(err.argmax()), err.max())
    assert n_all > 1000


def test_features16k_within_the_reference_own_envelope(ctx, oracle):
    """40 x 5 s of 16 kHz audio (19 520 rows) against the envelope of conftest.check_feature_envelope."""
    xs = [synth.utterance(7000 + i, 16000, 16000 * 5) for i in range(40)]
    offsets = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
    got = ctx.etsi16k_features(np.concatenate(xs), offsets)
    errs = []
    for u, x in enumerate(xs):
        ref, _ = oracle.etsi_features_16k(x)
        assert got[u].shape == ref.shape and np.array_equal(got[u][:, 14], ref[:, 14])
        errs.append(ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1))
    n_tail, n_flip = check_feature_envelope(np.concatenate(errs))
    print("16 kHz mode, fast mode: %d frames above 1e-4, %d above 1e-3 of %d" % (n_tail, n_flip, sum(map(len, errs))))


def test_features16k_use_the_high_band(ctx):
    """The high band must matter: replacing 4-8 kHz content changes c0/logE and the cepstra, the low-band stream not."""
    n = 48000
    x = synth.utterance(31, 16000, n).astype(np.float64)
    t = np.arange(n)
    y = x + 2500 * np.sin(2 * np.pi * 6100 * t / 16000.0) * (1 + np.sin(2 * np.pi * 3 * t / 16000.0))
    a, da = ctx.etsi16k_features(x.astype(np.int16), [0, n], denoised=True)
    b, db = ctx.etsi16k_features(np.clip(y, -32768, 32767).astype(np.int16), [0, n], denoised=True)
    assert a[0].shape == b[0].shape
    assert np.abs(a[0][50:, 13] - b[0][50:, 13]).mean() > 0.1          # logE carries the HP energy
    assert np.abs(a[0][50:, :12] - b[0][50:, :12]).max() > 0.5
    # the QMF stop band keeps a 6.1 kHz tone out of the low band: the denoised low-band stream barely moves
    assert np.abs(da[0].astype(np.int32) - db[0].astype(np.int32)).max() <= 40


def test_features16k_shards_and_chunks(ctx):
    """No cross-utterance state: any split, and the pipelined host path with tiny chunks, give identical rows."""
    samples, offsets = synth.corpus(1700, 40, 16000, min_s=0.3, max_s=1.5)
    whole, den = ctx.etsi16k_features(samples, offsets, denoised=True)
    for u in (0, 7, 39):
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        alone, d1 = ctx.etsi16k_features(samples[lo:hi], [0, hi - lo], denoised=True)
        assert np.array_equal(alone[0], whole[u])
        assert np.array_equal(d1[0], den[u])
    ctx.set_pipeline(32 << 10, 4)
    try:
        b, den_b = ctx.etsi16k_features(samples, offsets, denoised=True)
    finally:
        ctx.set_pipeline(512 << 20, 1024)
    for x, y in zip(whole, b):
        assert np.array_equal(x, y)
    for x, y in zip(den, den_b):
        assert np.array_equal(x, y)


def test_features16k_full_size(ctx, oracle):
    """1 000 x 10 s at 16 kHz: row counts from the closed form, spot-check 4 utterances against the oracle."""
    samples, offsets = synth.corpus(52000, 1000, 16000, seconds=10.0, threads=8)
    got = ctx.etsi16k_features(samples, offsets)
    rows = np.array([len(g) for g in got])
    assert (rows == 988).all()        # Z null rows + (1000 - Z - 12): latency + the VAD's delay line, for any Z leading zero frames
    for u in (0, 333, 998):
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        ref, _ = oracle.etsi_features_16k(samples[lo:hi])
        assert got[u].shape == ref.shape
        assert np.array_equal(got[u][:, 14], ref[:, 14])
        err = ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1)
        assert err.max() <= OUTLIER_TOL
