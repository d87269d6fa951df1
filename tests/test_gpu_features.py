"""the whole ETSI feature chain behind NoiseSup (WaveProc -> CompCeps -> PostProc -> VAD),
etsi/cpp/ParmInterface.c:274-311, against the oracle (pinned bit-for-bit on the reference build by
tests/test_oracle_golden.py).

Bars (default = fast arithmetic mode): row counts and VAD flags identical; denoised int16 within 1 LSB; cepstra within
1e-4 of the frame's scale (conftest.ceps_err) except where a 1-ulp difference in the denoised stream moved one of
WaveProc's integer peak positions -- bounded by the envelope the reference shows against ITSELF between two compiler
settings (conftest.check_feature_envelope, profiles/r02_ref_vs_ref_envelope.json).  In the exact arithmetic mode every
row is bit-identical (tests/test_gpu_exact.py::test_feature_chains_every_frame).
"""
import numpy as np
import pytest

from conftest import FEAT_OUTLIER_TOL, ceps_err, check_feature_envelope
from speech_enhancement_b200 import synth

pytestmark = pytest.mark.gpu

OUTLIER_TOL = FEAT_OUTLIER_TOL


def _batch(lens, zeros):
    xs = []
    for i, (n, z) in enumerate(zip(lens, zeros)):
        x = synth.utterance(300 + i, 8000, n)
        x[:80 * z] = 0
        xs.append(x)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    return xs, offsets


def test_features_vs_oracle(ctx, oracle):
    lens = [16000, 24000, 8000, 1000, 400, 12345, 960, 0, 79, 80 * 13, 40000]
    zeros = [0, 3, 7, 0, 0, 1, 0, 0, 0, 0, 2]
    xs, offsets = _batch(lens, zeros)
    got, den = ctx.etsi_features(np.concatenate(xs) if sum(lens) else np.zeros(0, np.int16), offsets, denoised=True)
    for u, x in enumerate(xs):
        ref, ref_den = oracle.etsi_features(x)
        g = got[u]
        assert g.shape == ref.shape, "utterance %d: %s vs %s" % (u, g.shape, ref.shape)
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        nb = (hi - lo) // 80 * 80
        d = np.abs(den[lo:lo + nb].astype(np.int32) - ref_den[:nb].astype(np.int32))
        assert d.size == 0 or d.max() <= 1
        if len(ref) == 0:
            continue
        assert np.array_equal(g[:, 14], ref[:, 14]), "utterance %d: VAD flags differ" % u
        z = zeros[u]
        assert not g[:z].any()
        err = ceps_err(ref[:, :14], g[:, :14]).max(axis=1)
        assert err.max() <= OUTLIER_TOL, "utterance %d frame %d: %g" % (u, int(err.argmax()), err.max())


def test_features_within_the_reference_own_envelope(ctx, oracle):
    """40 x 5 s (19 520 rows): shares of frames beyond 1e-4 / 1e-3 of the frame scale and the largest error stay inside
    what two builds of the reference show against each other."""
    xs = [synth.utterance(5000 + i, 8000, 8000 * 5) for i in range(40)]
    offsets = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
    got = ctx.etsi_features(np.concatenate(xs), offsets)
    errs = []
    for u, x in enumerate(xs):
        ref, _ = oracle.etsi_features(x)
        assert got[u].shape == ref.shape and np.array_equal(got[u][:, 14], ref[:, 14])
        errs.append(ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1))
    n_tail, n_flip = check_feature_envelope(np.concatenate(errs))
    print("8 kHz feature chain, fast mode: %d frames above 1e-4, %d above 1e-3 of %d" % (n_tail, n_flip, sum(map(len, errs))))


def test_features_device_memory_and_shards(ctx, oracle):
    """Any split of the batch gives identical rows (no cross-utterance state)."""
    lens = [8000, 5000, 12000, 3000]
    xs, offsets = _batch(lens, [0, 1, 0, 0])
    whole = ctx.etsi_features(np.concatenate(xs), offsets)
    for u, x in enumerate(xs):
        alone = ctx.etsi_features(x, [0, len(x)])[0]
        assert np.array_equal(alone, whole[u])


def test_wave_proc_known_answer(ctx, oracle):
    """DoWaveProc alone on identical input frames is bit-exact (integer peak picking, exact float products)."""
    rng = np.random.default_rng(0)
    frames = []
    for t in range(600):
        kind = t % 5
        n = np.arange(200)
        if kind == 0:
            f = rng.standard_normal(200) * rng.choice([0.1, 1, 30, 3000, 12000])
        elif kind == 1:
            f = 3000 * np.sin(2 * np.pi * n / rng.integers(30, 90) + rng.random()) * (1 + 0.5 * np.sin(2 * np.pi * n / 57))
        elif kind == 2:
            f = np.zeros(200)
            f[rng.integers(0, 200, rng.integers(1, 8))] = rng.integers(-32768, 32767)
        elif kind == 3:
            f = rng.integers(-32768, 32768, 200)
        else:
            f = np.round(rng.standard_normal(200) * 3)
        frames.append(f.astype(np.float32))
    frames = np.stack(frames)
    got = ctx.wave_proc(frames)
    for k in range(len(frames)):
        e = float(np.sum(frames[k].astype(np.float64) ** 2))
        if abs(e - 100.0) < 1e-3:
            continue                         # summation-order tie at the low-energy threshold
        assert np.array_equal(got[k], oracle.wave_proc(frames[k])), "frame %d" % k


def test_comp_ceps_known_answer(ctx, oracle):
    """DoCompCeps alone (se_comp_ceps_batch) on identical input frames: rel 1e-4 on the frame's cepstral scale."""
    rng = np.random.default_rng(2)
    n = np.arange(201)
    frames = np.stack([(rng.standard_normal(201) * a).astype(np.float32) for a in (1.0, 300.0, 9000.0)] +
                      [(4000 * np.sin(2 * np.pi * n * f / 8000.0)).astype(np.float32) for f in (100.0, 1234.5, 3900.0)] +
                      [np.zeros(201, np.float32), synth.utterance(3, 8000, 201).astype(np.float32)] * 20)
    got = ctx.comp_ceps(frames)
    for k in range(len(frames)):
        ref = oracle.comp_ceps(frames[k])
        assert ceps_err(ref, got[k]).max() <= 1e-4, k
    assert np.abs(got[6, :12]).max() < 1e-4 and got[6, 12] == -230.0 and got[6, 13] == -50.0   # all floors on a zero frame


def test_features_pipelined_chunks(ctx):
    """SE_MEM_HOST calls are cut into H2D/kernel/D2H chunks on several streams; tiny chunks must not change the rows."""
    samples, offsets = synth.corpus(900, 60, 8000, min_s=0.3, max_s=1.2)
    a, den_a = ctx.etsi_features(samples, offsets, denoised=True)
    ctx.set_pipeline(32 << 10, 4)
    try:
        b, den_b = ctx.etsi_features(samples, offsets, denoised=True)
    finally:
        ctx.set_pipeline(512 << 20, 1024)
    assert np.array_equal(den_a, den_b)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_flush_rows(ctx, oracle):
    """FlushAdvProcess / DoVADFlush (ParmInterface.c:347-355, VAD.c:342-431) behind se_ctx_set_feature_flush: six more
    rows per utterance -- the frames still in the VAD's delay line with the flush's own trigger logic, or, for
    utterances too short to have started the VAD, six copies of the stale FeatureBuffer, like the reference's driver
    loop.  Row counts and flags identical, coefficients of the flushed rows = rows already checked (delay line)."""
    lens = [16000, 24000, 8000, 80 * 9, 80 * 3, 0, 79, 80 * 16, 80 * 17, 12345]
    zeros = [0, 3, 7, 0, 0, 0, 0, 0, 2, 1]
    xs, offsets = _batch(lens, zeros)
    samples = np.concatenate(xs)
    plain = ctx.etsi_features(samples, offsets)
    ctx.set_feature_flush(True)
    try:
        got = ctx.etsi_features(samples, offsets)
        got16 = ctx.etsi16k_features(synth.utterance(31, 16000, 20000), [0, 20000])[0]
    finally:
        ctx.set_feature_flush(False)
    for u, x in enumerate(xs):
        ref = oracle.etsi_features_flush(x)
        g = got[u]
        assert g.shape == ref.shape == (len(plain[u]) + 6, 15), (u, g.shape, ref.shape)
        assert np.array_equal(g[:-6], plain[u])                       # the flush only appends
        assert np.array_equal(g[:, 14], ref[:, 14]), "utterance %d: VAD flags differ" % u
        err = ceps_err(ref[-6:, :14], g[-6:, :14]).max(axis=1)
        assert err.max() <= OUTLIER_TOL, (u, err)          # delay-line rows: the errors of the frames they are
    ref16 = oracle.etsi_features_flush(synth.utterance(31, 16000, 20000), True)
    assert got16.shape == ref16.shape and np.array_equal(got16[:, 14], ref16[:, 14])
    assert ceps_err(ref16[-6:, :14], got16[-6:, :14]).max() <= OUTLIER_TOL
    assert np.array_equal(ctx.etsi_features(samples, offsets)[0], plain[0])          # and it is off again
