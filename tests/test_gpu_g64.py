"""G64 parity: 64-channel gammatone sub-bands, IBM/IRM, masked zero-phase resynthesis.

Bars: EXACT mode is bit-identical to the oracle (int16 sub-bands, int16 resynth, IBM bits); the default
FMA-contracted mode must reach >= 60 dB SNR against the oracle's float output and +-1 LSB on int16.
"""
import numpy as np
import pytest

from conftest import snr_db
from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu

SNR_MIN_DB = 60.0


def test_gammatone_impulse_and_noise_exact(ctx, oracle):
    cf, bw, mid = oracle.channel_params()
    n = 4000
    rng = np.random.default_rng(3)
    x = np.zeros((64, n), np.float32)
    x[:, 0] = 1000.0                                   # impulse response of every channel
    x[32:] = (rng.standard_normal((32, n)) * 3000).astype(np.float32)
    ch = np.arange(64, dtype=np.int32)
    y = ctx.gammatone(x, ch, exact=True)
    yf = ctx.gammatone(x, ch, exact=False)
    for c in range(64):
        ref = oracle.gammatone(x[c], cf[c], bw[c], mid[c])
        assert np.array_equal(ref, y[c]), "channel %d not bit-exact" % c
        assert snr_db(ref, yf[c]) >= SNR_MIN_DB


def test_haircell_known_answer(ctx, oracle):
    """hairCell alone (Meddis model with the reference's double-promoted permeability term): bit-exact."""
    rng = np.random.default_rng(5)
    x = np.stack([rng.standard_normal(3000) * a for a in (0.5, 30.0, 2000.0, 60000.0)] +
                 [np.full(3000, -3.0), np.full(3000, -2.9999998), np.zeros(3000)]).astype(np.float32)
    y = ctx.haircell(x)
    for i in range(len(x)):
        assert np.array_equal(y[i], oracle.haircell(x[i])), i


def test_tables_match_oracle(ctx, oracle):
    cf, bw, mid = oracle.channel_params()
    assert np.array_equal(api.get_table("gt.cf", ctx), cf)
    assert np.array_equal(api.get_table("gt.bw", ctx), bw)
    assert np.array_equal(api.get_table("gt.mid", ctx), mid)


def test_subband_exact(ctx, oracle):
    lens = [3217, 1, 0, 6400, 777]
    xs = [synth.utterance(60 + i, 16000, n) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    out = ctx.subband(np.concatenate(xs), offsets, exact=True)
    for u, x in enumerate(xs):
        if len(x) == 0:
            continue
        got = out[64 * offsets[u]:64 * offsets[u + 1]].reshape(64, len(x))
        assert np.array_equal(got, oracle.subband(x)), "utterance %d" % u


def test_addnoise_exact(ctx, oracle):
    n = 16000
    _, clean, noise = synth.utterance(70, 16000, n, components=True)
    for db in (-5, 0, 5):
        rn, ry = oracle.addnoise(clean, noise, db)
        gn, gy = ctx.addnoise(clean, noise, [0, n], db)
        assert np.array_equal(rn, gn) and np.array_equal(ry, gy)


def test_ibm_bits_exact(ctx, oracle):
    """IBM bit-exact / IRM tight against OUR SpecInfo restatement (parity unpinned at that boundary)."""
    lens = [16000, 319, 320, 4801]
    comps = [synth.utterance(80 + i, 16000, n, components=True) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    pure = np.concatenate([c[1] for c in comps])
    noise = np.concatenate([c[2] for c in comps])
    irm, bits, fo = ctx.ibm(pure, noise, offsets)
    assert list(np.diff(fo)) == [99, 0, 1, 29]
    for u, (_, p, z) in enumerate(comps):
        if fo[u + 1] == fo[u]:
            continue
        rirm, ribm = oracle.single_ibm(oracle.subband(p), oracle.subband(z))
        gb = bits[fo[u]:fo[u + 1]]
        got = ((gb[:, None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.uint8)
        assert np.array_equal(got, ribm)
        gi = irm[fo[u]:fo[u + 1]]
        ok = np.isfinite(rirm)
        assert np.allclose(gi[ok], rirm[ok], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("mode", [api.RESYNTH_IBM, api.RESYNTH_ORI, api.RESYNTH_1DNN])
def test_resynth_modes(ctx, oracle, mode):
    """BASELINE configs 3 & 4 at small size: exact mode bit-identical, fast mode >= 60 dB / +-1 LSB."""
    lens = [8000, 4801, 320, 100, 0, 160]
    rng = np.random.default_rng(10 + mode)
    xs, masks, fo = [], [], [0]
    for i, n in enumerate(lens):
        xs.append(synth.utterance(90 + i, 16000, n))
        nf = api.resynth_num_frames(n, mode)
        assert nf == oracle.resynth_num_frames(n, mode) or n < 320
        m = rng.random((nf, 64)).astype(np.float32)
        if mode != api.RESYNTH_IBM:
            m[rng.random((nf, 64)) < 0.2] = 0.0
        masks.append(m)
        fo.append(fo[-1] + nf)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    samples = np.concatenate(xs)
    mask = np.concatenate(masks) if fo[-1] else np.zeros((0, 64), np.float32)
    exact = ctx.resynth(samples, offsets, mask, fo, mode, exact=True)
    fast = ctx.resynth(samples, offsets, mask, fo, mode, exact=False)
    fast_f = ctx.resynth(samples, offsets, mask, fo, mode, exact=False, f32=True)
    exact_f = ctx.resynth(samples, offsets, mask, fo, mode, exact=True, f32=True)
    for u, x in enumerate(xs):
        if len(x) == 0 or masks[u].shape[0] == 0:
            continue
        ref_i16, ref_f32, _ = oracle.resynth(x, masks[u], mode, want_f32=True)
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        assert np.array_equal(exact[lo:hi], ref_i16), "utterance %d exact mode" % u
        d = np.abs(fast[lo:hi].astype(np.int32) - ref_i16.astype(np.int32))
        assert d.max() <= 1
        # north_star: >= 60 dB SNR against the reference output -- float resynth_e[] against float resynth_e[]
        # (SE_FLAG_OUT_F32), so the bar is not blurred by the int16 quantisation of a quiet utterance
        if np.any(ref_f32):
            assert snr_db(ref_f32, fast_f[lo:hi]) >= SNR_MIN_DB
            assert np.array_equal(exact_f[lo:hi], ref_f32), "utterance %d exact mode, float output" % u


def test_resynth_all_ones_mask_scratch_chunking(ctx, oracle):
    """Small scratch limit forces the batch through several fwd/bwd launches; results must not change."""
    lens = [4000, 5000, 3000, 6000]
    xs = [synth.utterance(120 + i, 16000, n) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    fo = np.concatenate([[0], np.cumsum([api.resynth_num_frames(n, 0) for n in lens])]).astype(np.int64)
    mask = np.ones((int(fo[-1]), 64), np.float32)
    a = ctx.resynth(np.concatenate(xs), offsets, mask, fo, api.RESYNTH_IBM, exact=True)
    ctx.set_scratch_limit(2 << 20)        # 2 MiB = 8192 samples x 64 ch x 4 B
    b = ctx.resynth(np.concatenate(xs), offsets, mask, fo, api.RESYNTH_IBM, exact=True)
    ctx.set_scratch_limit(48 << 30)
    assert np.array_equal(a, b)
    ref, _ = oracle.resynth(xs[1], mask[fo[1]:fo[2]], 0)
    assert np.array_equal(a[offsets[1]:offsets[2]], ref)


@pytest.mark.parametrize("mode", [api.RESYNTH_IBM, api.RESYNTH_1DNN])
def test_resynth_time_chunked_default_mode(ctx, oracle, mode):
    """Default (non-EXACT) mode cuts long utterances into 1 s chunks with 2560 samples of filter run-in on
    either side (csrc/se_g64_resynth.cuh); the run-in error must stay under float rounding: +-1 LSB and
    >= 60 dB SNR against the oracle, including across chunk seams and with an impulsive input."""
    n = 16000 * 3 + 4321
    x = synth.utterance(130 + mode, 16000, n)
    x[20000] = 32767          # a click right before a seam excites the slow low-frequency channels
    x[31990:32010] = -30000
    nf = api.resynth_num_frames(n, mode)
    rng = np.random.default_rng(mode)
    m = rng.random((nf, 64)).astype(np.float32)
    m[:, :8] = 1.0            # keep the lowest (slowest-decaying) channels switched on everywhere
    got = ctx.resynth(x, [0, n], m, [0, nf], mode, exact=False)
    ref_i16, ref_f32, _ = oracle.resynth(x, m, mode, want_f32=True)
    d = np.abs(got.astype(np.int32) - ref_i16.astype(np.int32))
    assert d.max() <= 1, "max |d| = %d LSB at %d" % (d.max(), int(d.argmax()))
    assert snr_db(ref_i16, got) >= 60.0
    exact = ctx.resynth(x, [0, n], m, [0, nf], mode, exact=True)
    assert np.array_equal(exact, ref_i16)
    # SE_FLAG_LATENCY: 2 048-sample chunks (26 work items here instead of 4), same bars, float stream included
    lat = ctx.resynth(x, [0, n], m, [0, nf], mode, exact=False, latency=True)
    d = np.abs(lat.astype(np.int32) - ref_i16.astype(np.int32))
    assert d.max() <= 1, "latency chunks: max |d| = %d LSB at %d" % (d.max(), int(d.argmax()))
    lat_f = ctx.resynth(x, [0, n], m, [0, nf], mode, exact=False, f32=True, latency=True)
    assert snr_db(ref_f32, lat_f) >= 60.0
