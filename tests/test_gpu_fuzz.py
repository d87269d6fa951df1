"""Randomised edge-case sweep: signals the synthetic speech corpus never produces (silence, DC, full-scale square
waves, clicks, zero runs in odd places, tiny amplitudes, clipping) through P8, P16 and the feature chain.

Waveforms, counts and integer bookkeeping are held to the main parity bars on every signal.  Cepstra, gains and
the frame-dropping flags are ill-conditioned on some of these signals (log of band energies at the floor, DCT sums
that cancel to ~0, exactly periodic inputs with many equal Teager peaks, all-zero frames sitting exactly on a VAD
threshold): the reference disagrees WITH ITSELF there when it is merely compiled differently.  So those quantities
are self-calibrated per signal kind: the GPU's deviation from the reference (-O2, no FMA contraction) may not
exceed max(main bar, SLACK x the deviation of the reference's -O3/FMA build from that same -O2 build)."""
import numpy as np
import pytest

from conftest import ceps_err, rel_err, snr_db
from oracle import pyoracle
from speech_enhancement_b200 import synth

pytestmark = pytest.mark.gpu

SLACK = 8.0
KINDS = 10


@pytest.fixture(scope="module")
def oracle_fma():
    return pyoracle.Oracle(fast=True)


def _signals(rng, fs, count, max_s):
    out = []
    for i in range(count):
        n = int(rng.integers(0, int(max_s * fs)))
        t = np.arange(n)
        kind = i % 10
        if kind == 0:
            x = np.zeros(n)
        elif kind == 1:
            x = np.full(n, float(rng.integers(-20000, 20000)))                      # pure DC
        elif kind == 2:
            x = 32767.0 * np.sign(np.sin(2 * np.pi * t * rng.uniform(50, 3000) / fs))  # full-scale square wave
        elif kind == 3:
            x = rng.standard_normal(n) * rng.choice([0.6, 3.0, 40.0])               # near-silent noise
        elif kind == 4:
            x = rng.standard_normal(n) * 30000.0                                     # heavy clipping
        elif kind == 5:
            x = np.zeros(n)
            if n:
                x[rng.integers(0, n, max(1, n // 500))] = rng.choice([-32768.0, 32767.0, 1.0])  # clicks
        elif kind == 6:
            x = synth.utterance(7000 + i, fs, n).astype(np.float64)
            for _ in range(3):                                                       # zero runs anywhere
                if n > 400:
                    a = int(rng.integers(0, n - 400))
                    x[a:a + int(rng.integers(1, 400))] = 0
        elif kind == 7:
            x = 8000.0 * np.sin(2 * np.pi * t * rng.uniform(20, fs / 2 - 50) / fs) + 500.0   # tone + DC
        elif kind == 8:
            x = synth.utterance(7000 + i, fs, n).astype(np.float64) * rng.choice([0.001, 0.02, 4.0])
        else:
            x = np.where(t % 2 == 0, 32767.0, -32768.0)                              # Nyquist, full scale
        out.append(np.clip(np.round(x), -32768, 32767).astype(np.int16))
    return out


def _batch(xs):
    lens = [len(x) for x in xs]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    return (np.concatenate(xs) if offsets[-1] else np.zeros(0, np.int16)), offsets


def _flag_bits(tr, info):
    on = info["stage0_ran"] > 0
    return sum(int(np.sum(((tr[on] & bit) != 0) != (info[key][on] != 0)))
               for bit, key in ((1, "flag_vad"), (2, "sp_var"), (4, "sp_spec"), (8, "sp_mel"), (16, "sp_vadns"))), int(on.sum())


def test_fuzz_p8(ctx, oracle, oracle_fma):
    xs = _signals(np.random.default_rng(8), 8000, 120, 1.5)
    samples, offsets = _batch(xs)
    r = ctx.etsi_denoise(samples, offsets, ceps=True, den_f32=True, trace=True)
    g_ceps, s_ceps = np.zeros(KINDS), np.zeros(KINDS)
    g_flag, s_flag, frames = np.zeros(KINDS, int), np.zeros(KINDS, int), np.zeros(KINDS, int)
    for u, x in enumerate(xs):
        k = u % KINDS
        ref = oracle.etsi_denoise(x, with_ceps=True, trace=True)
        alt = oracle_fma.etsi_denoise(x, with_ceps=True, trace=True)
        lo = int(offsets[u])
        nb = len(x) // 80 * 80
        nc = len(ref["ceps"])
        assert int(r["n_ceps"][u]) == nc, u
        d = np.abs(r["den_i16"][lo:lo + nb].astype(np.int32) - ref["den_i16"][:nb].astype(np.int32))
        assert d.max(initial=0) <= 1, "utt %d kind %d: %d LSB" % (u, k, d.max())
        if np.any(ref["den_f32"][:nb]):
            assert snr_db(ref["den_f32"][:nb], r["den_f32"][lo:lo + nb]) >= 60.0, (u, k)
        co = int(r["ceps_offsets"][u])
        if nc:
            g_ceps[k] = max(g_ceps[k], ceps_err(ref["ceps"], r["ceps"][co:co + nc]).max())
            s_ceps[k] = max(s_ceps[k], ceps_err(ref["ceps"], alt["ceps"]).max())
        tr = r["trace"][int(r["trace_offsets"][u]):int(r["trace_offsets"][u]) + ref["calls"]]
        info = ref["info"]
        assert np.array_equal((tr & 128) != 0, info["produced"] < 0), u
        assert np.array_equal((tr & 64) != 0, info["produced"] > 0), u
        m, n = _flag_bits(tr, info)
        g_flag[k] += m
        frames[k] += n
        on = info["stage0_ran"] > 0
        s_flag[k] += sum(int(np.sum(info[key][on] != alt["info"][key][on]))
                         for key in ("flag_vad", "sp_var", "sp_spec", "sp_mel", "sp_vadns"))
    for k in range(KINDS):
        assert g_ceps[k] <= max(1e-4, SLACK * s_ceps[k]), "kind %d: cepstra %g (reference vs itself %g)" % (k, g_ceps[k], s_ceps[k])
        assert g_flag[k] <= max(2, 1e-3 * 5 * frames[k], SLACK * s_flag[k]), \
            "kind %d: %d flag mismatches (reference vs itself %d) in %d frames" % (k, g_flag[k], s_flag[k], frames[k])


def test_fuzz_p16(ctx, oracle, oracle_fma):
    xs = _signals(np.random.default_rng(16), 16000, 100, 1.2)
    samples, offsets = _batch(xs)
    r = ctx.p16_wiener(samples, offsets)
    bo = r["block_offsets"]
    g_gain, s_gain = np.zeros(KINDS), np.zeros(KINDS)
    g_flag, s_flag, blocks = np.zeros(KINDS, int), np.zeros(KINDS, int), np.zeros(KINDS, int)
    for u, x in enumerate(xs):
        k = u % KINDS
        s = oracle.p16_stream()
        ref = s.func_wiener(x.astype(np.float32))
        s.close()
        s = oracle_fma.p16_stream()
        alt = s.func_wiener(x.astype(np.float32))
        s.close()
        lo = int(offsets[u])
        nb = len(x) // 160
        nr = int(r["n_rows"][u])
        assert nr == len(ref["gains"]), (u, k)
        if np.any(ref["out"][:nb * 160]):
            assert snr_db(ref["out"][:nb * 160], r["out"][lo:lo + nb * 160]) >= 60.0, (u, k)
        else:
            assert not r["out"][lo:lo + nb * 160].any()
        if nr:
            g_gain[k] = max(g_gain[k], rel_err(ref["gains"], r["gains"][bo[u]:bo[u] + nr], 1e-3).max())
            if len(alt["gains"]) == nr:
                s_gain[k] = max(s_gain[k], rel_err(ref["gains"], alt["gains"], 1e-3).max())
        fl = r["flags"][:, bo[u]:bo[u] + nb]
        assert np.array_equal(fl[4], ref["frame_counter"]), u
        for i, key in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")):
            g_flag[k] += int(np.sum(fl[i] != ref[key]))
            s_flag[k] += int(np.sum(alt[key] != ref[key]))
        blocks[k] += nb
    for k in range(KINDS):
        assert g_gain[k] <= max(1e-4, SLACK * s_gain[k]), "kind %d: gains %g (reference vs itself %g)" % (k, g_gain[k], s_gain[k])
        assert g_flag[k] <= max(2, 1e-3 * 4 * blocks[k], SLACK * s_flag[k]), (k, g_flag[k], s_flag[k])


def test_fuzz_features(ctx, oracle, oracle_fma):
    xs = _signals(np.random.default_rng(15), 8000, 100, 1.5)
    samples, offsets = _batch(xs)
    got = ctx.etsi_features(samples, offsets)
    g_bad, s_bad, rows = np.zeros(KINDS, int), np.zeros(KINDS, int), np.zeros(KINDS, int)
    g_vad, s_vad, s_under = np.zeros(KINDS, int), np.zeros(KINDS, int), np.zeros(KINDS, int)
    for u, x in enumerate(xs):
        k = u % KINDS
        ref, _ = oracle.etsi_features(x)
        alt, _ = oracle_fma.etsi_features(x)
        # the VAD flag is a hangover-smoothed function of the NoiseSup speech flags: their own sensitivity counts
        ia = oracle.etsi_denoise(x, trace=True)["info"]
        ib = oracle_fma.etsi_denoise(x, trace=True)["info"]
        s_under[k] += sum(int(np.sum(ia[key] != ib[key])) for key in ("sp_var", "sp_spec", "sp_mel", "sp_vadns"))
        assert got[u].shape == ref.shape, (u, k)
        if not len(ref):
            continue
        g_vad[k] += int(np.sum(got[u][:, 14] != ref[:, 14]))
        g_bad[k] += int((ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1) > 1e-4).sum())
        if alt.shape == ref.shape:
            s_vad[k] += int(np.sum(alt[:, 14] != ref[:, 14]))
            s_bad[k] += int((ceps_err(ref[:, :14], alt[:, :14]).max(axis=1) > 1e-4).sum())
        rows[k] += len(ref)
    for k in range(KINDS):
        # rows above 1e-4 are WaveProc peak-position flips (tests/test_gpu_features.py): bounded by the main bar's
        # share on ordinary signals, by the reference's own flip rate on periodic / impulsive ones
        assert g_bad[k] <= max(0.03 * rows[k], 4 * s_bad[k] + 5), "kind %d: %d rows off (reference vs itself %d) of %d" % (k, g_bad[k], s_bad[k], rows[k])
        # kind 5 (isolated clicks in digital silence): almost every frame is all-zero input, where SpeechQVar's
        # variance E[W^2]-E[W]^2 (NoiseSup.c:813) is pure summation-order rounding noise compared against a running mean
        # of the same noise -- a tie by construction (the reference's two builds differ there too, test_fuzz_p8)
        tie_share = 2e-2 if k == 5 else 2e-3
        assert g_vad[k] <= max(2, tie_share * rows[k], SLACK * s_vad[k], s_under[k]), (k, g_vad[k], s_vad[k], s_under[k])


def test_fuzz_g64_exact_modes_stay_bit_exact(ctx, oracle):
    """EXACT-mode G64 kernels reproduce the reference's float arithmetic operation by operation, so even the
    degenerate signals must come out bit-identical (sub-band int16 rows, IBM bits, IBM-masked resynthesis);
    the default FMA mode must still reach +-1 LSB on them."""
    from speech_enhancement_b200 import api
    xs = [x for x in _signals(np.random.default_rng(64), 16000, 30, 0.25) if len(x) >= 400]
    samples, offsets = _batch(xs)
    sub = ctx.subband(samples, offsets, exact=True)
    rng = np.random.default_rng(3)
    for u, x in enumerate(xs):
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        ref_sub = oracle.subband(x)
        assert np.array_equal(sub[64 * lo:64 * hi].reshape(64, len(x)), ref_sub), (u, u % 10)
        nf = api.resynth_num_frames(len(x), api.RESYNTH_IBM)
        m = rng.random((nf, 64)).astype(np.float32)
        ref_i16, _ = oracle.resynth(x, m.copy(), api.RESYNTH_IBM)
        assert np.array_equal(ctx.resynth(x, [0, len(x)], m, [0, nf], api.RESYNTH_IBM, exact=True), ref_i16), u
        fast = ctx.resynth(x, [0, len(x)], m, [0, nf], api.RESYNTH_IBM, exact=False)
        # full-scale inputs wrap the int16 cast (Q4): compare modulo 2^16
        d = (fast.astype(np.int32) - ref_i16.astype(np.int32) + 32768) % 65536 - 32768
        assert np.abs(d).max(initial=0) <= 1, (u, u % 10, int(np.abs(d).max()))
    # IBM bits: clean = signal u, noise = signal u+1 (cut to the same length)
    for u in range(0, len(xs) - 1, 2):
        n = min(len(xs[u]), len(xs[u + 1]))
        p, z = xs[u][:n], xs[u + 1][:n]
        irm, bits, fo = ctx.ibm(p, z, [0, n])
        rirm, ribm = oracle.single_ibm(oracle.subband(p), oracle.subband(z))
        got = ((bits[:, None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.uint8)
        assert np.array_equal(got, ribm), u
