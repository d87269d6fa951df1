"""CUDA path vs the committed golden vectors, which were produced by the reference's OWN sources
(tests/golden/make_golden.py) — the GPU box has no /root/reference, so these fixtures are how the
kernels are checked against the real reference there."""
import os

import numpy as np
import pytest

from conftest import ceps_err, snr_db
from speech_enhancement_b200 import api

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name))


@pytest.mark.parametrize("prefix,xkey", [("", "x"), ("z_", "xz")])
def test_p8_against_reference_vectors(ctx, prefix, xkey):
    g = load("p8_etsi.npz")
    x = g[xkey]
    calls = len(x) // 80
    r = ctx.etsi_denoise(x, [0, len(x)], ceps=True, den_f32=True, trace=True)
    ref_i16 = g[prefix + "den_i16"][:calls * 80].astype(np.int32)
    assert np.abs(r["den_i16"][:calls * 80].astype(np.int32) - ref_i16).max() <= 1
    ret = g[prefix + "ret"]
    produced = ret > 0
    got_f = r["den_f32"][:calls * 80].reshape(calls, 80)
    assert snr_db(g[prefix + "out_f32"][produced], got_f[produced]) >= 60.0
    assert int(r["n_ceps"][0]) == len(g[prefix + "ceps"])                      # frame count: exact
    assert ceps_err(g[prefix + "ceps"], r["ceps"][:len(g[prefix + "ceps"])]).max() <= 1e-4
    tr = r["trace"][:calls]
    assert np.array_equal((tr & 64) != 0, produced)
    ran = (tr & 32) != 0
    mism = sum(int(np.sum(((tr[ran] & bit) != 0) != (g[prefix + k][ran] != 0)))
               for bit, k in ((1, "flag_vad"), (2, "sp_var"), (4, "sp_spec"), (8, "sp_mel"), (16, "sp_vadns")))
    assert mism == 0                                                           # no tie on these fixtures


SPEC_FLOOR = float(np.exp(-10.0))          # floor e^-10 for PSD-domain quantities


def spec_rel(ref, got):
    ref = np.asarray(ref, np.float64)
    got = np.asarray(got, np.float64)
    return np.abs(ref - got) / np.maximum(np.maximum(np.abs(ref), np.abs(got)), SPEC_FLOOR)


def test_spectra_against_reference_vectors(ctx):
    """a6/a7/a9 at the spectrum level: sqrt(PSD) (FFTtoPSD), noise estimate and denoised spectrum (PSDMean -> FilterCalc)
    of BOTH stages after every kept call, against the reference's own per-call state (default fast mode; bars as
    tests/test_gpu_spectra.py; the exact mode is bit-identical on the same vectors, tests/test_gpu_exact.py)."""
    g = load("spectra.npz")
    for prof, xk, ck, sk in ((8, "x8", "calls8", "sp8"), (16, "a16", "calls16", "sp16")):
        x = g[xk]
        sp, produced, _ = ctx.ns_spectra(prof, x, [0, len(x)])
        calls = g[ck]
        for st, bit in ((0, 2), (1, 1)):
            m = (produced[calls] & bit) != 0
            assert m.sum() >= len(calls) - 6
            ref, got = g[sk][m][:, 3 * st:3 * st + 3], sp[calls][m][:, 3 * st:3 * st + 3]
            err = spec_rel(ref, got)
            fr = np.abs(ref.astype(np.float64) - got) / np.maximum(np.abs(ref).max(axis=2, keepdims=True), SPEC_FLOOR)
            assert fr.max() <= 1e-4, (prof, st, fr.max())            # on the frame's scale: every bin
            if st == 0:
                assert err.max() <= 1e-4, (prof, st, err.max())      # element-wise: every stage-1 bin
            else:                                                    # stage 2: deep notches carry FFT rounding noise
                assert (err > 1e-4).sum() <= 1e-3 * err.size + 2, (prof, st, int((err > 1e-4).sum()))
    gp = load("p8_etsi.npz")                 # the fixture's last-call noise / denoised spectra
    sp, produced, _ = ctx.ns_spectra(8, gp["x"], [0, len(gp["x"])])
    assert spec_rel(gp["noise1_last"], sp[-1, 1]).max() <= 1e-4
    assert (np.abs(gp["den2_last"].astype(np.float64) - sp[-1, 5]) / np.abs(gp["den2_last"]).max()).max() <= 1e-4
    assert np.array_equal((produced & 1) != 0, gp["ret"] > 0)


def test_rfft_against_reference_vectors(ctx):
    g = load("rfft.npz")
    y = ctx.rfft(g["x256"], 8)
    assert np.max(np.abs(y - g["y256"])) <= 1e-4 * np.max(np.abs(g["y256"]))
    g = load("p16_wiener.npz")
    y = ctx.rfft(g["x512"], 8)
    assert np.max(np.abs(y - g["y512"])) <= 1e-4 * np.max(np.abs(g["y512"]))


def test_p16_against_reference_vectors(ctx):
    g = load("p16_wiener.npz")
    s = ctx.ns_stream(16)
    for tag in ("a", "b"):
        r = s.push(g[tag])
        assert snr_db(g[tag + "_out"], r["out"]) >= 60.0
        assert np.array_equal(r["flags"][4], g[tag + "_frame_counter"])
        for i, k in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")):
            assert np.array_equal(r["flags"][i], g[tag + "_" + k]), k
        assert r["gains"].shape == g[tag + "_gains"].shape
        assert np.max(np.abs(r["gains"] - g[tag + "_gains"])) <= 1e-6 + 1e-4 * np.abs(g[tag + "_gains"]).max()
    s.close()


def test_g64_against_reference_vectors(ctx):
    g = load("g64.npz")
    assert np.array_equal(api.get_table("gt.cf", ctx), g["cf"])
    x = np.tile(g["imp"], (len(g["chans"]), 1))
    assert np.array_equal(ctx.gammatone(x, g["chans"], exact=True), g["gt"])
    w = g["w"]
    assert np.array_equal(ctx.subband(w, [0, len(w)], exact=True).reshape(64, -1), g["sub"])
    an, ay = ctx.addnoise(g["clean"], g["noise"], [0, len(w)], 5)
    assert np.array_equal(an, g["an"]) and np.array_equal(ay, g["ay"])
    for mode in (0, 1, 2):
        m = g["mask%d" % mode]
        out = ctx.resynth(w, [0, len(w)], m, [0, len(m)], mode, exact=True)
        assert np.array_equal(out, g["out%d" % mode]), "mode %d" % mode
        fast = ctx.resynth(w, [0, len(w)], m, [0, len(m)], mode, exact=False)
        assert np.abs(fast.astype(np.int32) - g["out%d" % mode]).max() <= 1
