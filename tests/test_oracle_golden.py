"""Pins the plain-C oracle (oracle/se_oracle.c) bit-for-bit against golden vectors generated from the
reference's OWN sources (tests/golden/make_golden.py) and, where oracle/_ref is present, against the
reference build live on fresh inputs.  CPU only."""
import os

import numpy as np
import pytest

from oracle import pyoracle as po
from speech_enhancement_b200 import synth

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name))


def test_synth_generator_is_deterministic():
    """The fixtures embed their inputs; the generator must still produce the same samples."""
    g = load("p8_etsi.npz")
    assert np.array_equal(synth.utterance(3, 8000, 8033), g["x"])
    assert np.array_equal(synth.utterance(7, 8000, 4000), g["xz"])
    assert np.all(g["xz"][:400] == 0) and np.any(g["xz"][400:480] != 0)


@pytest.mark.parametrize("prefix,xkey", [("", "x"), ("z_", "xz")])
def test_p8_etsi_golden(oracle, prefix, xkey):
    g = load("p8_etsi.npz")
    x = g[xkey]
    r = oracle.etsi_denoise(x, with_ceps=True, trace=True)
    calls = len(x) // 80
    assert r["calls"] == calls
    assert np.array_equal(r["den_i16"][:calls * 80], g[prefix + "den_i16"][:calls * 80])
    assert np.array_equal(r["ceps"], g[prefix + "ceps"])             # bit-exact floats
    info = r["info"]
    ran = g[prefix + "ret"] >= 0
    assert np.array_equal(info["produced"] > 0, g[prefix + "ret"] > 0)
    for k in ("flag_vad", "hang_over", "nb_speech", "nb_frame0", "nb_frame1", "sp_var", "sp_spec", "sp_mel", "sp_vadns"):
        assert np.array_equal(info[k][ran], g[prefix + k][ran]), k
    for k in ("mean_en", "alfa_gf"):
        assert np.array_equal(info[k][ran], g[prefix + k][ran]), k
    n_ceps = calls - 6 - (5 if prefix else 0)                        # 5 leading all-zero frames are skipped
    assert len(r["ceps"]) == n_ceps


def test_spectra_golden(oracle):
    """Per-call SPECTRUM state (nSigSE / noiseSE / denSigSE of both Wiener stages, NoiseSup.c:102-115 as FFTtoPSD ->
    PSDMean -> FilterCalc leave them) bit-for-bit against the reference build, both profiles."""
    g = load("spectra.npz")
    for prof, xk, ck, sk in ((8, "x8", "calls8", "sp8"), (16, "a16", "calls16", "sp16")):
        sp, ran = oracle.ns_spectra(prof, g[xk])
        calls = g[ck]
        for st, bit in ((0, 2), (1, 1)):
            m = (ran[calls] & bit) != 0                       # the planes of a stage move only when it runs
            assert m.sum() >= len(calls) - 6
            assert np.array_equal(sp[calls][m][:, 3 * st:3 * st + 3], g[sk][m][:, 3 * st:3 * st + 3]), (prof, st)
    assert np.array_equal((oracle.ns_spectra(8, g["x8"])[1] & 1) != 0, g["ret8"] > 0)


def test_p8_features_golden(oracle):
    """WaveProc -> CompCeps -> PostProc -> VAD behind NoiseSup: bit-exact rows against the reference build."""
    g = load("p8_features.npz")
    for xk, fk, dk in (("x", "feat", "den"), ("xz", "z_feat", "z_den")):
        f, d = oracle.etsi_features(g[xk])
        assert f.shape == g[fk].shape and np.array_equal(f, g[fk])
        assert np.array_equal(d, g[dk])
    assert len(g["feat"]) == 100 - 12 and len(g["z_feat"]) == 5 + (50 - 5 - 12)
    assert not g["z_feat"][:5].any() and g["feat"][0, :14].all()        # null vectors, then real frames
    # FlushAdvProcess (DoVADFlush): 6 more rows; on an utterance too short to start the VAD, six stale FeatureBuffers
    assert np.array_equal(oracle.etsi_features_flush(g["x"])[-6:], g["flush"])
    assert np.array_equal(oracle.etsi_features_flush(g["x"])[:-6], g["feat"])
    assert np.array_equal(oracle.etsi_features_flush(g["xz"])[-6:], g["z_flush"])
    assert np.array_equal(oracle.etsi_features_flush(g["xs"]), g["s_flush"]) and g["s_flush"].shape == (6, 15)
    for v, w in zip(g["frames"], g["wp"]):
        assert np.array_equal(oracle.wave_proc(v), w)
    assert np.array_equal(g["wp"][2], g["frames"][2])                    # low-energy frames pass untouched


def test_etsi16k_golden(oracle):
    """ETSI 16 kHz mode: QMF split, HP-band coding/decoding, 26-band CompCeps -- bit-exact against the reference build."""
    g = load("etsi16k_features.npz")
    for xk, fk, dk in (("x", "feat", "den"), ("xz", "z_feat", "z_den")):
        f, d = oracle.etsi_features_16k(g[xk])
        assert f.shape == g[fk].shape and np.array_equal(f, g[fk])
        assert np.array_equal(d, g[dk])
    assert len(g["feat"]) == 100 - 12 and len(g["z_feat"]) == 5 + (50 - 5 - 12)
    assert not g["z_feat"][:5].any() and g["feat"][0, :14].all()
    assert np.array_equal(oracle.etsi_features_flush(g["x"], True)[-6:], g["flush"])
    lp, hp = oracle.qmf_split(g["q"])
    assert np.array_equal(lp, g["lp"]) and np.array_equal(hp, g["hp"])
    # the 16 kHz rows are not the 8 kHz rows of the decimated signal: the HP bands enter c0 / logE / the DCT
    f8, _ = oracle.etsi_features(np.ascontiguousarray(g["x"][::2]))
    assert f8.shape == g["feat"].shape and np.abs(f8[:, 12] - g["feat"][:, 12]).min() > 1.0


def test_rfft_golden(oracle):
    g = load("rfft.npz")
    for x, y in zip(g["x256"], g["y256"]):
        assert np.array_equal(oracle.rfft(x, 8), y)
    g = load("p16_wiener.npz")
    for x, y in zip(g["x512"], g["y512"]):
        assert np.array_equal(oracle.rfft(x, 8, cxx=True), y)        # truncated schedule, SURVEY Q1
        z = np.fft.rfft(x.astype(np.float64))
        assert np.max(np.abs(y[:200] - z.real[:200])) > 1.0           # ... which is NOT the true 512-pt FFT


def test_p8_tables_golden(oracle):
    g = load("p8_tables.npz")
    t = oracle.ns_tables(8)
    for k in ("sig_window", "ir_window", "bank_start", "bank_len", "idct"):
        assert np.array_equal(t[k], g[k]), k
    for j in range(25):
        n = g["bank_len"][j]
        assert np.array_equal(t["bank"][j, :n], g["bank"][j, :n])
    c = oracle.cc_tables()
    assert np.array_equal(c["hamming"], g["hamming"]) and np.array_equal(c["dct"], g["dct"])
    assert np.array_equal(c["start"], g["cc_start"]) and np.array_equal(c["len"], g["cc_len"])


def test_p16_wiener_golden(oracle):
    g = load("p16_wiener.npz")
    s = oracle.p16_stream()
    for tag in ("a", "b"):                                           # second call continues the first's state
        r = s.func_wiener(g[tag])
        for k in ("out", "sp_var", "sp_spec", "sp_mel", "sp_vadns", "frame_counter"):
            assert np.array_equal(r[k], g[tag + "_" + k]), (tag, k)
        # the reference only exposes the gains as "%f " text: equal to 6 decimals
        assert r["gains"].shape == g[tag + "_gains"].shape
        assert np.max(np.abs(r["gains"] - g[tag + "_gains"])) <= 5.1e-7
    s.close()
    rows = str(g["text_a"]).strip().split("\n")
    assert len(rows) == len(g["a_gains"]) == 100 - 2 - 4            # two zero blocks skipped, 4 blocks latency
    t = oracle.ns_tables(16)
    gt = load("p16_tables.npz")
    for k in ("sig_window", "ir_window", "bank", "idct"):
        assert np.array_equal(t[k], gt[k]), k


def test_g64_golden(oracle):
    g = load("g64.npz")
    cf, bw, mid = oracle.channel_params()
    assert np.array_equal(cf, g["cf"]) and np.array_equal(bw, g["bw"]) and np.array_equal(mid, g["mid"])
    for i, c in enumerate(g["chans"]):
        assert np.array_equal(oracle.gammatone(g["imp"], cf[c], bw[c], mid[c]), g["gt"][i])
    assert np.array_equal(oracle.haircell(g["gt"][1] * 50), g["hc"])
    assert np.array_equal(oracle.subband(g["w"]), g["sub"])
    an, ay = oracle.addnoise(g["clean"], g["noise"], 5)
    assert np.array_equal(an, g["an"]) and np.array_equal(ay, g["ay"])
    for mode in (0, 1, 2):
        out, m_after = oracle.resynth(g["w"], g["mask%d" % mode], mode)
        assert np.array_equal(out, g["out%d" % mode]), mode
        assert np.array_equal(m_after, g["mask_after%d" % mode])


def test_ibm_restatement_properties(oracle):
    """make_single_IBM sits on the un-vendored asdk::SpecInfo: PARITY UNPINNED.  What is
    pinned by the reference source: frame count (show_IBM.cpp:110), IRM = Sp/(Sp+Sn), IBM = Sn < Sp."""
    _, clean, noise = synth.utterance(12, 16000, 3200, components=True)
    p64, n64 = oracle.subband(clean), oracle.subband(noise)
    irm, ibm = oracle.single_ibm(p64, n64)
    assert irm.shape == ((3200 - 320) // 160 + 1, 64)
    ok = np.isfinite(irm)
    assert np.all((irm[ok] >= 0) & (irm[ok] <= 1))
    assert np.array_equal(ibm[ok] == 1, irm[ok] > 0.5) or np.mean((ibm[ok] == 1) != (irm[ok] > 0.5)) < 1e-3
    irm2, ibm2 = oracle.single_ibm(n64, p64)                         # swapping the roles flips the mask
    both = ok & np.isfinite(irm2) & (irm != 0.5)
    assert np.allclose(irm[both] + irm2[both], 1.0, atol=1e-5)


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_vs_reference_build_live():
    """Fresh inputs, bit-exact against the reference's own code (all paths)."""
    o = po.Oracle()
    x = synth.utterance(21, 8000, synth.ragged_length(21, 8000, 1.0, 2.0))
    a, b = o.etsi_denoise(x, with_ceps=True), po.RefEtsi().trace(x)
    n = len(x) // 80 * 80
    assert np.array_equal(a["den_i16"][:n], b["den_i16"][:n]) and np.array_equal(a["ceps"], b["ceps"])
    fa, fb = o.etsi_features(x), po.RefEtsi().etsi_features(x)
    assert np.array_equal(fa[0], fb[0]) and np.array_equal(fa[1], fb[1]) and fa[0][:, 14].any()
    for xx, m16 in ((x, False), (x[:80 * 8], False), (x[:0], False), (synth.utterance(24, 16000, 20000), True)):
        assert np.array_equal(o.etsi_features_flush(xx, m16), po.RefEtsi().etsi_features_flush(xx, m16))
    xw = synth.utterance(24, 16000, synth.ragged_length(24, 16000, 1.0, 2.0))
    ga, gb = o.etsi_features_16k(xw), po.RefEtsi().etsi_features_16k(xw)
    assert np.array_equal(ga[0], gb[0]) and np.array_equal(ga[1], gb[1]) and ga[0][:, 14].any()
    x16 = synth.utterance(22, 16000, 16000 + 77).astype(np.float32)
    s, R = o.p16_stream(), po.RefP16()
    A, B = s.func_wiener(x16), R.func_wiener(x16)
    for k in ("out", "sp_var", "sp_spec", "sp_mel", "sp_vadns", "frame_counter"):
        assert np.array_equal(A[k], B[k]), k
    s.close()
    R.close()
    w = synth.utterance(23, 16000, 4000)
    assert np.array_equal(o.subband(w), po.RefG64("subband").subband(w))
    rng = np.random.default_rng(2)
    for v, mode in (("resynth_ibm", 0), ("resynth_ori", 1), ("resynth_1dnn", 2)):
        m = rng.random((o.resynth_num_frames(len(w), mode), 64)).astype(np.float32)
        assert np.array_equal(o.resynth(w, m, mode)[0], po.RefG64(v).resynth(w, m)[0]), v


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_long_utterance_q2_wrap():
    """Past 32 767 frames FilterCalc's int16 copy of the frame counter goes negative (SURVEY Q2,
    etsi/cpp/NoiseSup.c:451-475): the restatement follows the reference's code across the wrap, bit for bit."""
    o = po.Oracle()
    n = 33000 * 80 + 5
    x = np.concatenate([synth.utterance(100 + u, 8000, 80000) for u in range(n // 80000 + 1)])[:n]
    a, b = o.etsi_denoise(x, with_ceps=True), po.RefEtsi().trace(x)
    assert np.array_equal(a["den_i16"][:33000 * 80], b["den_i16"][:33000 * 80]) and np.array_equal(a["ceps"], b["ceps"])
    n = 33000 * 160 + 77
    x16 = np.concatenate([synth.utterance(100 + u, 16000, 160000) for u in range(n // 160000 + 1)])[:n].astype(np.float32)
    s, R = o.p16_stream(), po.RefP16()
    A, B = s.func_wiener(x16), R.func_wiener(x16)
    for k in ("out", "sp_var", "sp_spec", "sp_mel", "sp_vadns", "frame_counter"):
        assert np.array_equal(A[k], B[k]), k
    s.close()
    R.close()


def test_reference_is_ill_conditioned_on_degenerate_signals(oracle):
    """Evidence for the self-calibrated bars of tests/test_gpu_fuzz.py: on DC and isolated-click inputs the
    reference algorithm, compiled -O3 with FMA contraction, disagrees with its own -O2/no-contraction build by far
    more than the 1e-4 cepstral bar (log of band energies at the floor, DCT sums that cancel), while on ordinary
    noisy speech the two builds agree to ~1e-5."""
    from conftest import ceps_err
    from test_gpu_fuzz import _signals
    fast = po.Oracle(fast=True)
    xs = _signals(np.random.default_rng(8), 8000, 120, 1.5)        # the sweep test_fuzz_p8 runs
    err = {1: 0.0, 5: 0.0, 8: 0.0}                                  # kinds: DC, isolated clicks, scaled speech
    for u, x in enumerate(xs):
        if u % 10 not in err:
            continue
        a = oracle.etsi_denoise(x, with_ceps=True)["ceps"]
        b = fast.etsi_denoise(x, with_ceps=True)["ceps"]
        assert a.shape == b.shape
        if len(a):
            err[u % 10] = max(err[u % 10], float(ceps_err(a, b).max()))
    assert err[8] < 1e-4 and err[1] > 3e-4 and err[5] > 1e-3, err


def test_ref_bench_loop_is_the_reference(oracle):
    """oracle/ref_bench.c (the timing loop of bench.py's reference arm): the outputs it computes on forked workers are
    those of the direct calls, for any worker count."""
    from speech_enhancement_b200 import api  # noqa: F401  (only for the frame-count formula below)
    xs8 = [synth.utterance(400 + i, 8000, 8000 + 37 * i) for i in range(5)]
    off8 = np.concatenate([[0], np.cumsum([len(x) for x in xs8])]).astype(np.int64)
    rb = po.RefBench("p8")
    t1, c1 = rb.run(np.concatenate(xs8), off8, 1)
    t3, c3 = rb.run(np.concatenate(xs8), off8, 3)
    want = 0.0
    for x in xs8:
        r = oracle.etsi_denoise(x, with_ceps=True)
        want += float(r["den_i16"].astype(np.float64).sum()) + float(r["ceps"].astype(np.float64).sum())
    # (the timed libraries are the -O3 / FMA builds: single LSBs may differ from the -O2 parity build)
    assert abs(c1 - c3) <= 1e-9 * max(1.0, abs(c1))                     # any worker count, same outputs
    assert abs(c1 - want) <= 50 + 1e-4 * abs(want)
    assert t1 > 0 and t3 > 0
    xs16 = [synth.utterance(410 + i, 16000, 16000 + 41 * i) for i in range(3)]
    off16 = np.concatenate([[0], np.cumsum([len(x) for x in xs16])]).astype(np.int64)
    rb = po.RefBench("p16")
    _, c = rb.run(np.concatenate(xs16), off16, 2)
    want = 0.0
    for x in xs16:
        s = oracle.p16_stream()
        r = s.func_wiener(x[:len(x) - len(x) % 160].astype(np.float32))
        s.close()
        want += float(r["out"].astype(np.float64).sum()) + float(r["gains"].astype(np.float64).sum())
    assert abs(c - want) <= 50 + 1e-4 * abs(want)
    rb = po.RefBench("resynth")
    nf = [oracle.resynth_num_frames(len(x), 0) for x in xs16]
    fo = np.concatenate([[0], np.cumsum(nf)]).astype(np.int64)
    mask = (np.random.default_rng(1).random((int(fo[-1]), 64)) > 0.4).astype(np.float32)
    _, c = rb.run(np.concatenate(xs16), off16, 2, mask, fo)
    want = sum(float(oracle.resynth(x, mask[fo[u]:fo[u + 1]].copy(), 0)[0].astype(np.float64).sum()) for u, x in enumerate(xs16))
    assert abs(c - want) <= 50
