"""The reference-named entry points of libse_b200_dropin.so (include/se_reference_api.h), driven the way
the reference's own drivers drive them (etsi/cpp/main.cpp:55-66, aurora_etsi_test.cpp:16-141)."""
import ctypes as C

import numpy as np
import pytest

from conftest import ceps_err, rel_err, snr_db
from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu
_vp = C.c_void_p


@pytest.fixture(scope="module")
def dropin(ctx):
    L = C.CDLL(api.DROPIN_PATH)
    L.etsi_denoise.argtypes = [_vp, _vp, C.c_long]
    L.etsi_denoise_synchronization.argtypes = [_vp, _vp, C.c_long]
    L.rfft.argtypes = [_vp, C.c_int, C.c_int]
    L.rfft.restype = None
    L.AdvProcessAlloc.restype = _vp
    L.AdvProcessAlloc.argtypes = [C.c_int]
    L.AdvProcessInit.argtypes = [_vp]
    L.DoAdvProcess.argtypes = [_vp, _vp, _vp, _vp]
    L.se_adv_feature_ready.argtypes = [_vp]
    L.DoNoiseSupInit.argtypes = [_vp]
    L.DoNoiseSup.argtypes = [_vp, _vp, _vp]
    L.DoCompCeps.argtypes = [_vp, _vp, _vp]
    L.AdvProcessDelete.argtypes = [C.POINTER(_vp)]
    L.FlushAdvProcess.argtypes = [_vp, _vp]
    L.DoNoiseSupAlloc.restype = _vp
    L.DoNoiseSupDelete.argtypes = [_vp]
    L.etsi_denoise_mapping_global_init.argtypes = [C.POINTER(_vp), _vp]
    L.etsi_denoise_mapping_thread_init.argtypes = [C.POINTER(_vp), _vp]
    L.etsi_denoise_mapping_func_Wiener.argtypes = [_vp] * 5
    L.etsi_denoise_mapping_thread_release.argtypes = [C.POINTER(_vp)]
    L.etsi_denoise_mapping_global_release.argtypes = [C.POINTER(_vp)]
    L.se_resynth.argtypes = [_vp, C.c_long, _vp, C.c_int, C.c_int, _vp]
    L.se_subband.argtypes = [_vp, C.c_long, _vp]
    return L


def test_etsi_denoise_signature_and_tail(dropin, oracle):
    x = synth.utterance(1, 8000, 80 * 120 + 33)
    out = np.full(len(x), 12345, np.int16)
    rc = dropin.etsi_denoise(x.ctypes.data, out.ctypes.data, len(x))
    assert rc == 0                                     # FALSE = success (AdvFrontEnd.c:205)
    ref = oracle.etsi_denoise(x)
    body = 80 * 120
    assert np.abs(out[:body].astype(np.int32) - ref["den_i16"][:body]).max() <= 1
    assert np.all(out[:320] == 0)                      # 4-call latency
    assert np.all(out[body:] == 12345)                 # tail untouched, like the reference
    assert dropin.etsi_denoise_16k(x.ctypes.data, out.ctypes.data, len(x)) == 1


def test_rfft_in_place(dropin, oracle):
    rng = np.random.default_rng(0)
    x = (rng.standard_normal(256) * 500).astype(np.float32)
    y = x.copy()
    dropin.rfft(y.ctypes.data, 256, 8)
    ref = oracle.rfft(x, 8)
    assert np.max(np.abs(ref - y)) <= 1e-4 * np.max(np.abs(ref))


def test_do_adv_process_frame_by_frame(dropin, oracle):
    x = synth.utterance(2, 8000, 80 * 40)
    ref = oracle.etsi_denoise(x, with_ceps=True)
    h = _vp(dropin.AdvProcessAlloc(8000))
    assert h.value
    dropin.AdvProcessInit(h)
    outs, feats, rets = [], [], []
    for b in range(40):
        frame = np.ascontiguousarray(x[b * 80:(b + 1) * 80])
        out = np.zeros(80, np.int16)
        feat = np.zeros(14, np.float32)
        rets.append(dropin.DoAdvProcess(frame.ctypes.data, out.ctypes.data, feat.ctypes.data, h))
        outs.append(out)
        if dropin.se_adv_feature_ready(h):
            feats.append(feat)
    dropin.AdvProcessDelete(C.byref(h))
    assert h.value is None
    assert rets == [0] * 4 + [1] * 36                  # TRUE once a shift is output (NoiseSup.c:1436-1439)
    got = np.concatenate(outs)
    assert np.abs(got.astype(np.int32) - ref["den_i16"]).max() <= 1
    assert len(feats) == len(ref["ceps"]) == 34
    assert ceps_err(ref["ceps"], np.array(feats)).max() <= 1e-4


@pytest.mark.parametrize("tag", ["z", "s", "m"])
def test_do_adv_process_and_flush_as_shipped(dropin, tag):
    """DoAdvProcess frame by frame and the FlushAdvProcess loop against the reference ITSELF driven the same way
    (tests/golden/adv_shipped.npz from oracle/_ref): which calls write DenoiseBuffer / FeatureBuffer and what they
    leave there.  z = 5 leading all-zero frames (TRUE, FeatureBuffer[0..13] = 0, DenoiseBuffer untouched,
    ParmInterface.c:314-327); s / m = 9 / 20 calls (FrameCounter below / above the flush's write threshold of 10)."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "adv_shipped.npz"))
    x = g[tag + "_x"]
    calls = len(x) // 80
    h = _vp(dropin.AdvProcessAlloc(8000))
    dropin.AdvProcessInit(h)
    rets = []
    for b in range(calls):
        frame = np.ascontiguousarray(x[b * 80:(b + 1) * 80])
        out = np.full(80, 0x7a7a, np.int16)
        feat = np.full(15, 7.0, np.float32)
        rets.append(dropin.DoAdvProcess(frame.ctypes.data, out.ctypes.data, feat.ctypes.data, h))
        wrote = bool(np.any(out != 0x7a7a))
        assert wrote == bool(g[tag + "_wrote"][b]), "call %d" % b
        ref_out = g[tag + "_out"][b * 80:(b + 1) * 80]
        if wrote:
            assert np.abs(out.astype(np.int32) - ref_out).max() <= 1
        ref_feat = g[tag + "_feat"][b]
        zero_frame = not np.any(frame) and not wrote and ref_feat[0] == 0.0
        if zero_frame:                       # null MFCC vector: [0..13] zeroed, [14] untouched; returns TRUE
            assert rets[-1] == 1 and not feat[:14].any() and feat[14] == 7.0
            assert np.array_equal(feat, ref_feat)
        elif not wrote:
            assert rets[-1] == 0 and np.array_equal(feat, ref_feat)      # latency: nothing touched
        else:
            assert rets[-1] == 1
    fr, ff = [], []
    for k in range(8):
        feat = np.full(15, 7.0, np.float32)
        fr.append(dropin.FlushAdvProcess(feat.ctypes.data, h))
        ff.append(feat)
        if not fr[-1]:
            break
    dropin.AdvProcessDelete(C.byref(h))
    assert fr == list(g[tag + "_flush_rets"])
    assert np.array_equal(np.array(ff), g[tag + "_flush_feat"])


def test_noise_sup_alloc_delete_by_name(dropin):
    """NoiseSupExports.h:20,26 by name: a NoiseSup state can be allocated and released on its own."""
    n = dropin.DoNoiseSupAlloc()
    assert n
    dropin.DoNoiseSupDelete(n)
    dropin.DoNoiseSupDelete(None)


def test_handles_used_from_other_threads(dropin, oracle):
    """Handles are created on one host thread and driven from others (the reference's handles are plain heap
    structs): every call holds its context's lock, so pushes from 4 threads on 4 handles of ONE context -- while the
    creating thread keeps calling etsi_denoise on the same context -- give the single-threaded results."""
    import threading
    xs = [synth.utterance(70 + i, 8000, 80 * 30) for i in range(4)]
    hs = [_vp(dropin.AdvProcessAlloc(8000)) for _ in xs]
    for h in hs:
        dropin.AdvProcessInit(h)
    got = [None] * 4
    err = []

    def work(i):
        try:
            outs = []
            for b in range(30):
                frame = np.ascontiguousarray(xs[i][b * 80:(b + 1) * 80])
                out = np.zeros(80, np.int16)
                dropin.DoAdvProcess(frame.ctypes.data, out.ctypes.data, None, hs[i])
                outs.append(out)
            got[i] = np.concatenate(outs)
        except Exception as e:      # pragma: no cover
            err.append(e)

    ths = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in ths:
        t.start()
    big = synth.utterance(80, 8000, 80000)
    big_out = np.zeros_like(big)
    for _ in range(3):                                   # the creating thread's own context stays busy meanwhile
        assert dropin.etsi_denoise(big.ctypes.data, big_out.ctypes.data, len(big)) == 0
    for t in ths:
        t.join()
    assert not err
    for i, x in enumerate(xs):
        ref = oracle.etsi_denoise(x)["den_i16"]
        assert np.abs(got[i].astype(np.int32) - ref).max() <= 1
    assert np.abs(big_out.astype(np.int32) - oracle.etsi_denoise(big)["den_i16"]).max() <= 1
    for h in hs:
        dropin.AdvProcessDelete(C.byref(h))


def test_plugin_api_like_extract_wiener(dropin, oracle, tmp_path):
    glb, thd = _vp(), _vp()
    assert dropin.etsi_denoise_mapping_global_init(C.byref(glb), None) == 1
    assert dropin.etsi_denoise_mapping_thread_init(C.byref(thd), glb) == 1

    class In(C.Structure):
        _fields_ = [("inData", _vp), ("dataNum", C.c_int)]

    class Out(C.Structure):
        _fields_ = [("outData", _vp)] + [(k, _vp) for k in ("var", "spec", "mel", "vadns", "fc")]

    libc = C.CDLL(None)
    libc.fopen.restype = _vp
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [_vp]
    s = oracle.p16_stream()
    for chan in range(2):                              # state carries over between channel files
        wav = synth.utterance(40 + chan, 16000, 160 * 30 + 17)
        num = len(wav) - len(wav) % 160                # "wipe the tail", aurora_etsi_test.cpp:100-101
        x = wav[:num].astype(np.float32)
        nb = num // 160
        out = np.zeros(num, np.float32)
        fl = [np.zeros(nb, np.int32) for _ in range(5)]
        path = str(tmp_path / ("c%d.wiener" % chan)).encode()
        fp = libc.fopen(path, b"w")
        i, o = In(x.ctypes.data, num), Out(out.ctypes.data, *[f.ctypes.data for f in fl])
        rc = dropin.etsi_denoise_mapping_func_Wiener(glb, thd, C.addressof(i), C.addressof(o), fp)
        libc.fclose(fp)
        assert rc == 0
        ref = s.func_wiener(x)
        assert snr_db(ref["out"], out) >= 60.0
        assert np.array_equal(fl[4], ref["frame_counter"])
        rows = [l.split() for l in open(path.decode()).read().strip().split("\n")]
        assert len(rows) == len(ref["gains"]) and all(len(r) == 25 for r in rows)
        txt = np.array(rows, np.float64)
        assert np.abs(txt - ref["gains"]).max() <= 2e-6 + 1e-4 * np.abs(ref["gains"]).max()
    s.close()
    dropin.etsi_denoise_mapping_thread_release(C.byref(thd))
    dropin.etsi_denoise_mapping_global_release(C.byref(glb))
    assert thd.value is None and glb.value is None


def test_resynth_c_twin_marks_mask(dropin, oracle):
    x = synth.utterance(3, 16000, 3200)
    nf = api.resynth_num_frames(len(x), 0)
    m = np.random.default_rng(5).random((nf, 64)).astype(np.float32)
    m0 = m.copy()
    out = np.zeros(len(x), np.int16)
    assert dropin.se_resynth(x.ctypes.data, len(x), m.ctypes.data, nf, 0, out.ctypes.data) == 0
    ref, mref = oracle.resynth(x, m0, 0)
    assert np.array_equal(out, ref)
    assert np.array_equal(m, mref)                     # > 0.5 entries overwritten with 1.0 (extractwav.cpp:99)


def test_module_entry_points_do_noise_sup_and_comp_ceps(dropin, oracle):
    """DoNoiseSup / DoCompCeps called directly, the way DoAdvProcess reaches them through FEParamsX
    (etsi/cpp/ParmInterface.c:260-293): float shifts in, BOOLEAN latency semantics, cepstra from frameBuf + 1."""
    from conftest import ceps_err, snr_db
    x = synth.utterance(9, 8000, 80 * 40)
    x[:160] = 0                                              # no zero-frame gate at this level: processed like any frame
    h = _vp(dropin.AdvProcessAlloc(8000))
    assert h.value
    dropin.DoNoiseSupInit(h)
    s = oracle.L.so_ns_create(8)
    hist = np.zeros(241, np.float32)                          # what BufInShiftToPut / BufInGetLast keep (BufferIn.c:157-205)
    n_out = 0
    for c in range(len(x) // 80):
        frame = x[c * 80:(c + 1) * 80].astype(np.float32)
        out = np.full(80, np.nan, np.float32)
        ref = np.zeros(80, np.float32)
        got = dropin.DoNoiseSup(frame.ctypes.data, out.ctypes.data, h)
        exp = oracle.L.so_ns_block(s, frame.ctypes.data, ref.ctypes.data, None)
        assert got == exp, c
        if not got:
            assert np.all(np.isnan(out))                     # latency: OutData untouched
            continue
        assert snr_db(ref, out) >= 60.0 or not ref.any()
        n_out += 1
        hist = np.concatenate([hist[80:], ref])
        if n_out >= 3:                                        # frameBuf + 1 = the last 200 samples, [-1] valid
            buf = np.ascontiguousarray(hist[241 - 201:])
            coef = np.zeros(14, np.float32)
            assert dropin.DoCompCeps(buf.ctypes.data + 4, coef.ctypes.data, h) == 1
            assert ceps_err(oracle.comp_ceps(buf), coef).max() <= 1e-4
    assert n_out == 40 - 4
    oracle.L.so_ns_destroy(s)
    dropin.AdvProcessDelete(C.byref(h))


def _buffer_denoise_restated(etsi_denoise, src, bl, work):
    """aurora_speech_enhancement.cpp:25-79 line by line over any etsi_denoise(x) -> int16[len - len%80]."""
    n = len(src)
    des = np.full(n, 777, np.int16)
    times, left = n // bl, n % bl

    def run(lo, ln):
        y = etsi_denoise(src[lo:lo + ln])
        work[:len(y)] = y                                # the tail ln % 80 of `work` keeps what it held

    if times == 0:
        run(0, n)
        des[:n - 320] = work[320:n]
    else:
        run(0, bl)
        des[:bl - 320] = work[320:bl]
        for i in range(1, times):
            run(i * bl - 320, bl + 320)
            des[i * bl - 320:(i + 1) * bl - 320] = work[320:bl + 320]
        run(times * bl - 320, left + 320)
        des[times * bl - 320:times * bl - 320 + left] = work[320:320 + left]
    des[n - 320:] = 0
    return des


@pytest.mark.parametrize("n,bl", [(8000 * 7 + 3333, 16000), (8000 * 7 + 3333, 80000), (16000 * 3, 16000), (4321, 8000)])
def test_buffer_denoise_windows_as_one_batch(dropin, oracle, n, bl):
    """BufferDenoise (the Win32 tool's 3-minute windows with 320 samples of overlap) with small windows: all
    windows run as ONE ragged GPU batch; result = the reference's sequential loop, stale `work` tail included."""
    dropin.BufferDenoise.argtypes = [_vp, _vp, C.c_int, C.c_int, _vp]
    dropin.BufferDenoise.restype = None
    dropin.se_buffer_denoise.argtypes = [_vp, _vp, C.c_int, C.c_int]
    x = synth.utterance(9, 8000, n)
    des = np.full(n, 777, np.int16)
    work = np.full(bl + 320, -5, np.int16)
    dropin.BufferDenoise(x.ctypes.data, des.ctypes.data, n, bl, work.ctypes.data)

    def ours(seg):                                      # the drop-in's own per-call etsi_denoise (GPU)
        y = np.zeros(len(seg), np.int16)
        seg = np.ascontiguousarray(seg)
        assert dropin.etsi_denoise(seg.ctypes.data, y.ctypes.data, len(seg)) == 0
        return y[:len(seg) - len(seg) % 80]

    seq = _buffer_denoise_restated(ours, x, bl, np.full(bl + 320, -5, np.int16))
    assert np.array_equal(des, seq)                     # batched windows == sequential windows, bit for bit
    ref = _buffer_denoise_restated(lambda s: oracle.etsi_denoise(s)["den_i16"][:len(s) - len(s) % 80], x, bl,
                                   np.full(bl + 320, -5, np.int16))
    assert np.abs(des.astype(np.int32) - ref.astype(np.int32)).max() <= 1
    assert np.all(des[n - 320:] == 0)
    des2 = np.zeros(n, np.int16)
    assert dropin.se_buffer_denoise(x.ctypes.data, des2.ctypes.data, n, bl) == 0
    stale = (n % bl + 320) % 80 if n >= bl else n % 80   # only the < 80 stale samples depend on `work`'s history
    keep = np.ones(n, bool)
    keep[n - 320 - stale:n - 320] = False
    assert np.array_equal(des2[keep], des[keep])
    assert dropin.se_buffer_denoise(x.ctypes.data, des2.ctypes.data, 100, bl) != 0      # < 320 samples: rejected


@pytest.mark.parametrize("mode", [0, 2])
def test_cxx_signatures_drive_the_gpu(ctx, tmp_path, mode):
    """include/se_reference_api.hpp: subbband(CWave*, path, tail, id), resynth(CWave*, FILESNAME, vector<vector<float>>&, id)
    and addnoise(CWave*, CWave*, CWave*, id, FILESNAME) compiled into a C++ driver; the files it writes (the reference's
    names: <id><tail>_<chan>.wav, <id>_e_resynth.wav, <id>_noisy.wav) hold exactly what the batch API returns."""
    import subprocess
    from test_abi_host import build_cxx_driver
    exe = build_cxx_driver(tmp_path)
    n = 6000
    x, clean, noise = synth.utterance(41, 16000, n, components=True)
    nf = api.resynth_num_frames(n, mode)
    mask = np.random.default_rng(3).random((nf, 64)).astype(np.float32)
    for d in ("sub", "res", "mix", "pure", "subp", "subn"):
        (tmp_path / d).mkdir()
    x.tofile(tmp_path / "x.i16"); noise.tofile(tmp_path / "n.i16"); mask.tofile(tmp_path / "m.f32")
    clean.tofile(tmp_path / "c.i16")
    subprocess.check_call([exe, str(tmp_path / "x.i16"), str(tmp_path / "n.i16"), str(tmp_path / "m.f32"), str(nf), str(mode),
                           str(tmp_path) + "/"] + ([str(tmp_path / "c.i16")] if mode == 0 else []))
    if mode == 0:
        # make_single_IBM(char*, FILE*, FILESNAME) (make_IBM.h:9): the text matrix it appends holds the IRM of the fused
        # batch call, digit for digit; gammaToneFilter / hairCell (HuWang.h:49,62) with a by-value, non-preset channel
        irm, _, _ = ctx.ibm(clean, noise, [0, n])
        txt = open(tmp_path / "irm.txt").read()
        assert txt.startswith("utt1_noisy [\n") and txt.endswith("]\n")
        rows = txt[len("utt1_noisy [\n"):-2].split("\n ")
        assert len(rows) == len(irm)
        want_rows = ["".join("%.7f " % v for v in r) for r in irm]
        assert rows == want_rows
        from oracle import pyoracle
        o = pyoracle.Oracle()
        gt = np.fromfile(tmp_path / "gt.f32", np.float32)
        ref_gt = o.gammatone(x.astype(np.float32), 1234.5, 150.0, 0.7)
        assert np.array_equal(gt, ref_gt)
        assert np.array_equal(np.fromfile(tmp_path / "hc.f32", np.float32), o.haircell(ref_gt))
    sub = ctx.subband(x, [0, n], exact=True).reshape(64, n)
    for chan in (0, 17, 63):
        got = np.fromfile(tmp_path / "sub" / ("utt1_t_%d.wav" % chan), np.int16)
        assert np.array_equal(got, sub[chan])
    want = ctx.resynth(x, [0, n], mask, [0, nf], mode)
    assert np.array_equal(np.fromfile(tmp_path / "res" / "utt1_e_resynth.wav", np.int16), want)
    after = np.fromfile(tmp_path / "res" / "mask_after.f32", np.float32).reshape(nf, 64)
    assert np.array_equal(after, np.where(mask > 0.5, np.float32(1.0), mask) if mode == 0 else mask)
    an, ay = ctx.addnoise(x, noise, [0, n], 5)
    assert np.array_equal(np.fromfile(tmp_path / "mix" / "utt1_noisy.wav", np.int16), ay)
    assert np.array_equal(np.fromfile(tmp_path / "mix" / "noise_scaled.i16", np.int16), an)


def test_concurrent_callers_get_their_own_context(dropin, oracle):
    """The reference's batch tool calls etsi_denoise from N worker threads (aurora_speech_enhancement.cpp:110-120,
    312-328).  Every calling thread of the drop-in gets its own context and stream: results are those of the
    single-threaded calls, bit for bit, and the calls overlap on the GPU."""
    import threading
    import time
    xs = [synth.utterance(60 + i, 8000, 80000) for i in range(8)]
    want = []
    t0 = time.perf_counter()
    for x in xs:
        y = np.zeros_like(x)
        assert dropin.etsi_denoise(x.ctypes.data, y.ctypes.data, len(x)) == 0
        want.append(y)
    t_seq = time.perf_counter() - t0
    got = [np.zeros_like(x) for x in xs]
    rcs = [None] * len(xs)

    def work(i):
        for _ in range(3):                               # the first call of a thread creates its context
            rcs[i] = dropin.etsi_denoise(xs[i].ctypes.data, got[i].ctypes.data, len(xs[i]))

    th = [threading.Thread(target=work, args=(i,)) for i in range(len(xs))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert rcs == [0] * len(xs)
    for a, b in zip(want, got):
        assert np.array_equal(a, b)
    # warmed-up threads (the first call of a thread creates its context): 8 concurrent calls vs 8 sequential ones
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(8) as ex:
        list(ex.map(work, range(8)))
        t0 = time.perf_counter()
        list(ex.map(lambda i: dropin.etsi_denoise(xs[i].ctypes.data, got[i].ctypes.data, len(xs[i])), range(8)))
        t_par = time.perf_counter() - t0
        # the same for the sub-band path: se_resynth per thread
        w = [synth.utterance(70 + i, 16000, 48000) for i in range(8)]
        nf = api.resynth_num_frames(48000, 0)
        masks = [np.random.default_rng(i).random((nf, 64)).astype(np.float32) for i in range(8)]
        outs = [np.zeros(48000, np.int16) for _ in range(8)]

        def rs(i):
            m = masks[i].copy()
            return dropin.se_resynth(w[i].ctypes.data, 48000, m.ctypes.data, nf, 0, outs[i].ctypes.data)

        assert list(ex.map(rs, range(8))) == [0] * 8
        par = [o.copy() for o in outs]
    for i in range(8):
        assert rs(i) == 0 and np.array_equal(outs[i], par[i])
    print("8 x 10 s utterances: sequential %.1f ms, 8 threads %.1f ms" % (t_seq * 1e3, t_par * 1e3))
    assert t_par < 0.6 * t_seq


def test_resynth_adapter_fast_mode_env(dropin, tmp_path):
    """SE_B200_FAST=1 (read once per process) switches se_resynth to the time-chunked default-mode kernels: within
    1 LSB of the bit-exact adapter output."""
    import os
    import subprocess
    import sys
    n = 64000
    x = synth.utterance(81, 16000, n)
    nf = api.resynth_num_frames(n, 1)
    mask = np.random.default_rng(8).random((nf, 64)).astype(np.float32)
    exact = np.zeros(n, np.int16)
    assert dropin.se_resynth(x.ctypes.data, n, mask.copy().ctypes.data, nf, 1, exact.ctypes.data) == 0
    x.tofile(tmp_path / "x.i16"); mask.tofile(tmp_path / "m.f32")
    code = ("import ctypes as C, numpy as np, sys\n"
            "D = C.CDLL(sys.argv[1]); D.se_resynth.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_int, C.c_int, C.c_void_p]\n"
            "x = np.fromfile(sys.argv[2] + '/x.i16', np.int16); m = np.fromfile(sys.argv[2] + '/m.f32', np.float32)\n"
            "o = np.zeros(len(x), np.int16)\n"
            "assert D.se_resynth(x.ctypes.data, len(x), m.ctypes.data, len(m) // 64, 1, o.ctypes.data) == 0\n"
            "o.tofile(sys.argv[2] + '/o.i16')\n")
    subprocess.check_call([sys.executable, "-c", code, api.DROPIN_PATH, str(tmp_path)], env=dict(os.environ, SE_B200_FAST="1"))
    fast = np.fromfile(tmp_path / "o.i16", np.int16)
    d = np.abs(fast.astype(np.int32) - exact.astype(np.int32))
    assert d.max() <= 1
