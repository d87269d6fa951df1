"""CPU-side checks of the product library: it loads, exports every symbol the headers declare, builds
bit-identical constant tables, fails loudly without a GPU (no CPU fallback), and plans shards."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from speech_enhancement_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{}]*\)\s*;", src)
    return sorted(set(n for n in names if n not in ("defined",)))


def test_c_abi_exports_every_declared_symbol():
    L = C.CDLL(api.LIB_PATH)
    names = _declared("se_b200.h")
    assert len(names) >= 25 and "se_etsi_denoise_batch" in names and "se_resynth_batch" in names
    for n in names:
        assert hasattr(L, n), "libse_b200.so does not export %s" % n


def test_dropin_exports_reference_names():
    L = C.CDLL(api.DROPIN_PATH)
    names = sorted(set(_declared("se_reference_api.h") + _declared("se_io.h")))
    for must in ("etsi_denoise", "rfft", "DoAdvProcess", "etsi_denoise_mapping_func_Wiener", "se_resynth", "se_subband", "se_wav_read",
                 "se_mask_reader_rows", "se_write_kaldi_matrix"):
        assert must in names
    for n in names:
        assert hasattr(L, n), "libse_b200_dropin.so does not export %s" % n


def test_no_cpu_fallback():
    """Without a CUDA device every compute entry point refuses to run (SE_ENODEVICE)."""
    L = api.lib()
    if L.se_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(api.SeError) as e:
        api.Context(0)
    assert "(-2)" in str(e.value) and "no CPU path" in str(e.value)
    # the drop-in adapters report failure the reference's way (TRUE = failure)
    D = C.CDLL(api.DROPIN_PATH)
    x = np.zeros(800, np.int16)
    y = np.zeros(800, np.int16)
    D.etsi_denoise.argtypes = [C.c_void_p, C.c_void_p, C.c_long]
    assert D.etsi_denoise(x.ctypes.data, y.ctypes.data, 800) == 1


def test_product_does_not_link_the_oracle():
    import subprocess
    for so in (api.LIB_PATH, api.DROPIN_PATH):
        out = subprocess.run(["ldd", so], capture_output=True, text=True).stdout
        assert "oracle" not in out and "libref_" not in out
        sym = subprocess.run(["nm", "-D", so], capture_output=True, text=True).stdout
        assert " so_" not in sym and "ref_etsi" not in sym


def test_tables_bit_identical_to_oracle(oracle):
    """Host-built tables (csrc/se_tables.cpp) == the oracle's == the reference's (golden p8/p16 tables)."""
    for prof, key in ((8, "p8"), (16, "p16")):
        t = oracle.ns_tables(prof)
        assert np.array_equal(api.get_table(key + ".sig_window"), t["sig_window"])
        assert np.array_equal(api.get_table(key + ".ir_window"), t["ir_window"])
        assert np.array_equal(api.get_table(key + ".idct").reshape(25, 25), t["idct"])
        bank = api.get_table(key + ".bank").reshape(25, -1)
        if prof == 8:
            ln = api.get_table("p8.bank_len").astype(int)
            assert np.array_equal(ln, t["bank_len"])
            assert np.array_equal(api.get_table("p8.bank_start").astype(int), t["bank_start"])
            for j in range(25):
                assert np.array_equal(bank[j, :ln[j]], t["bank"][j, :ln[j]])
        else:
            assert np.array_equal(bank, t["bank"])
    c = oracle.cc_tables()
    assert np.array_equal(api.get_table("cc.hamming"), c["hamming"])
    assert np.array_equal(api.get_table("cc.dct").reshape(12, 23), c["dct"])
    st, ln = api.get_table("cc.start").astype(int), api.get_table("cc.len").astype(int)
    assert np.array_equal(st, c["start"]) and np.array_equal(ln, c["len"])
    tri = api.get_table("cc.tri").reshape(23, 129)
    for j in range(23):
        assert np.array_equal(tri[j, :ln[j]], c["tri"][j, :ln[j]])
    h = oracle.hp16k_tables()                                     # ETSI 16 kHz mode
    assert np.array_equal(api.get_table("hp16.dp"), h["dp"])
    assert np.array_equal(api.get_table("hp16.fb").reshape(5, 65), h["fb"])
    assert np.array_equal(api.get_table("hp16.fb_start").astype(int), h["start"])
    assert np.array_equal(api.get_table("hp16.fb_len").astype(int), h["len"])
    assert np.array_equal(api.get_table("hp16.dct26").reshape(12, 26), h["dct26"])
    assert np.array_equal(h["hp"], -h["dp"] * (-1.0) ** np.arange(118))       # the high-pass is the sign-alternated prototype
    cf, bw, mid = oracle.channel_params()
    assert np.array_equal(api.get_table("gt.cf"), cf)
    assert np.array_equal(api.get_table("gt.bw"), bw)
    assert np.array_equal(api.get_table("gt.mid"), mid)


def test_frame_count_helpers(oracle):
    for n in (0, 79, 80, 479, 480, 560, 561, 80000):
        assert api.etsi_num_ceps_frames(n) == max(0, n // 80 - 6)
    for n in (0, 319, 320, 479, 480, 160000):
        assert api.ibm_num_frames(n) == (0 if n < 320 else (n - 320) // 160 + 1)
        assert api.resynth_num_frames(n, api.RESYNTH_1DNN) == n // 160
    for n in (320, 481, 160000):
        for mode in (0, 1, 2):
            assert api.resynth_num_frames(n, mode) == oracle.resynth_num_frames(n, mode)


def test_shard_plan_properties():
    rng = np.random.default_rng(0)
    lens = rng.integers(1, 160000, 1000)
    off = np.concatenate([[0], np.cumsum(lens)])
    for k in (1, 2, 4, 8):
        p = api.shard_plan(off, k)
        assert p[0] == 0 and p[-1] == 1000 and np.all(np.diff(p) >= 0)
        loads = np.array([off[p[i + 1]] - off[p[i]] for i in range(k)])
        assert loads.sum() == off[-1]
        assert loads.max() - loads.min() <= 2 * 160000          # balanced to within one utterance each side
    assert list(api.shard_plan([0, 10], 4)) == [0, 0, 1, 1, 1]  # fewer utterances than shards: empty shards
    assert list(api.shard_plan([0], 3)) == [0, 0, 0, 0]


def test_rms_normalise_matches_the_tools_arithmetic():
    """se_rms_normalise = aurora_speech_enhancement.cpp:172-184 (host-side caller code, no GPU needed)."""
    D = C.CDLL(api.DROPIN_PATH)
    D.se_rms_normalise.restype = C.c_double
    D.se_rms_normalise.argtypes = [C.c_void_p, C.c_int, C.c_float]
    rng = np.random.default_rng(5)
    for scale in (30.0, 900.0, 9000.0):
        p = np.clip(rng.standard_normal(5000) * scale, -32768, 32767).astype(np.int16)
        q = p.copy()
        f = D.se_rms_normalise(q.ctypes.data, len(q), 3000.0)
        energy = np.sqrt(np.sum(p.astype(np.int64) ** 2).astype(np.float64) / len(p))
        assert abs(f - 3000.0 / energy) <= 1e-12 * f
        want = np.trunc(p.astype(np.float32).astype(np.float64) * f).astype(np.int64)
        want = ((want + 32768) % 65536 - 32768).astype(np.int16)       # (short) wraps like the reference's cast
        assert np.array_equal(q, want)
    z = np.zeros(16, np.int16)
    D.se_rms_normalise(z.ctypes.data, 16, 3000.0)                      # 0 * inf = NaN -> cvttsd2si -> low 16 bits 0
    assert np.all(z == 0)


def build_cxx_driver(out_dir):
    """tests/cxx/dropin_driver.cpp: a stand-in for the reference's C++ drivers over include/se_reference_api.hpp."""
    import subprocess
    exe = os.path.join(str(out_dir), "dropin_driver")
    lib = os.path.dirname(api.DROPIN_PATH)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cxx", "dropin_driver.cpp"), "-L" + lib, "-lse_b200_dropin",
                           "-lse_b200", "-Wl,-rpath," + lib, "-o", exe])
    return exe


def test_cxx_adapters_compile_and_link(tmp_path):
    """subbband / resynth / addnoise with the reference's C++ signatures (CWave*, FILESNAME, vector<vector<float>>&)
    compile against a stand-in wave class and resolve against libse_b200_dropin.so."""
    exe = build_cxx_driver(tmp_path)
    assert os.path.exists(exe)
