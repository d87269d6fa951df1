"""Host-side file formats around the hot path (include/se_io.h): RIFF PCM16, the Kaldi-style mask text the
resynthesis drivers parse (resyth_64sub_IBM/cpp/main.cpp:84-146) and the feature text writers.  CPU only."""
import ctypes as C
import os
import re
import wave

import numpy as np
import pytest

from speech_enhancement_b200 import api, synth

_vp = C.c_void_p


@pytest.fixture(scope="module")
def io():
    L = C.CDLL(api.DROPIN_PATH)
    L.se_wav_read.argtypes = [C.c_char_p, C.c_int, C.POINTER(_vp), C.POINTER(C.c_long), C.POINTER(C.c_int)]
    L.se_wav_write.argtypes = [C.c_char_p, _vp, C.c_long, C.c_int]
    L.se_io_free.argtypes = [_vp]
    L.se_mask_reader_open.restype = _vp
    L.se_mask_reader_open.argtypes = [C.c_char_p]
    L.se_mask_reader_next.argtypes = [_vp, C.c_char_p, C.c_int]
    L.se_mask_reader_rows.argtypes = [_vp, C.c_int, _vp]
    L.se_mask_reader_close.argtypes = [_vp]
    L.se_write_kaldi_matrix.argtypes = [_vp, C.c_char_p, _vp, C.c_int, C.c_int]
    L.se_write_wiener.argtypes = [_vp, _vp, C.c_int]
    L.se_write_sibm.argtypes = [_vp, C.c_char_p, C.c_int, _vp, C.c_int]
    L.se_write_frames_f.argtypes = [_vp, _vp, C.c_int, C.c_int]
    libc = C.CDLL(None)
    libc.fopen.restype = _vp
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [_vp]
    L.libc = libc
    return L


def _read_wav(io, path, channel=0):
    p, n, fs = _vp(), C.c_long(), C.c_int()
    rc = io.se_wav_read(path.encode(), channel, C.byref(p), C.byref(n), C.byref(fs))
    if rc:
        return rc, None, None
    a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_short)), (max(n.value, 1),))[:n.value].copy()
    io.se_io_free(p)
    return 0, a, fs.value


def test_wav_roundtrip_and_python_interop(io, tmp_path):
    x = synth.utterance(0, 16000, 12345)
    p = str(tmp_path / "FDHC0_SI1559_noisy.wav")
    assert io.se_wav_write(p.encode(), x.ctypes.data, len(x), 16000) == 0
    with wave.open(p) as w:                                        # an independent reader agrees
        assert (w.getnchannels(), w.getsampwidth(), w.getframerate(), w.getnframes()) == (1, 2, 16000, len(x))
        assert np.array_equal(np.frombuffer(w.readframes(len(x)), np.int16), x)
    rc, y, fs = _read_wav(io, p)
    assert rc == 0 and fs == 16000 and np.array_equal(x, y)
    # stereo file with an extra chunk before "data": left / right channel selection
    p2 = str(tmp_path / "st.wav")
    with wave.open(p2, "w") as w:
        w.setnchannels(2); w.setsampwidth(2); w.setframerate(8000)
        w.writeframes(np.stack([x[:100], -x[:100]], 1).tobytes())
    assert np.array_equal(_read_wav(io, p2, 0)[1], x[:100]) and np.array_equal(_read_wav(io, p2, 1)[1], -x[:100])
    assert _read_wav(io, p2, 2)[0] != 0 and _read_wav(io, str(tmp_path / "missing.wav"))[0] != 0


def test_kaldi_mask_write_then_parse_like_the_resynth_driver(io, tmp_path):
    rng = np.random.default_rng(1)
    mats = {"FDHC0_SI1559": rng.random((7, 64)).astype(np.float32), "MBPM0_SX317": rng.random((1, 64)).astype(np.float32),
            "FJLM0_SI1043": rng.random((3, 64)).astype(np.float32)}
    path = str(tmp_path / "result.txt")
    fp = io.libc.fopen(path.encode(), b"w")
    for k, m in mats.items():
        assert io.se_write_kaldi_matrix(fp, k.encode(), m.ctypes.data, m.shape[0], 64) == 0
    io.libc.fclose(fp)
    text = open(path).read()
    assert text.startswith("FDHC0_SI1559_noisy [\n") and text.count("]\n") == 3           # show_IBM.cpp:194-209 layout
    assert re.fullmatch(r"(\d\.\d{7} ){64}", text.split("\n")[1])
    r = io.se_mask_reader_open(path.encode())
    assert r
    name = C.create_string_buffer(256)
    for k, m in mats.items():
        assert io.se_mask_reader_next(r, name, 256) == 1 and name.value.decode() == k + "_noisy"
        got = np.zeros_like(m)
        assert io.se_mask_reader_rows(r, m.shape[0], got.ctypes.data) == m.shape[0]
        assert np.max(np.abs(got - m)) <= 1.2e-7                       # "%.7f" text
    assert io.se_mask_reader_next(r, name, 256) == 0
    io.se_mask_reader_close(r)


def test_mask_reader_quirks(io, tmp_path):
    """Row count comes from the caller (the wav length), short rows read as zeros, a header ends a matrix."""
    path = str(tmp_path / "r.txt")
    open(path, "w").write("a_noisy [\n" + " ".join(["0.5"] * 64) + "\n 1 2 3\n" + "b_noisy  [\n" + " ".join(["0.25"] * 64) + " ]\n")
    r = io.se_mask_reader_open(path.encode())
    name = C.create_string_buffer(64)
    assert io.se_mask_reader_next(r, name, 64) == 1 and name.value == b"a_noisy"
    m = np.full((5, 64), -1, np.float32)
    assert io.se_mask_reader_rows(r, 5, m.ctypes.data) == 2                # stopped by the next header
    assert np.all(m[0] == 0.5) and list(m[1, :4]) == [1, 2, 3, 0] and np.all(m[1, 3:] == 0)
    assert io.se_mask_reader_next(r, name, 64) == 1 and name.value == b"b_noisy"
    m = np.zeros((1, 64), np.float32)
    assert io.se_mask_reader_rows(r, 1, m.ctypes.data) == 1 and np.all(m == 0.25)
    io.se_mask_reader_close(r)
    assert not io.se_mask_reader_open(str(tmp_path / "nope.txt").encode())


def test_feature_text_writers(io, tmp_path):
    g = np.random.default_rng(2).random((4, 25)).astype(np.float32)
    ibm = (np.random.default_rng(3).random((6, 64)) > 0.5).astype(np.uint8)
    p = str(tmp_path / "f.txt")
    fp = io.libc.fopen(p.encode(), b"w")
    assert io.se_write_wiener(fp, g.ctypes.data, 4) == 0
    assert io.se_write_sibm(fp, b"ID", 17, ibm.ctypes.data, 6) == 0
    assert io.se_write_frames_f(fp, g.ctypes.data, 2, 25) == 0
    io.libc.fclose(fp)
    lines = open(p).read().split("\n")
    assert lines[0] == "".join("%f " % v for v in g[0])                   # NoiseSup.cpp:1318-1326
    assert lines[4] == "ID_noisy_17 " + "".join("%d " % v for v in ibm[:, 17])
    assert lines[5] == "".join("%f " % v for v in g[0]) and len(lines) == 8
