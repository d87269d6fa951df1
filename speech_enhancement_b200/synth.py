"""Deterministic synthetic noisy-speech corpus — host-side input generator for
tests and bench.py.  Thin ctypes view of csrc/synth.c (xorshift64*, seed 0x5EED0000+u)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "csrc", "libse_synth.so")
_lib = None


def _load():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "csrc", "synth.c")
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-o", _SO, src, "-lm"])
        _lib = C.CDLL(_SO)
        _lib.se_synth_length.restype = C.c_long
        _lib.se_synth_length.argtypes = [C.c_long, C.c_int, C.c_double, C.c_double]
        _lib.se_synth_utterance.argtypes = [C.c_long, C.c_int, C.c_long, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.se_synth_soft_mask.argtypes = [C.c_long, C.c_void_p, C.c_long, C.c_void_p]
    return _lib


def utterance(u, fs, n, components=False):
    """int16 noisy utterance u of n samples at fs Hz; with components=True also (clean, noise)."""
    L = _load()
    noisy = np.empty(n, np.int16)
    if components:
        clean = np.empty(n, np.int16)
        noise = np.empty(n, np.int16)
        L.se_synth_utterance(u, fs, n, noisy.ctypes.data, clean.ctypes.data, noise.ctypes.data)
        return noisy, clean, noise
    L.se_synth_utterance(u, fs, n, noisy.ctypes.data, None, None)
    return noisy


def ragged_length(u, fs, min_s=2.0, max_s=10.0):
    return int(_load().se_synth_length(u, fs, min_s, max_s))


def corpus(first_u, count, fs, seconds=None, min_s=2.0, max_s=10.0, threads=1):
    """Concatenated ragged batch: (samples int16[total], offsets int64[count+1]).  threads > 1 generates
    utterances in parallel (the C generator releases the GIL; results do not depend on the thread count)."""
    L = _load()
    lens = [int(seconds * fs) if seconds is not None else ragged_length(first_u + i, fs, min_s, max_s)
            for i in range(count)]
    offsets = np.zeros(count + 1, np.int64)
    offsets[1:] = np.cumsum(lens)
    samples = np.empty(int(offsets[-1]), np.int16)

    def gen(i):
        L.se_synth_utterance(first_u + i, fs, lens[i], samples[offsets[i]:].ctypes.data, None, None)

    if threads > 1:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(gen, range(count)))
    else:
        for i in range(count):
            gen(i)
    return samples, offsets


def soft_mask(u, irm):
    irm = np.ascontiguousarray(irm, np.float32)
    out = np.empty_like(irm)
    _load().se_synth_soft_mask(u, irm.ctypes.data, irm.size, out.ctypes.data)
    return out
