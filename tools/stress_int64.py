"""More than 2^31 samples in ONE batch call (64-bit offsets end to end): 30 000 x 10 s 8 kHz through
se_etsi_denoise_batch with host buffers (pipelined chunks); spot-check utterances against single-utterance calls."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_enhancement_b200 import api, synth      # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
T = max(1, min(32, os.cpu_count() or 1))
samples, offsets = synth.corpus(100000, N, 8000, seconds=10.0, threads=T)
print("total samples", int(offsets[-1]), "> 2^31:", int(offsets[-1]) > 2 ** 31)
ctx = api.Context(0)
r = ctx.etsi_denoise(samples, offsets, ceps=True)
assert int(r["n_ceps"].min()) >= 80000 // 80 - 11
for u in (0, 1, N // 2, N - 27000 if N > 27000 else N // 3, N - 2, N - 1):
    lo, hi = int(offsets[u]), int(offsets[u + 1])
    a = ctx.etsi_denoise(samples[lo:hi], [0, hi - lo], ceps=True)
    assert np.array_equal(a["den_i16"], r["den_i16"][lo:hi]), u
    co = int(r["ceps_offsets"][u])
    assert np.array_equal(a["ceps"][:int(a["n_ceps"][0])], r["ceps"][co:co + int(r["n_ceps"][u])]), u
print("ok: utterances beyond sample 2^31 equal their single-utterance results")
