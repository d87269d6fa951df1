"""Per-call latency of the reference-named adapters (one 10 s utterance per call, the way the reference's unchanged
single-threaded drivers would use them) next to the reference's own code on one host core (oracle/_ref, -O2).
Run on the GPU box:  python tests/tools/percall_latency.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle                         # noqa: E402
from speech_enhancement_b200 import api, synth      # noqa: E402

_vp = C.c_void_p
D = C.CDLL(api.DROPIN_PATH)
D.etsi_denoise.argtypes = [_vp, _vp, C.c_long]
D.se_subband.argtypes = [_vp, C.c_long, _vp]
D.se_resynth.argtypes = [_vp, C.c_long, _vp, C.c_int, C.c_int, _vp]
D.se_single_ibm.argtypes = [_vp, _vp, C.c_long, _vp, _vp]


def med(fn, reps=5):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(ts))


x8 = synth.utterance(1, 8000, 80000)
y8 = np.zeros_like(x8)
x16, clean, noise = synth.utterance(2, 16000, 160000, components=True)
sub = np.zeros(64 * 160000, np.int16)
out16 = np.zeros(160000, np.int16)
nf = api.resynth_num_frames(160000, 0)
mask = (np.random.default_rng(0).random((nf, 64)) > 0.4).astype(np.float32)
irm = np.zeros((nf, 64), np.float32)
ibm = np.zeros((nf, 64), np.uint8)
rows = []
re, rs, rr = pyoracle.RefEtsi(), pyoracle.RefG64("subband"), pyoracle.RefG64("resynth_ibm")
rows.append(("etsi_denoise (8 kHz)", med(lambda: D.etsi_denoise(x8.ctypes.data, y8.ctypes.data, 80000)),
             med(lambda: re.etsi_denoise(x8), 3)))
with api.Context(0) as ctx:
    rows.append(("func_Wiener via se_p16_wiener_batch, 1 utterance", med(lambda: ctx.p16_wiener(x16, [0, 160000])), None))
rp = pyoracle.RefP16()
rows[-1] = (rows[-1][0], rows[-1][1], med(lambda: rp.func_wiener(x16.astype(np.float32)), 3))
rows.append(("subbband (64 rows, exact)", med(lambda: D.se_subband(x16.ctypes.data, 160000, sub.ctypes.data)),
             med(lambda: rs.subband(x16), 2)))
rows.append(("resynth IBM (exact)", med(lambda: D.se_resynth(x16.ctypes.data, 160000, mask.ctypes.data, nf, 0, out16.ctypes.data)),
             med(lambda: rr.resynth(x16, mask), 2)))
with api.Context(0) as ctx:
    fo = [0, nf]
    rows.append(("resynth IBM, default mode (1 s chunks), se_resynth_batch", med(lambda: ctx.resynth(x16, [0, 160000], mask, fo, 0, exact=False)), None))
    rows.append(("resynth IBM, SE_FLAG_LATENCY (2 048-sample chunks)", med(lambda: ctx.resynth(x16, [0, 160000], mask, fo, 0, exact=False, latency=True)), None))
rows.append(("make_single_IBM incl. both subbband passes", med(lambda: D.se_single_ibm(clean.ctypes.data, noise.ctypes.data, 160000, irm.ctypes.data, ibm.ctypes.data)),
             None))
print("| call (one 10 s utterance) | drop-in on the GPU, ms | reference on one host core, ms |")
print("|---|---|---|")
for name, g, c in rows:
    print("| %s | %.1f | %s |" % (name, g, "%.0f" % c if c else "-"))
