"""Per-signal-kind parity report of the edge-case sweep in tests/test_gpu_fuzz.py (run on the GPU box)."""
import os
import sys
from collections import defaultdict

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ceps_err, rel_err, snr_db      # noqa: E402
from oracle import pyoracle                         # noqa: E402
from speech_enhancement_b200 import api             # noqa: E402
from test_gpu_fuzz import _batch, _signals          # noqa: E402

o = pyoracle.Oracle()
ctx = api.Context(0)
KIND = ["zeros", "dc", "square", "quiet noise", "clipped noise", "clicks", "speech+zero runs", "tone+dc", "scaled speech", "nyquist"]

xs = _signals(np.random.default_rng(8), 8000, 120, 1.5)
samples, offsets = _batch(xs)
r = ctx.etsi_denoise(samples, offsets, ceps=True, den_f32=True, trace=True)
st = defaultdict(lambda: dict(lsb=0, snr=999.0, ceps=0.0, flags=0, frames=0))
for u, x in enumerate(xs):
    ref = o.etsi_denoise(x, with_ceps=True, trace=True)
    lo = int(offsets[u]); nb = len(x) // 80 * 80
    k = st[u % 10]
    d = np.abs(r["den_i16"][lo:lo + nb].astype(np.int32) - ref["den_i16"][:nb].astype(np.int32))
    k["lsb"] = max(k["lsb"], int(d.max(initial=0)))
    if np.any(ref["den_f32"][:nb]):
        k["snr"] = min(k["snr"], snr_db(ref["den_f32"][:nb], r["den_f32"][lo:lo + nb]))
    co = int(r["ceps_offsets"][u])
    if len(ref["ceps"]):
        k["ceps"] = max(k["ceps"], float(ceps_err(ref["ceps"], r["ceps"][co:co + len(ref["ceps"])]).max()))
    tr = r["trace"][int(r["trace_offsets"][u]):int(r["trace_offsets"][u]) + ref["calls"]]
    info = ref["info"]; on = info["stage0_ran"] > 0
    for bit, key in ((1, "flag_vad"), (2, "sp_var"), (4, "sp_spec"), (8, "sp_mel"), (16, "sp_vadns")):
        k["flags"] += int(np.sum(((tr[on] & bit) != 0) != (info[key][on] != 0)))
    k["frames"] += int(on.sum())
print("P8")
for i in range(10):
    print("  %-18s" % KIND[i], st[i])

xs = _signals(np.random.default_rng(16), 16000, 100, 1.2)
samples, offsets = _batch(xs)
r = ctx.p16_wiener(samples, offsets)
bo = r["block_offsets"]
st = defaultdict(lambda: dict(snr=999.0, gains=0.0, flags=0, blocks=0, rows_bad=0))
for u, x in enumerate(xs):
    s = o.p16_stream(); ref = s.func_wiener(x.astype(np.float32)); s.close()
    lo = int(offsets[u]); nb = len(x) // 160; nr = int(r["n_rows"][u]); k = st[u % 10]
    if nr != len(ref["gains"]):
        k["rows_bad"] += 1
        continue
    if np.any(ref["out"][:nb * 160]):
        k["snr"] = min(k["snr"], snr_db(ref["out"][:nb * 160], r["out"][lo:lo + nb * 160]))
    if nr:
        k["gains"] = max(k["gains"], float(rel_err(ref["gains"], r["gains"][bo[u]:bo[u] + nr], 1e-3).max()))
    fl = r["flags"][:, bo[u]:bo[u] + nb]
    for i, key in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")):
        k["flags"] += int(np.sum(fl[i] != ref[key]))
    k["blocks"] += nb
print("P16")
for i in range(10):
    print("  %-18s" % KIND[i], st[i])

xs = _signals(np.random.default_rng(15), 8000, 100, 1.5)
samples, offsets = _batch(xs)
got = ctx.etsi_features(samples, offsets)
st = defaultdict(lambda: dict(rows=0, bad=0, vad=0, max=0.0, shape_bad=0))
for u, x in enumerate(xs):
    ref, _ = o.etsi_features(x); k = st[u % 10]
    if got[u].shape != ref.shape:
        k["shape_bad"] += 1
        continue
    if not len(ref):
        continue
    e = ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1)
    k["rows"] += len(e); k["bad"] += int((e > 1e-4).sum()); k["max"] = max(k["max"], float(e.max()))
    k["vad"] += int(np.sum(got[u][:, 14] != ref[:, 14]))
print("features")
for i in range(10):
    print("  %-18s" % KIND[i], st[i])
