"""How often does the GPU feature chain differ from the oracle by more than 1e-4 of the frame scale?
(WaveProc peak-position flips on 1-ulp differences of the denoised stream.)  Run on the GPU box."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ceps_err, ceps_err_survey     # noqa: E402
from oracle import pyoracle                        # noqa: E402
from speech_enhancement_b200 import api, synth     # noqa: E402

n_utts = int(sys.argv[1]) if len(sys.argv) > 1 else 200
o = pyoracle.Oracle(fast=False)
ctx = api.Context(0)
if "--exact" in sys.argv:                           # se_ctx_set_exact: the reference's own arithmetic
    ctx.set_exact(True)
    print("exact arithmetic mode")
xs = [synth.utterance(5000 + i, 8000, 8000 * 5) for i in range(n_utts)]
offsets = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
got = ctx.etsi_features(np.concatenate(xs), offsets)
errs, vad_diff, rows, sv = [], 0, 0, []
for u, x in enumerate(xs):
    ref, _ = o.etsi_features(x)
    assert ref.shape == got[u].shape
    errs.append(ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1))
    sv.append(ceps_err_survey(ref[:, :14], got[u][:, :14]).ravel())
    vad_diff += int((ref[:, 14] != got[u][:, 14]).sum())
    rows += len(ref)
e = np.concatenate(errs)
print("rows", rows, "vad flag mismatches", vad_diff, "values not bit-identical",
      sum(int(np.sum(o.etsi_features(x)[0].view(np.uint32) != got[u].view(np.uint32))) for u, x in enumerate(xs)))
for t in (1e-5, 1e-4, 1e-3, 1e-2, 5e-2):
    print("share of frames with err > %g: %.5f" % (t, float((e > t).mean())))
print("max err %.4g, median %.3g" % (e.max(), np.median(e)))
sv = np.concatenate(sv)
print("SURVEY 8(c) element-wise metric: %d of %d values above 1e-4 (share %.5f), max %.3g" % ((sv > 1e-4).sum(), sv.size, (sv > 1e-4).mean(), sv.max()))

# ---- the same statistics for the ETSI 16 kHz mode, with the flush rows on ---------------------------------------
ctx.set_feature_flush(True)
xs = [synth.utterance(7000 + i, 16000, 16000 * 5) for i in range(n_utts)]
offsets = np.concatenate([[0], np.cumsum([len(x) for x in xs])]).astype(np.int64)
got, den = ctx.etsi16k_features(np.concatenate(xs), offsets, denoised=True)
errs, vad_diff, rows, lsb, sv = [], 0, 0, 0, []
for u, x in enumerate(xs):
    ref = o.etsi_features_flush(x, True)
    _, rden = o.etsi_features_16k(x)
    assert ref.shape == got[u].shape, (u, ref.shape, got[u].shape)
    errs.append(ceps_err(ref[:, :14], got[u][:, :14]).max(axis=1))
    sv.append(ceps_err_survey(ref[:, :14], got[u][:, :14]).ravel())
    vad_diff += int((ref[:, 14] != got[u][:, 14]).sum())
    lsb = max(lsb, int(np.abs(den[u].astype(np.int32) - rden.astype(np.int32)).max()))
    rows += len(ref)
e = np.concatenate(errs)
print("16 kHz mode (+flush): rows", rows, "vad flag mismatches", vad_diff, "low-band int16 max |d|", lsb,
      "values not bit-identical",
      sum(int(np.sum(o.etsi_features_flush(x, True).view(np.uint32) != got[u].view(np.uint32))) for u, x in enumerate(xs)))
for t in (1e-5, 1e-4, 1e-3, 1e-2, 5e-2):
    print("share of frames with err > %g: %.5f" % (t, float((e > t).mean())))
print("max err %.4g, median %.3g" % (e.max(), np.median(e)))
sv = np.concatenate(sv)
print("SURVEY 8(c) element-wise metric: %d of %d values above 1e-4 (share %.5f), max %.3g" % ((sv > 1e-4).sum(), sv.size, (sv > 1e-4).mean(), sv.max()))
