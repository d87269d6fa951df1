"""The reference against ITSELF: oracle/_ref built -O2 -ffp-contract=off (libref_etsi.so, the parity build) vs the same
sources built -O3 -march=x86-64-v3 (libref_etsi_fast.so: AVX2, FMA contraction) on the corpora
tests/tools/feature_outliers.py runs on the GPU (8 kHz: 200 x 5 s = 97 600 rows; 16 kHz mode + flush: 200 x 5 s =
98 800 rows) and on the P8 NoiseSup+CompCeps corpus.  CPU only; needs oracle/_ref (dev container).

What it answers: how far do two builds of the reference's own code drift apart under
  (a) SURVEY 8(c)'s element-wise cepstral metric  |a-b| <= 1e-4 max(|a|,|b|,1e-3), and
  (b) the frame-scale metric of tests/conftest.py:ceps_err,
and how often does a 1-ulp difference in the denoised stream move one of WaveProc's integer peak positions (a "flip",
frame error > 1e-3) -- the envelope the GPU feature-chain bars in tests/test_gpu_features.py / test_gpu_etsi16k.py are
set to.  Writes profiles/r02_ref_vs_ref_envelope.json.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ceps_err, ceps_err_survey     # noqa: E402
from oracle import pyoracle as po                  # noqa: E402
from speech_enhancement_b200 import synth          # noqa: E402


def stats(e_frame, e_survey, vad_diff, rows, lsb=None):
    d = {"rows": int(rows), "vad_flag_mismatches": int(vad_diff),
         "frame_scale": {"share_gt_1e-4": float((e_frame > 1e-4).mean()), "share_gt_1e-3": float((e_frame > 1e-3).mean()),
                         "share_gt_1e-2": float((e_frame > 1e-2).mean()), "max": float(e_frame.max()),
                         "median": float(np.median(e_frame))},
         "survey_metric": {"elements": int(e_survey.size), "elements_gt_1e-4": int((e_survey > 1e-4).sum()),
                           "share_gt_1e-4": float((e_survey > 1e-4).mean()), "max": float(e_survey.max())}}
    if lsb is not None:
        d["int16_max_abs_diff"] = int(lsb)
    return d


def main():
    n_utts = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    assert po.build_ref(), "needs /root/reference"
    a, b = po.RefEtsi(fast=False), po.RefEtsi(fast=True)
    out = {"builds": {"a": "gcc -O2 -ffp-contract=off (libref_etsi.so)", "b": "gcc -O3 -march=x86-64-v3 (libref_etsi_fast.so)"},
           "n_utts": n_utts}

    # ---- 8 kHz feature chain (NoiseSup -> WaveProc -> CompCeps -> PostProc -> VAD) ----------------------------------
    ef, es, vd, rows, lsb = [], [], 0, 0, 0
    for i in range(n_utts):
        x = synth.utterance(5000 + i, 8000, 8000 * 5)
        fa, da = a.etsi_features(x)
        fb, db = b.etsi_features(x)
        assert fa.shape == fb.shape
        ef.append(ceps_err(fa[:, :14], fb[:, :14]).max(axis=1))
        es.append(ceps_err_survey(fa[:, :14], fb[:, :14]).ravel())
        vd += int((fa[:, 14] != fb[:, 14]).sum())
        lsb = max(lsb, int(np.abs(da.astype(np.int32) - db.astype(np.int32)).max()))
        rows += len(fa)
    out["features_8k"] = stats(np.concatenate(ef), np.concatenate(es), vd, rows, lsb)
    print("8 kHz feature chain:", json.dumps(out["features_8k"]))

    # ---- ETSI 16 kHz mode, flush rows on ----------------------------------------------------------------------------
    ef, es, vd, rows, lsb = [], [], 0, 0, 0
    for i in range(n_utts):
        x = synth.utterance(7000 + i, 16000, 16000 * 5)
        fa, fb = a.etsi_features_flush(x, True), b.etsi_features_flush(x, True)
        assert fa.shape == fb.shape
        ef.append(ceps_err(fa[:, :14], fb[:, :14]).max(axis=1))
        es.append(ceps_err_survey(fa[:, :14], fb[:, :14]).ravel())
        vd += int((fa[:, 14] != fb[:, 14]).sum())
        _, da = a.etsi_features_16k(x)
        _, db = b.etsi_features_16k(x)
        lsb = max(lsb, int(np.abs(da.astype(np.int32) - db.astype(np.int32)).max()))
        rows += len(fa)
    out["features_16k_mode"] = stats(np.concatenate(ef), np.concatenate(es), vd, rows, lsb)
    print("16 kHz mode:", json.dumps(out["features_16k_mode"]))

    # ---- P8 NoiseSup + raw CompCeps (no WaveProc): the metric's own path --------------------------------------------
    ef, es, rows, lsb = [], [], 0, 0
    for i in range(n_utts):
        x = synth.utterance(1000 + i, 8000, synth.ragged_length(1000 + i, 8000, 2.0, 10.0))
        ta, tb = a.trace(x, with_ceps=True), b.trace(x, with_ceps=True)
        assert ta["ceps"].shape == tb["ceps"].shape
        if len(ta["ceps"]):
            ef.append(ceps_err(ta["ceps"], tb["ceps"]).max(axis=1))
            es.append(ceps_err_survey(ta["ceps"], tb["ceps"]).ravel())
        lsb = max(lsb, int(np.abs(ta["den_i16"].astype(np.int32) - tb["den_i16"].astype(np.int32)).max()))
        rows += len(ta["ceps"])
    out["p8_noisesup_compceps"] = stats(np.concatenate(ef), np.concatenate(es), 0, rows, lsb)
    print("P8 NoiseSup+CompCeps:", json.dumps(out["p8_noisesup_compceps"]))

    with open(os.path.join(ROOT, "profiles", "r02_ref_vs_ref_envelope.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
