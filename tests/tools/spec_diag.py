import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, 'tests')
import numpy as np
from oracle import pyoracle
from speech_enhancement_b200 import api, synth
o = pyoracle.Oracle(); ctx = api.Context(0)
FL = float(np.exp(-10.0))
for prof, fs in ((8, 8000), (16, 16000)):
    tot = np.zeros(8); bad = np.zeros(8); worst_el = np.zeros(8); worst_fr = np.zeros(8)
    for u in range(6):
        x = synth.utterance(100 + u, fs, fs * 5)
        sp, pr, _ = ctx.ns_spectra(prof, x, [0, len(x)])
        rsp, ran = o.ns_spectra(prof, x)
        for w in range(8):
            st = 0 if w in (0, 1, 2, 6) else 1
            m = (ran & (2 if st == 0 else 1)) != 0
            a = rsp[m, w].astype(np.float64); b = sp[m, w].astype(np.float64)
            el = np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), FL)
            fr = np.abs(a - b) / np.maximum(np.abs(a).max(axis=1, keepdims=True), FL)
            tot[w] += el.size; bad[w] += (el > 1e-4).sum(); worst_el[w] = max(worst_el[w], el.max()); worst_fr[w] = max(worst_fr[w], fr.max())
    print("P%d" % prof)
    for w in range(8):
        print("  plane %d: %d of %d above 1e-4 elementwise (share %.2e), max elementwise %.3g, max on the frame's scale %.3g" % (w, bad[w], tot[w], bad[w] / tot[w], worst_el[w], worst_fr[w]))
