import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from conftest import ceps_err, snr_db
from oracle import pyoracle
from speech_enhancement_b200 import api, synth
o = pyoracle.Oracle()
FR = 33500  # see tests/test_gpu_long.py
n = FR*80+33
x = np.concatenate([synth.utterance(100+u, 8000, 80000) for u in range(n//80000+1)])[:n]
with api.Context(0) as ctx:
    r = ctx.etsi_denoise(x, [0, len(x)], ceps=True, den_f32=True, trace=True)
ref = o.etsi_denoise(x, with_ceps=True, trace=True)
nc = len(ref["ceps"])
e = ceps_err(ref["ceps"], r["ceps"][:nc])
em = e.max(axis=1)
bad = np.nonzero(em > 1e-4)[0]
print("bad frames", len(bad), bad[:40], bad[-10:])
for lo in range(0, nc, 2000):
    print(lo, "max err %.2e" % em[lo:lo+2000].max(), "p99 %.2e" % np.percentile(em[lo:lo+2000], 99))
# float stream error per segment
df = np.abs(ref["den_f32"][:FR*80].astype(np.float64) - r["den_f32"][:FR*80])
for lo in range(0, FR, 2000):
    seg = slice(lo*80, (lo+2000)*80)
    print(lo, "den max abs err %.3e" % df[seg].max(), "snr %.1f" % snr_db(ref["den_f32"][seg], r["den_f32"][seg]))
if len(bad):
    f = bad[0]
    print("frame", f, "ref", ref["ceps"][f], "got", r["ceps"][f])
