"""Where does the exact arithmetic mode (se_ctx_set_exact) still differ from the oracle?  Run on the GPU box.
Prints, stage by stage, how many values are not bit-identical and the first place they differ."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ceps_err                       # noqa: E402
from oracle import pyoracle                         # noqa: E402
from speech_enhancement_b200 import api, synth      # noqa: E402

o = pyoracle.Oracle()
ctx = api.Context(0)
ctx.set_exact(True)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20


def neq(a, b):
    a = np.asarray(a); b = np.asarray(b)
    return int(np.sum(a.view(np.uint32) != b.view(np.uint32))) if a.dtype == np.float32 else int(np.sum(a != b))


def first(a, b):
    d = np.flatnonzero(np.asarray(a).ravel().view(np.uint32) != np.asarray(b).ravel().view(np.uint32))
    return None if len(d) == 0 else (int(d[0]), float(np.asarray(a).ravel()[d[0]]), float(np.asarray(b).ravel()[d[0]]))


rng = np.random.default_rng(5)
x = (rng.standard_normal((8, 256)) * 1000).astype(np.float32)
y = ctx.rfft(x, 8)
print("rfft 256/8 : %d of %d values differ" % (neq(y, np.stack([o.rfft(v, 8) for v in x])), y.size))
x5 = (rng.standard_normal((8, 512)) * 1000).astype(np.float32)
y5 = ctx.rfft(x5, 8)
print("rfft 512/8 : %d of %d values differ" % (neq(y5, np.stack([o.rfft(v, 8, cxx=True) for v in x5])), y5.size))

fr = (rng.standard_normal((64, 201)) * 800).astype(np.float32)
cc = ctx.comp_ceps(fr)
rc = np.stack([o.comp_ceps(f) for f in fr])
print("comp_ceps  : %d of %d values differ, first %s" % (neq(cc, rc), cc.size, first(cc, rc)))

q = synth.utterance(9, 16000, 160 * 40)
lp, hp = ctx.qmf_split(q)
rlp, rhp = o.qmf_split(q)
print("qmf_split  : lp %d hp %d of %d differ" % (neq(lp.ravel(), rlp.ravel()), neq(hp.ravel(), rhp.ravel()), lp.size))

# ---- P8 spectra / stream / ceps
tot = dict(sp=0, spn=0, den=0, denn=0, ce=0, cen=0, i16=0, flags=0)
for u in range(N):
    xs = synth.utterance(1000 + u, 8000, synth.ragged_length(1000 + u, 8000, 2.0, 6.0))
    sp, produced, _ = ctx.ns_spectra(8, xs, [0, len(xs)])
    rsp, ran = o.ns_spectra(8, xs)
    assert np.array_equal(produced, ran)
    for w in range(8):
        st = 0 if w in (0, 1, 2, 6) else 1
        m = (ran & (2 if st == 0 else 1)) != 0
        d = neq(sp[m, w], rsp[m, w])
        tot["sp"] += d
        tot["spn"] += int(m.sum()) * sp.shape[2]
        if d and u < 3:
            bad = np.flatnonzero((sp[:, w].view(np.uint32) != rsp[:, w].view(np.uint32)).any(axis=1) & m)
            print("   utt %d plane %d: %d values differ, first block %d" % (u, w, d, int(bad[0])))
    r = ctx.etsi_denoise(xs, [0, len(xs)], ceps=True, den_f32=True, trace=True)
    ref = o.etsi_denoise(xs, with_ceps=True, trace=True)
    nb = len(xs) // 80 * 80
    tot["den"] += neq(r["den_f32"][:nb], ref["den_f32"][:nb]); tot["denn"] += nb
    tot["i16"] += neq(r["den_i16"][:nb], ref["den_i16"][:nb])
    nc = len(ref["ceps"])
    tot["ce"] += neq(r["ceps"][:nc], ref["ceps"]); tot["cen"] += nc * 14
    if neq(r["ceps"][:nc], ref["ceps"]) and u < 3:
        print("   utt %d ceps first diff %s, max frame-scale err %.3g" % (u, first(r["ceps"][:nc], ref["ceps"]), ceps_err(ref["ceps"], r["ceps"][:nc]).max()))
    tr = r["trace"][:ref["calls"]]
    info = ref["info"]; on = info["stage0_ran"] > 0
    tot["flags"] += sum(int(np.sum(((tr[on] & bit) != 0) != (info[key][on] != 0)))
                        for bit, key in ((1, "flag_vad"), (2, "sp_var"), (4, "sp_spec"), (8, "sp_mel"), (16, "sp_vadns")))
print("P8 exact   : spectra %d of %d differ; den_f32 %d of %d; den_i16 %d; ceps %d of %d; flag mismatches %d"
      % (tot["sp"], tot["spn"], tot["den"], tot["denn"], tot["i16"], tot["ce"], tot["cen"], tot["flags"]))

# ---- P16
tot = dict(out=0, outn=0, g=0, gn=0, fl=0, sp=0, spn=0)
for u in range(max(2, N // 2)):
    xs = synth.utterance(2000 + u, 16000, synth.ragged_length(2000 + u, 16000, 2.0, 5.0))
    r = ctx.p16_wiener(xs, [0, len(xs)])
    s = o.p16_stream(); ref = s.func_wiener(xs.astype(np.float32)); s.close()
    nb = len(xs) // 160
    tot["out"] += neq(r["out"][:nb * 160], ref["out"][:nb * 160]); tot["outn"] += nb * 160
    nr = int(r["n_rows"][0])
    assert nr == len(ref["gains"])
    tot["g"] += neq(r["gains"][:nr], ref["gains"]); tot["gn"] += nr * 25
    tot["fl"] += sum(int(np.sum(r["flags"][i][:nb] != ref[k])) for i, k in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns", "frame_counter")))
    sp, produced, _ = ctx.ns_spectra(16, xs, [0, len(xs)])
    rsp, ran = o.ns_spectra(16, xs)
    for w in range(8):
        st = 0 if w in (0, 1, 2, 6) else 1
        m = (ran & (2 if st == 0 else 1)) != 0
        d = neq(sp[m, w], rsp[m, w])
        tot["sp"] += d; tot["spn"] += int(m.sum()) * sp.shape[2]
        if d and u < 2:
            bad = np.flatnonzero((sp[:, w].view(np.uint32) != rsp[:, w].view(np.uint32)).any(axis=1) & m)
            print("   P16 utt %d plane %d: %d values differ, first block %d" % (u, w, d, int(bad[0])))
print("P16 exact  : out %d of %d differ; gains %d of %d; flags %d; spectra %d of %d"
      % (tot["out"], tot["outn"], tot["g"], tot["gn"], tot["fl"], tot["sp"], tot["spn"]))

# ---- feature chains
for mode16, name in ((False, "features 8k"), (True, "features 16k-mode")):
    ctx.set_feature_flush(True)
    fs = 16000 if mode16 else 8000
    xs = [synth.utterance((7000 if mode16 else 5000) + i, fs, fs * 5) for i in range(N)]
    offsets = np.concatenate([[0], np.cumsum([len(v) for v in xs])]).astype(np.int64)
    if mode16:
        got, den = ctx.etsi16k_features(np.concatenate(xs), offsets, denoised=True)
    else:
        got, den = ctx.etsi_features(np.concatenate(xs), offsets, denoised=True)
    nd = nv = rows = 0
    worst = 0.0
    for u, v in enumerate(xs):
        ref = o.etsi_features_flush(v, mode16)
        assert ref.shape == got[u].shape, (ref.shape, got[u].shape)
        d = neq(got[u], ref)
        nd += d; nv += ref.size; rows += len(ref)
        worst = max(worst, float(ceps_err(ref[:, :14], got[u][:, :14]).max()))
        if d and u < 3:
            bad = np.flatnonzero((got[u].view(np.uint32) != ref.view(np.uint32)).any(axis=1))
            print("   %s utt %d: %d values differ, first row %d: %s" % (name, u, d, int(bad[0]), first(got[u][bad[0]], ref[bad[0]])))
    print("%-18s: %d rows, %d of %d values differ, max frame-scale err %.3g" % (name, rows, nd, nv, worst))
    ctx.set_feature_flush(False)
