"""Generates tests/golden/*.npz from oracle/_ref = the reference's OWN sources compiled in place from
/root/reference (oracle/Makefile `make ref`).  Run in the dev container (needs /root/reference):

    python tests/golden/make_golden.py

The vectors pin the plain-C oracle (tests/test_oracle_golden.py) and, on the GPU box where
/root/reference does not exist, the CUDA path (tests/test_gpu_golden.py).
Inputs come from the deterministic generator speech_enhancement_b200/csrc/synth.c.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import pyoracle as po  # noqa: E402
from speech_enhancement_b200 import synth  # noqa: E402


def features():
    """Full feature chain (WaveProc -> CompCeps -> PostProc -> VAD, ParmInterface.c:274-311) + DoWaveProc alone."""
    assert po.build_ref(), "needs /root/reference"
    r = po.RefEtsi()
    x = synth.utterance(3, 8000, 8000 + 33)
    xz = synth.utterance(7, 8000, 4000)                      # 5 leading all-zero frames -> 5 null vectors
    f, d = r.etsi_features(x)
    fz, dz = r.etsi_features(xz)
    rng = np.random.default_rng(386)
    n = np.arange(200)
    frames = np.stack([(rng.standard_normal(200) * 2000).astype(np.float32),
                       (3000 * np.sin(2 * np.pi * n / 61.0) * (1 + 0.5 * np.sin(2 * np.pi * n / 57))).astype(np.float32),
                       np.round(rng.standard_normal(200) * 0.5).astype(np.float32),     # below the energy gate
                       rng.integers(-32768, 32768, 200).astype(np.float32)])           # int32 wrap-around case
    xs = synth.utterance(11, 8000, 80 * 9)                   # too short to start the VAD: the flush emits stale rows
    np.savez_compressed(os.path.join(HERE, "p8_features.npz"), x=x, xz=xz, feat=f, z_feat=fz, den=d, z_den=dz,
                        frames=frames, wp=np.stack([r.wave_proc(v) for v in frames]),
                        # + FlushAdvProcess: the last 6 rows of each
                        flush=r.etsi_features_flush(x)[-6:], z_flush=r.etsi_features_flush(xz)[-6:],
                        xs=xs, s_flush=r.etsi_features_flush(xs))


def etsi16k():
    """ETSI 16 kHz mode (16kHzProcessing.c + hooks): feature rows, low-band stream and the bare QMF split."""
    assert po.build_ref(), "needs /root/reference"
    r = po.RefEtsi()
    x = synth.utterance(5, 16000, 16000 + 71)
    xz = synth.utterance(7, 16000, 8000)                     # 5 leading all-zero calls -> 5 null vectors
    f, d = r.etsi_features_16k(x)
    fz, dz = r.etsi_features_16k(xz)
    q = synth.utterance(9, 16000, 160 * 12)
    lp, hp = r.qmf_split(q)
    np.savez_compressed(os.path.join(HERE, "etsi16k_features.npz"), x=x, xz=xz, feat=f, z_feat=fz, den=d, z_den=dz,
                        q=q, lp=lp, hp=hp, flush=r.etsi_features_flush(x, True)[-6:])


def spectra():
    """Per-call SPECTRUM state of the reference itself (NoiseSupStructX.nsVar.spectrum after each call): the P8 utterance
    of p8_etsi.npz through ref_etsi_trace, the P16 signal `a` of p16_wiener.npz through func_Wiener one 160-sample block
    at a time on one thread instance.  Calls 0..15 and every 4th after that are kept (planes nSigSE1, noiseSE1,
    denSigSE1, nSigSE2, noiseSE2, denSigSE2)."""
    assert po.build_ref(), "needs /root/reference"
    keys = ("nsig1", "noise1", "den1", "nsig2", "noise2", "den2")
    r = po.RefEtsi()
    x = synth.utterance(3, 8000, 8000 + 33)
    t = r.trace(x, with_ceps=False)
    n8 = len(x) // 80
    keep8 = np.array([c for c in range(n8) if c < 16 or c % 4 == 3 or c == n8 - 1])
    p8 = np.stack([t[k][keep8] for k in keys], axis=1)
    p = po.RefP16()
    a = synth.utterance(5, 16000, 16000).astype(np.float32)
    a[160 * 20:160 * 22] = 0.0
    n16 = len(a) // 160
    rows = []
    for b in range(n16):
        p.func_wiener(a[160 * b:160 * (b + 1)])
        rows.append(np.stack(p.state()[0]))
    p.close()
    keep16 = np.array([c for c in range(n16) if c < 16 or c % 4 == 3 or c == n16 - 1])
    np.savez_compressed(os.path.join(HERE, "spectra.npz"), x8=x, calls8=keep8, sp8=p8, ret8=t["ret"], a16=a,
                        calls16=keep16, sp16=np.stack(rows)[keep16])


def adv_shipped():
    """DoAdvProcess / FlushAdvProcess exactly as the reference ships them (ParmInterface.c:208-355), frame by frame:
    leading all-zero frames (null vector, TRUE), latency calls, and the flush of a never-fed VAD delay line."""
    assert po.build_ref(), "needs /root/reference"
    r = po.RefEtsi()
    d = {}
    for tag, seed, n in (("z", 7, 4000), ("s", 3, 80 * 9), ("m", 3, 80 * 20)):
        x = synth.utterance(seed, 8000, n)
        t = r.adv_process_as_shipped(x)
        d[tag + "_x"] = x
        for k, v in t.items():
            d[tag + "_" + k] = v
    np.savez_compressed(os.path.join(HERE, "adv_shipped.npz"), **d)


def main():
    assert po.build_ref(), "needs /root/reference"
    if "--only-adv" in sys.argv:
        return adv_shipped()
    if "--only-spectra" in sys.argv:
        return spectra()
    if "--only-features" in sys.argv:
        return features()
    if "--only-16k" in sys.argv:
        return etsi16k()
    features()
    etsi16k()
    spectra()
    adv_shipped()
    rng = np.random.default_rng(20141106)

    # ---- P8: etsi_denoise + re-enabled CompCeps step, with the per-call VAD trace ------------------
    r = po.RefEtsi()
    x = synth.utterance(3, 8000, 8000 + 33)                 # 1 s + ragged tail
    t = r.trace(x, with_ceps=True)
    out, rc = r.etsi_denoise(x)
    assert rc == 0 and np.array_equal(out[:8000], t["den_i16"][:8000])
    xz = synth.utterance(7, 8000, 4000)                      # utterance with 50 ms of leading zeros
    tz = r.trace(xz, with_ceps=True)
    keys = ("den_i16", "ceps", "ret", "flag_vad", "hang_over", "nb_speech", "nb_frame0", "nb_frame1", "sp_var",
            "sp_spec", "sp_mel", "sp_vadns", "mean_en", "alfa_gf", "out_f32")
    np.savez_compressed(os.path.join(HERE, "p8_etsi.npz"), x=x, xz=xz,
                        **{k: t[k] for k in keys}, **{"z_" + k: tz[k] for k in keys},
                        noise1_last=t["noise1"][-1], den2_last=t["den2"][-1])
    tb = r.tables()
    np.savez_compressed(os.path.join(HERE, "p8_tables.npz"), **tb)
    fx = (rng.standard_normal((4, 256)) * 1000).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "rfft.npz"), x256=fx, y256=np.stack([r.rfft(v, 8) for v in fx]))

    # ---- P16: plugin func_Wiener, two calls on one thread instance (state carry-over) --------------
    p = po.RefP16()
    a = synth.utterance(5, 16000, 16000).astype(np.float32)
    a[160 * 20:160 * 22] = 0.0                               # zero blocks in the middle
    b = synth.utterance(6, 16000, 8000).astype(np.float32)
    ra, rb = p.func_wiener(a, want_text=True), p.func_wiener(b)
    fx5 = (rng.standard_normal((4, 512)) * 1000).astype(np.float32)
    y5 = np.stack([p.rfft(v, 8) for v in fx5])               # the truncated (512, 8) schedule, SURVEY Q1
    np.savez_compressed(os.path.join(HERE, "p16_wiener.npz"), a=a, b=b, text_a=np.array(ra["text"]),
                        x512=fx5, y512=y5,
                        **{"a_" + k: v for k, v in ra.items() if k != "text"},
                        **{"b_" + k: v for k, v in rb.items()})
    np.savez_compressed(os.path.join(HERE, "p16_tables.npz"), **p.tables())
    p.close()

    # ---- G64: channel parameters, gammatone, hair cell, subband, addnoise, resynth x3 --------------
    g = po.RefG64("subband")
    cf, bw, mid = g.channel_params()
    w = synth.utterance(9, 16000, 1600)
    imp = np.zeros(600, np.float32)
    imp[0] = 1000.0
    chans = [0, 17, 40, 63]
    gt = np.stack([g.gammatone(imp, cf[c], bw[c], mid[c]) for c in chans])
    hc = g.haircell(gt[1] * 50)
    sub = g.subband(w)
    _, clean, noise = synth.utterance(10, 16000, 1600, components=True)
    an, ay = g.addnoise(clean, noise, 5)
    res = {}
    for name, mode in (("resynth_ibm", 0), ("resynth_ori", 1), ("resynth_1dnn", 2)):
        gg = po.RefG64(name)
        nf = 1600 // 160 if mode == 2 else (1600 - 320) // 160 + 1
        m = rng.random((nf, 64)).astype(np.float32)
        if mode:
            m[rng.random((nf, 64)) < 0.25] = 0.0
        o, m_after = gg.resynth(w, m)
        res["mask%d" % mode], res["out%d" % mode], res["mask_after%d" % mode] = m, o, m_after
    np.savez_compressed(os.path.join(HERE, "g64.npz"), cf=cf, bw=bw, mid=mid, w=w, imp=imp, chans=np.array(chans),
                        gt=gt, hc=hc, sub=sub, clean=clean, noise=noise, an=an, ay=ay, **res)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
