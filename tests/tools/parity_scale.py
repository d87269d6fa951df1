"""Parity statistics at scale (run on the GPU box): P8 and P16 on N ragged synthetic utterances against the -O2 oracle.
Prints worst-case errors and the number of frame-dropping / VAD flag mismatches (threshold ties)."""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ceps_err, ceps_err_survey, rel_err, snr_db      # noqa: E402
from oracle import pyoracle                         # noqa: E402
from speech_enhancement_b200 import api, synth      # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
T = max(1, min(32, os.cpu_count() or 1))
o = pyoracle.Oracle()
ctx = api.Context(0)

samples, offsets = synth.corpus(30000, N, 8000, min_s=2.0, max_s=6.0, threads=T)
r = ctx.etsi_denoise(samples, offsets, ceps=True, den_f32=True, trace=True)


def p8(u):
    x = samples[offsets[u]:offsets[u + 1]]
    ref = o.etsi_denoise(x, with_ceps=True, trace=True)
    lo = int(offsets[u]); nb = len(x) // 80 * 80; nc = len(ref["ceps"]); co = int(r["ceps_offsets"][u])
    lsb = int(np.abs(r["den_i16"][lo:lo + nb].astype(np.int32) - ref["den_i16"][:nb].astype(np.int32)).max(initial=0))
    snr = snr_db(ref["den_f32"][:nb], r["den_f32"][lo:lo + nb])
    ce = float(ceps_err(ref["ceps"], r["ceps"][co:co + nc]).max()) if nc else 0.0
    sv = ceps_err_survey(ref["ceps"], r["ceps"][co:co + nc]) if nc else np.zeros((0, 14))
    tr = r["trace"][int(r["trace_offsets"][u]):int(r["trace_offsets"][u]) + ref["calls"]]
    info = ref["info"]; on = info["stage0_ran"] > 0
    mm = sum(int(np.sum(((tr[on] & bit) != 0) != (info[key][on] != 0)))
             for bit, key in ((1, "flag_vad"), (2, "sp_var"), (4, "sp_spec"), (8, "sp_mel"), (16, "sp_vadns")))
    return lsb, snr, ce, mm, int(on.sum()), nc == int(r["n_ceps"][u]), int((sv > 1e-4).sum()), sv.size, float(sv.max(initial=0))


with ThreadPoolExecutor(T) as ex:
    res = list(ex.map(p8, range(N)))
print("P8  %d utterances, %d frames: max |d| %d LSB, min SNR %.1f dB, max ceps err %.3g, flag mismatches %d of %d, counts ok %s"
      % (N, sum(x[4] for x in res), max(x[0] for x in res), min(x[1] for x in res), max(x[2] for x in res),
         sum(x[3] for x in res), 5 * sum(x[4] for x in res), all(x[5] for x in res)))
print("P8  SURVEY 8(c) element-wise metric (floor 1e-3): %d of %d cepstral values above 1e-4 (share %.5f), max %.3g"
      % (sum(x[6] for x in res), sum(x[7] for x in res), sum(x[6] for x in res) / max(1, sum(x[7] for x in res)),
         max(x[8] for x in res)))

samples, offsets = synth.corpus(60000, N, 16000, min_s=2.0, max_s=6.0, threads=T)
r = ctx.p16_wiener(samples, offsets)
bo = r["block_offsets"]


def p16(u):
    x = samples[offsets[u]:offsets[u + 1]]
    s = o.p16_stream(); ref = s.func_wiener(x.astype(np.float32)); s.close()
    lo = int(offsets[u]); nb = len(x) // 160; nr = int(r["n_rows"][u])
    snr = snr_db(ref["out"][:nb * 160], r["out"][lo:lo + nb * 160])
    ge = float(rel_err(ref["gains"], r["gains"][bo[u]:bo[u] + nr], 1e-3).max()) if nr and nr == len(ref["gains"]) else 0.0
    fl = r["flags"][:, bo[u]:bo[u] + nb]
    mm = sum(int(np.sum(fl[i] != ref[k])) for i, k in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")))
    return snr, ge, mm, nb, nr == len(ref["gains"]) and np.array_equal(fl[4], ref["frame_counter"])


with ThreadPoolExecutor(T) as ex:
    res = list(ex.map(p16, range(N)))
print("P16 %d utterances, %d blocks: min SNR %.1f dB, max gains err %.3g, flag mismatches %d of %d, counts ok %s"
      % (N, sum(x[3] for x in res), min(x[0] for x in res), max(x[1] for x in res), sum(x[2] for x in res),
         4 * sum(x[3] for x in res), all(x[4] for x in res)))

# ---- G64: default-mode resynthesis (both mask kinds), exact sub-bands and IBM bits ---------------------------
M = max(16, N // 10)
xs = [synth.utterance(90000 + i, 16000, synth.ragged_length(90000 + i, 16000, 1.0, 3.0), components=True) for i in range(M)]
lens = [len(x[0]) for x in xs]
offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
noisy = np.concatenate([x[0] for x in xs])
for mode, name in ((api.RESYNTH_IBM, "IBM"), (api.RESYNTH_1DNN, "1dnn")):
    nfs = [api.resynth_num_frames(n, mode) for n in lens]
    fo = np.concatenate([[0], np.cumsum(nfs)]).astype(np.int64)
    rng = np.random.default_rng(mode)
    mask = rng.random((int(fo[-1]), 64), dtype=np.float32)
    if mode != api.RESYNTH_IBM:
        mask[mask < 0.2] = 0.0
    fast = ctx.resynth(noisy, offs, mask, fo, mode, exact=False)
    exact = ctx.resynth(noisy, offs, mask, fo, mode, exact=True)

    def rs(u):
        ref, _ = o.resynth(xs[u][0], mask[fo[u]:fo[u + 1]].copy(), mode)
        lo, hi = int(offs[u]), int(offs[u + 1])
        d = int(np.abs(fast[lo:hi].astype(np.int32) - ref.astype(np.int32)).max(initial=0))
        return d, snr_db(ref, fast[lo:hi]), bool(np.array_equal(exact[lo:hi], ref))

    with ThreadPoolExecutor(T) as ex:
        res = list(ex.map(rs, range(M)))
    print("G64 resynth %-4s %d utterances, %d samples: default mode max |d| %d LSB, min SNR %.1f dB; exact mode bit-identical: %s"
          % (name, M, int(offs[-1]), max(x[0] for x in res), min(x[1] for x in res), all(x[2] for x in res)))

pure = np.concatenate([x[1] for x in xs]); noise = np.concatenate([x[2] for x in xs])
sub = ctx.subband(noisy, offs, exact=True)
irm, bits, fo = ctx.ibm(pure, noise, offs)


def g64(u):
    lo, hi = int(offs[u]), int(offs[u + 1])
    ok_sub = bool(np.array_equal(sub[64 * lo:64 * hi].reshape(64, hi - lo), o.subband(xs[u][0])))
    rirm, ribm = o.single_ibm(o.subband(xs[u][1]), o.subband(xs[u][2]))
    gb = ((bits[fo[u]:fo[u + 1], None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.uint8)
    return ok_sub, int(np.sum(gb != ribm)), ribm.size


with ThreadPoolExecutor(T) as ex:
    res = list(ex.map(g64, range(M)))
print("G64 subband rows bit-identical: %s; IBM bit mismatches %d of %d" % (all(x[0] for x in res), sum(x[1] for x in res), sum(x[2] for x in res)))
