"""BASELINE.json's configurations at FULL size, checked through size-independent properties:
every utterance of the big batch is bit-identical to the same utterance processed alone or in a small batch
(no cross-utterance state, any queue order / chunking / shard), a sample of utterances meets the oracle
tolerances, and structural invariants hold for all of them (frame counts, latency zeros, tails)."""
import os

import numpy as np
import pytest

from conftest import rel_err, snr_db
from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu
THREADS = max(1, min(32, os.cpu_count() or 1))


def test_c2_p16_10k_utterances(ctx, oracle):
    """configs[1]: 10 000 ragged 16 kHz utterances through the plugin NoiseSup (the bench's headline workload)."""
    n_utts = 10000
    samples, offsets = synth.corpus(0, n_utts, 16000, min_s=2.0, max_s=10.0, threads=THREADS)
    r = ctx.p16_wiener(samples, offsets)
    bo = r["block_offsets"]
    lens = np.diff(offsets)
    assert np.array_equal(np.diff(bo), lens // 160)
    # rows = stage-2 frames = processed blocks - 4 of latency (zero blocks are skipped)
    assert np.all(r["n_rows"] <= np.maximum(lens // 160 - 4, 0)) and r["n_rows"].sum() > 0.99 * (lens // 160 - 4).sum()
    rng = np.random.default_rng(1)
    picks = list(rng.choice(n_utts, 12, replace=False)) + [int(np.argmax(lens)), int(np.argmin(lens)), 0, n_utts - 1]
    for u in picks:
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        alone = ctx.p16_wiener(samples[lo:hi], [0, hi - lo])
        assert np.array_equal(alone["out"], r["out"][lo:hi]), u
        nb = int(bo[u + 1] - bo[u])
        assert np.array_equal(alone["flags"][:, :nb], r["flags"][:, bo[u]:bo[u + 1]]), u
        assert alone["n_rows"][0] == r["n_rows"][u]
        assert np.array_equal(alone["gains"][:nb], r["gains"][bo[u]:bo[u + 1]])
        assert not r["out"][hi - (hi - lo) % 160:hi].any()              # tail is defined as 0
    for u in picks[:4]:                                                 # oracle tolerances on a sample
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        s = oracle.p16_stream()
        ref = s.func_wiener(samples[lo:hi].astype(np.float32))
        s.close()
        n = (hi - lo) // 160 * 160
        assert snr_db(ref["out"][:n], r["out"][lo:lo + n]) >= 60.0
        nr = int(r["n_rows"][u])
        assert nr == len(ref["gains"])
        assert rel_err(ref["gains"], r["gains"][bo[u]:bo[u] + nr], 1e-3).max() <= 1e-4
        fl = r["flags"][:, bo[u]:bo[u + 1]]
        same = sum(int(np.sum(fl[i] == ref[k])) for i, k in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")))
        assert same >= 0.999 * 4 * fl.shape[1]                          # flags equal except threshold ties
        assert np.array_equal(fl[4], ref["frame_counter"])


@pytest.mark.parametrize("mode", [api.RESYNTH_IBM, api.RESYNTH_1DNN])
def test_c3_c4_resynth_1k_utterances(ctx, oracle, mode):
    """configs[2] / configs[3]: 1 000 x 10 s masked resynthesis; a utterance inside the big batch equals the same
    utterance alone, and a sample meets the oracle bar (+-1 LSB, >= 60 dB)."""
    n_utts, n = 1000, 160000
    samples, offsets = synth.corpus(20000, n_utts, 16000, seconds=10.0, threads=THREADS)
    nf = api.resynth_num_frames(n, mode)
    fo = np.arange(n_utts + 1, dtype=np.int64) * nf
    rng = np.random.default_rng(mode)
    mask = rng.random((n_utts * nf, 64), dtype=np.float32)
    if mode == api.RESYNTH_1DNN:
        mask[mask < 0.2] = 0.0
    out = ctx.resynth(samples, offsets, mask, fo, mode, exact=False)
    assert out.shape == samples.shape
    for u in (0, 499, 999):
        lo, hi = u * n, (u + 1) * n
        alone = ctx.resynth(samples[lo:hi], [0, n], mask[fo[u]:fo[u + 1]], [0, nf], mode, exact=False)
        assert np.array_equal(alone, out[lo:hi])
    u = 123
    ref_i16, _ = oracle.resynth(samples[u * n:(u + 1) * n], mask[fo[u]:fo[u + 1]].copy(), mode)
    d = np.abs(out[u * n:(u + 1) * n].astype(np.int32) - ref_i16.astype(np.int32))
    assert d.max() <= 1 and snr_db(ref_i16, out[u * n:(u + 1) * n]) >= 60.0


def test_c5_p8_3k_utterances(ctx, oracle):
    """configs[4] shape for the P8 half: 3 000 x 10 s 8 kHz through NoiseSup+CompCeps (one launch)."""
    from test_gpu_p8 import _compare_utt, TIE_FRACTION
    n_utts, n = 3000, 80000
    samples, offsets = synth.corpus(40000, n_utts, 8000, seconds=10.0, threads=THREADS)
    r = ctx.etsi_denoise(samples, offsets, ceps=True, den_f32=True, trace=True)
    # 5 % of the synthetic utterances start with 50 ms of exact zeros: 5 skipped frames (ParmInterface.c:244-250)
    assert set(np.unique(r["n_ceps"]).tolist()) <= {n // 80 - 6, n // 80 - 6 - 5} and np.any(r["n_ceps"] == n // 80 - 11)
    den = r["den_i16"].reshape(n_utts, n)
    assert not den[:, :4 * 80].any()                                   # NoiseSup latency: first 4 shifts are 0
    for u in (0, 1499, 2999):
        alone = ctx.etsi_denoise(samples[u * n:(u + 1) * n], [0, n], ceps=True, den_f32=True, trace=True)
        assert np.array_equal(alone["den_i16"], den[u])
        assert np.array_equal(alone["ceps"], r["ceps"][r["ceps_offsets"][u]:r["ceps_offsets"][u + 1]])
    u = 77
    mism, ran = _compare_utt(oracle, samples[u * n:(u + 1) * n], r, u * n, (u + 1) * n, int(r["ceps_offsets"][u]),
                             int(r["n_ceps"][u]), int(r["trace_offsets"][u]))
    assert mism <= max(1, int(TIE_FRACTION * ran * 5))
