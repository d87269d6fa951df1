"""the training-data flow chained ON THE DEVICE through the C ABI with SE_MEM_DEVICE pointers
(enhancement_extract_subband_linux/cpp/main.cpp:91-274 -> resyth_64sub_IBM): addnoise at an SNR -> scaled noise +
noisy mixture -> fused subbband x2 + make_single_IBM -> IRM / packed IBM bits -> IBM-masked resynthesis of the
mixture.  No host copy between the steps; the result must equal the same chain run through the oracle."""
import numpy as np
import pytest

from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu


def test_mix_mask_resynth_chain_device_resident(ctx, oracle):
    torch = pytest.importorskip("torch")
    lens = [6400, 4801, 3217]
    comps = [synth.utterance(400 + i, 16000, n, components=True) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    pure = np.concatenate([c[1] for c in comps])
    noise = np.concatenate([c[2] for c in comps])
    nu = len(lens)
    fo = np.concatenate([[0], np.cumsum([api.ibm_num_frames(n) for n in lens])]).astype(np.int64)
    assert [api.resynth_num_frames(n, api.RESYNTH_IBM) for n in lens] == list(np.diff(fo))
    d_pure = torch.from_numpy(pure).cuda()
    d_noise = torch.from_numpy(noise).cuda()
    d_noisy = torch.empty_like(d_pure)
    d_irm = torch.empty(int(fo[-1]) * 64, dtype=torch.float32, device="cuda")
    d_bits = torch.empty(int(fo[-1]), dtype=torch.int64, device="cuda")
    d_out = torch.empty_like(d_pure)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    try:
        api._check(ctx.L.se_addnoise_batch(ctx.h, d_pure.data_ptr(), d_noise.data_ptr(), d_noisy.data_ptr(),
                                           offsets.ctypes.data, nu, 0, api.SE_MEM_DEVICE), "se_addnoise_batch")
        ctx.ibm_raw(d_pure.data_ptr(), d_noise.data_ptr(), offsets, nu, d_irm.data_ptr(), d_bits.data_ptr(), fo,
                    mem=api.SE_MEM_DEVICE)
        # the IRM doubles as the soft mask of the IBM-mode resynthesis (> 0.5 -> 1, resyth_64sub_IBM/cpp/extractwav.cpp:95-108)
        ctx.resynth_raw(d_noisy.data_ptr(), offsets, nu, d_irm.data_ptr(), fo, api.RESYNTH_IBM, api.SE_FLAG_EXACT,
                        d_out.data_ptr(), mem=api.SE_MEM_DEVICE)
        ctx.sync()
    finally:
        ctx.set_stream(None)
    noisy, irm, out = d_noisy.cpu().numpy(), d_irm.cpu().numpy().reshape(-1, 64), d_out.cpu().numpy()
    bits = d_bits.cpu().numpy().view(np.uint64)
    for u, (_, p, z) in enumerate(comps):
        lo, hi = int(offsets[u]), int(offsets[u + 1])
        zn, mix = oracle.addnoise(p, z, 0)
        assert np.array_equal(mix, noisy[lo:hi])
        rirm, ribm = oracle.single_ibm(oracle.subband(p), oracle.subband(zn))
        gb = ((bits[fo[u]:fo[u + 1], None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.uint8)
        assert np.array_equal(gb, ribm)
        ok = np.isfinite(rirm)
        gi = irm[fo[u]:fo[u + 1]]
        assert np.allclose(gi[ok], rirm[ok], rtol=1e-4, atol=1e-6)
        m = np.where(ok, gi, 0.0).astype(np.float32)
        if not ok.all():
            continue          # 0/0 frames (digital silence in both signals) make the mask undefined in the reference too
        ref_out, _ = oracle.resynth(mix, m.copy(), api.RESYNTH_IBM)
        assert np.array_equal(out[lo:hi], ref_out), u


def test_extract_wiener_chain(ctx, oracle):
    """extract_Wiener (aurora_etsi_test.cpp:16-141) as one device-resident chain: subbband() -> 64 channel signals, each
    cut to a multiple of 160 samples, all 64 through func_Wiener on ONE thread instance (state carries over from channel
    to channel).  Against the oracle driven exactly that way: FrameCounter and row counts exact, flags equal except
    ties, 25-band gains rel 1e-4.  Includes a scratch limit that forces the batch into several groups."""
    from conftest import rel_err
    lens = [160 * 12 + 77, 160 * 7, 100, 160 * 20 + 159, 0, 160 * 9 + 1]
    xs = [synth.utterance(500 + i, 16000, n) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    samples = np.concatenate(xs)
    got = ctx.extract_wiener(samples, offsets)
    ctx.set_scratch_limit(1 << 20)                          # ~8 000 samples per group: one or two utterances each
    try:
        grouped = ctx.extract_wiener(samples, offsets)
    finally:
        ctx.set_scratch_limit(48 << 30)
    mism = blocks = 0
    for u, x in enumerate(xs):
        n = len(x) - len(x) % 160
        nb = n // 160
        g, fl, rows = got[u]
        assert g.shape == (64 * nb, 25) and fl.shape == (5, 64 * nb)
        assert np.array_equal(g, grouped[u][0]) and np.array_equal(fl, grouped[u][1]) and rows == grouped[u][2]
        if nb == 0:
            assert rows == 0
            continue
        sub = oracle.subband(x[:n]).reshape(64, n)
        s = oracle.p16_stream()
        ref_rows = 0
        r0 = 0
        for chan in range(64):
            ref = s.func_wiener(sub[chan].astype(np.float32))
            k = len(ref["gains"])
            if k:
                e = rel_err(ref["gains"], g[r0:r0 + k], 1e-3)
                assert e.max() <= 1e-4, (u, chan, e.max())
            r0 += k
            ref_rows += k
            f = fl[:, chan * nb:(chan + 1) * nb]
            assert np.array_equal(f[4], ref["frame_counter"]), (u, chan)
            for i, key in enumerate(("sp_var", "sp_spec", "sp_mel", "sp_vadns")):
                mism += int(np.sum(f[i] != ref[key]))
            blocks += nb
        s.close()
        assert rows == ref_rows
        assert not g[ref_rows:].any()
    assert mism <= max(1, int(1e-3 * blocks * 4))
