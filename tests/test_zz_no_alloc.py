"""se_resynth_noalloc (include/se_b200.h): the resynthesis of ONE utterance with no allocation in the call -- scratch in
a file-scope static __device__ array, work items planned on the device.  No counterpart in the reference.
Checked through the C ABI against the existing entry point se_resynth_batch on the
same inputs, on two calls in a row (the second one shorter, so the shared array still holds the first call's rows
beyond it): int16 within 1 LSB (the bar of tests/test_gpu_g64.py) -- and in fact identical, because both entries run the
same kernels over the same work-item plan.

Collected last on purpose (test_zz_*): a failure here must not hide other tests.
Passes alone in a fresh process and after the other tests (it creates its own context)."""
import numpy as np
import pytest

from speech_enhancement_b200 import api, synth

pytestmark = pytest.mark.gpu


@pytest.fixture()
def own_ctx():
    c = api.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("mode,exact", [(api.RESYNTH_IBM, False), (api.RESYNTH_1DNN, False), (api.RESYNTH_IBM, True)])
def test_noalloc_matches_the_batch_entry_twice_in_a_row(own_ctx, mode, exact):
    torch = pytest.importorskip("torch")
    ctx = own_ctx
    flags = api.SE_FLAG_EXACT if exact else 0
    rng = np.random.default_rng(40 + mode)
    for call, n in enumerate((16000 * 3 + 123, 16000 + 77) if not exact else (16000 + 41, 8000 + 5)):
        x = synth.utterance(500 + 10 * mode + call, 16000, n)
        nf = api.resynth_num_frames(n, mode)
        m = rng.random((nf, 64)).astype(np.float32)
        if mode != api.RESYNTH_IBM:
            m[rng.random((nf, 64)) < 0.2] = 0.0
        want = ctx.resynth(x, [0, n], m, [0, nf], mode, exact=exact)            # existing entry point, host buffers
        d_x, d_m = torch.from_numpy(x).cuda(), torch.from_numpy(m).cuda()
        d_out = torch.full((n,), 12345, dtype=torch.int16, device="cuda")
        torch.cuda.synchronize()
        ctx.resynth_noalloc_raw(d_x.data_ptr(), n, d_m.data_ptr(), mode, flags, d_out.data_ptr())
        ctx.sync()
        got = d_out.cpu().numpy()
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 1, "call %d: max |d| = %d LSB" % (call, d.max())
        assert np.array_equal(got, want), "call %d: same kernels, same plan -- expected identical output" % call
        assert np.any(got != 0)


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(own_ctx):
    torch = pytest.importorskip("torch")
    n = 200001                                    # SE_RESYNTH_NOALLOC_MAX_SAMPLES + 1
    d_x = torch.zeros(n, dtype=torch.int16, device="cuda")
    d_m = torch.zeros(api.resynth_num_frames(n, 0) * 64, dtype=torch.float32, device="cuda")
    d_out = torch.zeros(n, dtype=torch.int16, device="cuda")
    with pytest.raises(api.SeError, match="exceed the fixed scratch"):
        own_ctx.resynth_noalloc_raw(d_x.data_ptr(), n, d_m.data_ptr(), 0, 0, d_out.data_ptr())
    # the largest accepted size runs
    n = 200000
    own_ctx.resynth_noalloc_raw(d_x.data_ptr(), n, d_m.data_ptr(), 0, 0, d_out.data_ptr())
    own_ctx.sync()
    assert not d_out[:n].any().item()             # zero input -> zero output
