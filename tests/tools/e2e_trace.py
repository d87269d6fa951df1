"""Timeline of one host-buffer (SE_MEM_HOST) batch call: when each chunk's H2D, kernel and D2H finish (SE_B200_TRACE).
Run on the GPU box:  python tests/tools/e2e_trace.py [p8|p16] [n_utts]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench                                         # noqa: E402
from speech_enhancement_b200 import api              # noqa: E402

prof = sys.argv[1] if len(sys.argv) > 1 else "p8"
nu = int(sys.argv[2]) if len(sys.argv) > 2 else (6000 if prof == "p8" else 10000)
if prof == "p8":
    samples, offsets = bench.make_corpus(0, nu, 8000, 0, 0, seconds=10.0)
else:
    samples, offsets = bench.make_corpus(0, nu, 16000, 2.0, 10.0)
total = int(offsets[-1])
with api.Context(0) as ctx:
    h_in = bench.pinned(api, samples)
    if prof == "p8":
        co = np.zeros(nu + 1, np.int64)
        co[1:] = np.cumsum(np.maximum(np.diff(offsets) // 80 - 6, 0))
        h_den = api.PinnedArray(total, np.int16)
        h_ceps = api.PinnedArray(int(co[-1]) * 14, np.float32)
        h_nc = api.PinnedArray(nu, np.int32)
        call = lambda: ctx.etsi_denoise_raw(h_in.ptr, offsets, nu, h_den.ptr, h_ceps.ptr, co, h_nc.ptr, mem=api.SE_MEM_HOST)
    else:
        bo = np.zeros(nu + 1, np.int64)
        bo[1:] = np.cumsum(np.diff(offsets) // 160)
        tb = int(bo[-1])
        h_out = api.PinnedArray(total, np.float32)
        h_flags = api.PinnedArray(5 * tb, np.int32)
        h_gains = api.PinnedArray(tb * 25, np.float32)
        h_rows = api.PinnedArray(nu, np.int32)
        call = lambda: ctx.p16_wiener_raw(h_in.ptr, 0, offsets, nu, h_out.ptr, h_flags.ptr, h_gains.ptr, h_rows.ptr, bo, mem=api.SE_MEM_HOST)
    for _ in range(3):
        call()
    t0 = time.perf_counter()
    call()
    print("call %.2f ms" % ((time.perf_counter() - t0) * 1e3))
    os.environ["SE_B200_TRACE"] = "1"
    call()
