"""Kernel time of ONE 10 s utterance through the pipelined NoiseSup kernel (and, for comparison, the batch kernel), from
the library's own CUDA-event counters; also the command to put under ncu for the per-role stall picture.
Run on the GPU box:  python tests/tools/pipe_profile.py [reps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from speech_enhancement_b200 import api, synth      # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
x8 = synth.utterance(1, 8000, 80000)
x16 = synth.utterance(2, 16000, 160000)
with api.Context(0) as ctx:
    for mode, name in ((1, "batch kernel (1 warp)"), (2, "pipelined kernel")):
        ctx.set_ns_pipeline(mode)
        for prof, kid, fn in ((8, api.K_NS_P8, lambda: ctx.etsi_denoise(x8, [0, len(x8)], ceps=True)),
                              (16, api.K_NS_P16, lambda: ctx.p16_wiener(x16, [0, len(x16)]))):
            fn()
            ctx.reset_kernel_times()
            ctx.profiling(True)
            for _ in range(reps):
                fn()
            ms, n = ctx.kernel_time(kid)
            ctx.profiling(False)
            print("P%d %-24s kernel %.3f ms per 10 s utterance (%d launches)" % (prof, name, ms / max(n, 1), n))
    # per-role wait cycles (only in a -DSE_PIPE_STATS build: SE_B200_LIB=.../variants/pstats/libse_b200.so)
    import ctypes as C
    L = api.lib()
    if hasattr(L, "se_debug_pipe_stats"):
        buf = (C.c_ulonglong * 32)()
        for prof, fn in ((8, lambda: ctx.etsi_denoise(x8, [0, len(x8)], ceps=True)), (16, lambda: ctx.p16_wiener(x16, [0, len(x16)]))):
            ctx.set_ns_pipeline(2)
            fn()
            L.se_debug_pipe_stats(buf, 1)
            fn()
            ctx.sync()
            L.se_debug_pipe_stats(buf, 1)
            v = np.array(list(buf), np.float64).reshape(8, 4)
            for w in range(8 if prof == 8 else 5):
                tot = v[w, 2]
                print("P%d warp %d: total %.0f cycles, waiting for input %.1f %%, for output room %.1f %%, busy %.1f %% (%.0f cycles per block)"
                      % (prof, w, tot, 100 * v[w, 0] / tot, 100 * v[w, 1] / tot, 100 * (tot - v[w, 0] - v[w, 1]) / tot,
                         (tot - v[w, 0] - v[w, 1]) / 1000.0))
