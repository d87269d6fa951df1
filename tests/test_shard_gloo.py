"""World-size-2 run of the host-side sharding logic over gloo (CPU): each rank takes its contiguous
utterance range from se_shard_plan, processes it (with the oracle standing in for the GPU — this is a
test of the partitioning and gathering, not of the kernels), and the gathered per-utterance results
must equal the single-process results.  There is no collective on the data path; gloo is only used
here to gather the results for the check."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    from oracle import pyoracle
    from speech_enhancement_b200 import api, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    samples, offsets = synth.corpus(300, 9, 8000, min_s=0.3, max_s=1.0)
    plan = api.shard_plan(offsets, world)
    lo, hi = int(plan[rank]), int(plan[rank + 1])
    o = pyoracle.Oracle()
    mine = {}
    for u in range(lo, hi):
        r = o.etsi_denoise(samples[offsets[u]:offsets[u + 1]], with_ceps=True)
        mine[u] = (int(r["den_i16"].astype(np.int64).sum()), len(r["ceps"]))
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        q.put((gathered, plan.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_covers_every_utterance_once():
    sys.path.insert(0, ROOT)
    from oracle import pyoracle
    from speech_enhancement_b200 import synth
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, plan = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    merged = {}
    for d in gathered:
        for k, v in d.items():
            assert k not in merged, "utterance %d processed twice" % k
            merged[k] = v
    assert sorted(merged) == list(range(9)) and plan[0] == 0 and plan[-1] == 9
    samples, offsets = synth.corpus(300, 9, 8000, min_s=0.3, max_s=1.0)
    o = pyoracle.Oracle()
    for u in range(9):
        r = o.etsi_denoise(samples[offsets[u]:offsets[u + 1]], with_ceps=True)
        assert merged[u] == (int(r["den_i16"].astype(np.int64).sum()), len(r["ceps"]))
