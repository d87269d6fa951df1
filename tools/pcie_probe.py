"""Host<->device copy ceiling of the box, measured independently of the product: pinned D2H / H2D bandwidth alone and
both directions at once, on N ranks CONCURRENTLY (one per GPU).

  python tools/pcie_probe.py                                                     one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
         tools/pcie_probe.py [--gpu-map stride|dense]                            N GPUs at once

--gpu-map stride puts rank k on GPU k * (visible / N) (N = 2 -> GPUs 0 and 4 of 8), dense on GPU k.  Rank 0 prints one
JSON line: per-rank and aggregate GB/s for D2H alone, H2D alone, and each direction while both run.  The end-to-end
numbers of bench.py (e2e.host_link_gbs_per_gpu) are to be read against these."""
import argparse
import json
import os

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--gpu-map", default="stride", choices=["stride", "dense"])
ap.add_argument("--mb", type=int, default=1024)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
n_vis = torch.cuda.device_count()
stride = n_vis // world if (a.gpu_map == "stride" and world > 1 and n_vis >= world) else 1
device = local * max(stride, 1)
torch.cuda.set_device(device)
dist = None
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", device))

n = a.mb << 20
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if dist:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn):
    """GB/s per direction of this rank, all ranks copying at the same time."""
    fn()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    barrier()
    return n / 1e9 / (ms * 1e-3)


def d2h():
    h_out.copy_(d_a, non_blocking=True)


def h2d():
    d_b.copy_(h_in, non_blocking=True)


def both():
    s1.wait_stream(torch.cuda.current_stream())
    s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        h_out.copy_(d_a, non_blocking=True)
    with torch.cuda.stream(s2):
        d_b.copy_(h_in, non_blocking=True)


res = {}
for name, fn in (("d2h_alone", d2h), ("h2d_alone", h2d), ("each_direction_with_both_running", both)):
    v = timed(fn)
    if dist:
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        allv = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allv, t)
        per = [float(x.item()) for x in allv]
    else:
        per = [v]
    res[name] = {"per_rank_gbs": [round(x, 2) for x in per], "aggregate_gbs": round(sum(per), 2)}
if rank == 0:
    print(json.dumps({"n_gpus": world, "gpu_map": a.gpu_map, "devices": [k * max(stride, 1) for k in range(world)],
                      "visible_gpus": n_vis, "buffer_mb": a.mb, **res}))
if dist:
    dist.destroy_process_group()
