"""PCIe floor of the end-to-end number: pinned D2H / H2D bandwidth alone and both directions at once."""
import torch

n = 1 << 30
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def d2h():
    h_out.copy_(d_a, non_blocking=True)


def h2d():
    d_b.copy_(h_in, non_blocking=True)


def both():
    with torch.cuda.stream(s1):
        h_out.copy_(d_a, non_blocking=True)
    with torch.cuda.stream(s2):
        d_b.copy_(h_in, non_blocking=True)


for name, fn, gb in (("D2H", d2h, 1), ("H2D", h2d, 1), ("both (each direction)", both, 1)):
    ms = timed(fn)
    print("%s: %.1f GB/s" % (name, gb * n / 1e9 / (ms * 1e-3)))
