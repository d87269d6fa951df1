#!/usr/bin/env python
"""bench.py — audio-seconds per second (xRT) of the speech-enhancement hot path on B200.

  python bench.py [--gpus N --steps K --warmup W] [--workload p16|p8|resynth]      our arm (CUDA, through the C ABI)
  python bench.py --impl reference [--workload p16|p8|resynth] [...]               the reference's own CPU code

A "step" is one pass of the hot path over one batch of synthetic utterances.  Headline workload (the default) =
BASELINE.json configs[1]: 10k synthetic 16 kHz noisy utterances (ragged 2-10 s) through the plugin NoiseSup (framed
rfft + two-stage Wiener filter, etsi_denoise_mapping_func_Wiener) -> denoised stream, 25 band gains per frame,
frame-dropping flags.  The two paths BASELINE.json's metric names -- P8 ETSI NoiseSup+CompCeps (etsi_denoise + the
CompCeps step) and the 64-sub IBM resynthesis (resyth_64sub_IBM) -- are measured in the same run with the same key set
(value, e2e, roofline, cpu_baseline) under "workloads"; --workload makes one of them the headline.  Everything else on
the path (feature chains, sub-band / IBM kernels, extract_Wiener chain) is under "other_workloads" (device-resident).

Weak scaling: every rank (one per GPU) processes its own batch of the same size; there is no collective on the data
path (utterances are independent).  torch is used for device memory, the stream/events and torch.distributed only.
The reference arm and the cpu_baseline legs time oracle/_ref (the reference's own sources compiled in place) through
oracle/ref_bench.c: forked C workers over the in-memory batch, no Python in the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

P16_BYTES_PER_BLOCK = 320 + 640 + 100 + 20      # int16 in, f32 denoised out, 25 f32 gains, 5 int32 flags
P8_BYTES_PER_HOP = 160 + 160 + 56               # int16 in, int16 out, 14 f32 cepstra
G64_BYTES_PER_SAMPLE = 2 + 2 + 64 * 4 / 160.0   # int16 in, int16 out, float mask row per 160 samples


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="p16", choices=["p16", "p8", "resynth"],
                    help="headline workload: p16 = configs[1] (default), p8 = ETSI NoiseSup+CompCeps, resynth = resyth_64sub_IBM")
    ap.add_argument("--utts", type=int, default=10000, help="utterances per GPU in the p16 batch")
    ap.add_argument("--gpu-map", default="stride", choices=["stride", "dense"],
                    help="rank k -> GPU k * (visible / N) (stride: ranks spread over the box's PCIe root ports) or GPU k")
    ap.add_argument("--exact", action="store_true", help="NoiseSup family in the exact arithmetic mode (se_ctx_set_exact); for --profile-only runs")
    ap.add_argument("--no-extra", action="store_true", help="skip the other_workloads section")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--chunk-mb", type=int, default=0, help="override the SE_MEM_HOST pipeline chunk size (MiB)")
    ap.add_argument("--chunk-utts", type=int, default=0, help="override the pipeline's minimum utterances per chunk")
    ap.add_argument("--pipe-mode", type=int, default=-1, help="se_ctx_set_pipeline_lead bits (A/B): 1 lead chunk, 2 upload ahead")
    ap.add_argument("--profile-only", default="", choices=["", "p16", "p8", "p8feat", "etsi16k", "resynth", "ibm", "subband", "wiener"],
                    help="run just one of the other_workloads (for ncu) and print its dict")
    return ap.parse_args()


def make_corpus(first_u, count, fs, min_s, max_s, seconds=None, components=False):
    """Ragged synthetic corpus, generated with all host threads (the C generator releases the GIL)."""
    from speech_enhancement_b200 import synth
    lens = np.array([int(seconds * fs) if seconds else synth.ragged_length(first_u + i, fs, min_s, max_s)
                     for i in range(count)], np.int64)
    offsets = np.zeros(count + 1, np.int64)
    offsets[1:] = np.cumsum(lens)
    total = int(offsets[-1])
    noisy = np.empty(total, np.int16)
    L = synth._load()

    def gen(i):
        L.se_synth_utterance(first_u + i, fs, int(lens[i]), noisy[offsets[i]:].ctypes.data, None, None)

    with ThreadPoolExecutor(max_workers=max(1, min(32, os.cpu_count() or 1))) as ex:
        list(ex.map(gen, range(count)))
    return noisy, offsets


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.stop = threading.Event()
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except ValueError:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
class Dist:
    """torch.distributed plumbing (barrier + max over ranks); a no-op at N=1 without torchrun."""

    def __init__(self, torch, n_gpus, device=None):
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0")) if device is None else device    # CUDA device index of this rank
        self.on = self.world > 1
        if self.on:
            import torch.distributed as dist
            self.dist = dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            # stdout carries exactly one JSON line: NCCL's version banner / debug log goes to stderr
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        if self.on:
            self.dist.barrier()

    def max(self, v):
        if not self.on:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, v):
        if not self.on:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.on:
            self.dist.destroy_process_group()


def timed_steps(torch, dist, stream, step, warmup, steps):
    """W untimed + exactly K timed steps, CUDA events on the launching stream, barrier+sync both sides;
    returns total milliseconds (max over ranks)."""
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    return dist.max(e0.elapsed_time(e1))


def host_timed_steps(torch, dist, step, warmup, steps):
    """Same bracket for the end-to-end leg (synchronous host-buffer calls): wall clock between syncs."""
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    dist.barrier()
    return dist.max((t1 - t0) * 1e3)


def e2e_timed(torch, dist, step, steps):
    """End-to-end leg: (solo_ms, all_ms).  With several ranks, rank 0 first runs the same K steps ALONE while the others
    wait (solo_ms: what one GPU reaches on this box when it has the host<->device links to itself), then all ranks run
    them together (all_ms, max over ranks) -- so the line can state e2e_efficiency = e2e_N / (N * e2e_solo) from one run."""
    solo = None
    if dist.on:
        if dist.rank == 0:
            step()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(steps):
                step()
            torch.cuda.synchronize()
            solo = (time.perf_counter() - t0) * 1e3
        dist.barrier()
    return solo, host_timed_steps(torch, dist, step, 1, steps)


def ncu_capture(workload, alg_bytes):
    """The committed ncu capture of the workload's dominant kernel(s) (profiles/traffic.json), if it was taken on this
    workload size; None otherwise."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)[workload]
        if abs(t["algorithmic_bytes_per_launch"] - alg_bytes) <= 0.001 * alg_bytes:
            return t
    except (OSError, KeyError, ValueError):
        pass
    return None


def roofline(workload, kernel, alg_bytes, kernel_ms, sm_mhz, sm_count=148):
    """roofline object of one workload: ALGORITHMIC bytes per launch / live-timed kernel duration against the measured
    HBM peak; traffic = DRAM bytes of the same launch from the committed ncu capture; issue_bound = executed
    warp-instructions (ncu) against the issue slots the SMs offered during the live-timed launch (what actually bounds
    these kernels)."""
    hbm, src = peaks()
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    t = ncu_capture(workload, alg_bytes)
    r = {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
         "traffic": t["traffic_bytes"] if t else None, "peak_source": src, "kernel": kernel, "kernel_ms": kernel_ms,
         "algorithmic_bytes_per_launch": alg_bytes, "issue_bound": None}
    if t and sm_mhz:
        slots = sm_count * 4 * sm_mhz * 1e6 * kernel_ms * 1e-3
        r["issue_bound"] = {"warp_instructions_per_launch": t["inst_executed"], "issue_slots_in_launch": slots,
                            "frac": t["inst_executed"] / slots}
    if t and sm_mhz and t.get("shared_wavefronts"):
        # the resource that actually binds the NoiseSup kernels: LDS / STS / SHFL wavefronts (ncu) against
        # one wavefront per SM and clock during the live-timed launch
        cycles = sm_count * sm_mhz * 1e6 * kernel_ms * 1e-3
        r["shared_memory_pipe"] = {"wavefronts_per_launch": t["shared_wavefronts"], "wavefront_slots_in_launch": cycles,
                                   "frac": t["shared_wavefronts"] / cycles}
    return r


def dev(torch, a):
    return torch.from_numpy(a).cuda()


def pinned(api, a):
    p = api.PinnedArray(a.shape, a.dtype)
    p.array[...] = a
    return p


def bench_p16(args, torch, api, ctx, dist, stream, first_u, corpus=None):
    """Headline: BASELINE configs[1]."""
    samples, offsets = corpus or make_corpus(first_u, args.utts, 16000, 2.0, 10.0)
    nu = args.utts
    bo = np.zeros(nu + 1, np.int64)
    bo[1:] = np.cumsum(np.diff(offsets) // 160)
    tb, total = int(bo[-1]), int(offsets[-1])
    audio_s = total / 16000.0
    d_in = dev(torch, samples)
    d_out = torch.empty(total, dtype=torch.float32, device="cuda")
    d_flags = torch.empty(5 * tb, dtype=torch.int32, device="cuda")
    d_gains = torch.empty(tb * 25, dtype=torch.float32, device="cuda")
    d_rows = torch.empty(nu, dtype=torch.int32, device="cuda")

    def step():
        ctx.p16_wiener_raw(d_in.data_ptr(), 0, offsets, nu, d_out.data_ptr(), d_flags.data_ptr(), d_gains.data_ptr(),
                           d_rows.data_ptr(), bo, mem=api.SE_MEM_DEVICE)

    ctx.profiling(False)
    for _ in range(args.warmup):
        step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    l0 = ctx.launch_count()
    with ClockSampler(dist.local) as clk:
        ms = timed_steps(torch, dist, stream, step, 0, args.steps)
    launches = ctx.launch_count() - l0
    k_ms, k_n = ctx.kernel_time(api.K_NS_P16)
    ctx.profiling(False)
    rows_total = int(d_rows.sum().item())
    if args.profile_only == "p16":          # device-resident leg only (ncu captures, kernel A/B runs)
        return {"workload": workload_name(args.utts), "xrt": audio_s * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
                "kernel_ms": k_ms / max(k_n, 1), "blocks": tb}

    # end to end: pinned host buffers through the same C-ABI call, H2D + D2H inside the timed region
    h_in = pinned(api, samples)
    h_out = api.PinnedArray(total, np.float32)
    h_flags = api.PinnedArray(5 * tb, np.int32)
    h_gains = api.PinnedArray(tb * 25, np.float32)
    h_rows = api.PinnedArray(nu, np.int32)

    def step_e2e():
        ctx.p16_wiener_raw(h_in.ptr, 0, offsets, nu, h_out.ptr, h_flags.ptr, h_gains.ptr, h_rows.ptr, bo,
                           mem=api.SE_MEM_HOST)

    solo_ms, e2e_ms = e2e_timed(torch, dist, step_e2e, args.steps)
    same = bool(np.array_equal(h_rows.array, d_rows.cpu().numpy()))

    # What extract_Wiener actually keeps (aurora_etsi_test.cpp:66-133 frees outData unused): gains + flags only.
    def step_feat():
        ctx.p16_wiener_raw(h_in.ptr, 0, offsets, nu, None, h_flags.ptr, h_gains.ptr, h_rows.ptr, bo, mem=api.SE_MEM_HOST)

    feat_ms = host_timed_steps(torch, dist, step_feat, 1, args.steps)
    h2d = samples.nbytes
    d2h = h_out.array.nbytes + h_flags.array.nbytes + h_gains.array.nbytes + h_rows.array.nbytes
    feat_d2h = h_flags.array.nbytes + h_gains.array.nbytes + h_rows.array.nbytes
    for p in (h_in, h_out, h_flags, h_gains, h_rows):
        p.free()
    del d_in, d_out, d_flags, d_gains, d_rows
    torch.cuda.empty_cache()
    return dict(name=workload_name(args.utts), audio_s=audio_s, ms=ms, launches=launches, k_ms=k_ms / max(k_n, 1),
                alg_bytes=tb * P16_BYTES_PER_BLOCK, kernel="ns_stream_kernel<16,int16>", blocks=tb, rows=rows_total,
                e2e_ms=e2e_ms, solo_ms=solo_ms, feat_ms=feat_ms, feat_d2h=feat_d2h,
                h2d=h2d, d2h=d2h, clocks=clk.summary(), e2e_matches_device=same,
                samples=samples, offsets=offsets, mask=None, frame_offsets=None)


def bench_p8(args, torch, api, ctx, dist, stream, first_u, n_utts=6000, features=False, full=False, corpus=None):
    """P8: ETSI NoiseSup + CompCeps on 10 s 8 kHz utterances (C1-type work, batched); features=True runs the
    whole feature chain (WaveProc -> CompCeps -> PostProc -> VAD, 15-column rows) behind NoiseSup.
    full=True (the metric's own path): also the end-to-end leg through se_etsi_denoise_batch with pinned HOST buffers
    and the raw numbers run_ours() needs for value / e2e / roofline / cpu_baseline."""
    samples, offsets = corpus or make_corpus(first_u, n_utts, 8000, 0, 0, seconds=10.0)
    total = int(offsets[-1])
    co = np.zeros(n_utts + 1, np.int64)
    co[1:] = np.cumsum(np.diff(offsets) // 80 if features else np.maximum(np.diff(offsets) // 80 - 6, 0))
    d_in = dev(torch, samples)
    d_den = torch.empty(total, dtype=torch.int16, device="cuda")
    d_ceps = torch.empty(int(co[-1]) * (15 if features else 14), dtype=torch.float32, device="cuda")
    d_nc = torch.empty(n_utts, dtype=torch.int32, device="cuda")

    def step():
        if features:
            ctx.etsi_features_raw(d_in.data_ptr(), offsets, n_utts, d_ceps.data_ptr(), co, d_nc.data_ptr(),
                                  d_den.data_ptr(), mem=api.SE_MEM_DEVICE)
        else:
            ctx.etsi_denoise_raw(d_in.data_ptr(), offsets, n_utts, d_den.data_ptr(), d_ceps.data_ptr(), co,
                                 d_nc.data_ptr(), mem=api.SE_MEM_DEVICE)

    for _ in range(args.warmup):
        step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    l0 = ctx.launch_count()
    with ClockSampler(dist.local) as clk:
        ms = timed_steps(torch, dist, stream, step, 0, args.steps)
    launches = ctx.launch_count() - l0
    k_ms, k_n = ctx.kernel_time(api.K_NS_P8)
    ctx.profiling(False)
    audio_s = total / 8000.0
    hops = total // 80
    hbm, _ = peaks()
    alg_bytes = hops * (P8_BYTES_PER_HOP + (4 if features else 0))
    gbs = alg_bytes / (k_ms / max(k_n, 1) * 1e-3) / 1e9
    name = "P8 etsi NoiseSup+WaveProc+CompCeps+PostProc+VAD" if features else "P8 etsi NoiseSup+CompCeps"
    out = {"workload": "%s, %d x 10 s 8 kHz" % (name, n_utts), "xrt": audio_s * args.steps / (ms * 1e-3),
           "ms_per_step": ms / args.steps, "kernel_ms": k_ms / max(k_n, 1), "achieved_gbs": gbs, "frac": gbs / hbm}
    if not full:
        return out
    # end to end: pinned host buffers through the same C-ABI call, H2D + D2H inside the timed region
    rows_dev = d_nc.cpu().numpy()
    h_in = pinned(api, samples)
    h_den = api.PinnedArray(total, np.int16)
    h_ceps = api.PinnedArray(int(co[-1]) * 14, np.float32)
    h_nc = api.PinnedArray(n_utts, np.int32)

    def step_e2e():
        ctx.etsi_denoise_raw(h_in.ptr, offsets, n_utts, h_den.ptr, h_ceps.ptr, co, h_nc.ptr, mem=api.SE_MEM_HOST)

    solo_ms, e2e_ms = e2e_timed(torch, dist, step_e2e, args.steps)
    same = bool(np.array_equal(h_nc.array, rows_dev))
    d2h = h_den.array.nbytes + h_ceps.array.nbytes + h_nc.array.nbytes
    for q in (h_in, h_den, h_ceps, h_nc):
        q.free()
    del d_in, d_den, d_ceps, d_nc
    torch.cuda.empty_cache()
    return dict(name=out["workload"], audio_s=audio_s, ms=ms, launches=launches, k_ms=k_ms / max(k_n, 1), alg_bytes=alg_bytes,
                kernel="ns_stream_kernel<8,int16>", e2e_ms=e2e_ms, solo_ms=solo_ms, h2d=samples.nbytes, d2h=d2h, clocks=clk.summary(),
                e2e_matches_device=same, samples=samples, offsets=offsets, mask=None, frame_offsets=None)


def bench_etsi16k(args, torch, api, ctx, dist, stream, first_u, n_utts=3000):
    """ETSI 16 kHz mode (QMF split -> low-band NoiseSup/WaveProc/CompCeps + HP-band spectral subtraction and coding
    -> 26-band cepstra -> PostProc -> VAD) on 10 s 16 kHz utterances; 15-column feature rows + low-band int16 out."""
    samples, offsets = make_corpus(first_u, n_utts, 16000, 0, 0, seconds=10.0)
    total = int(offsets[-1])
    fo = np.zeros(n_utts + 1, np.int64)
    fo[1:] = np.cumsum(np.diff(offsets) // 160)
    do = fo * 80
    d_in = dev(torch, samples)
    d_feat = torch.empty(int(fo[-1]) * 15, dtype=torch.float32, device="cuda")
    d_den = torch.empty(int(do[-1]), dtype=torch.int16, device="cuda")
    d_nr = torch.empty(n_utts, dtype=torch.int32, device="cuda")

    def step():
        ctx.etsi16k_features_raw(d_in.data_ptr(), offsets, n_utts, d_feat.data_ptr(), fo, d_nr.data_ptr(),
                                 d_den.data_ptr(), do, mem=api.SE_MEM_DEVICE)

    for _ in range(args.warmup):
        step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    ms = timed_steps(torch, dist, stream, step, 0, args.steps)
    k_ms, k_n = ctx.kernel_time(api.K_NS_P8)
    ctx.profiling(False)
    audio_s = total / 16000.0
    calls = total // 160
    hbm, _ = peaks()
    gbs = calls * (320 + 160 + 60) / (k_ms / max(k_n, 1) * 1e-3) / 1e9      # 160 int16 in, 80 int16 + 15 f32 out
    return {"workload": "ETSI 16 kHz mode feature chain, %d x 10 s 16 kHz" % n_utts,
            "xrt": audio_s * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
            "kernel_ms": k_ms / max(k_n, 1), "achieved_gbs": gbs, "frac": gbs / hbm}


def bench_ibm(args, torch, api, ctx, dist, stream, first_u, n_utts=2000, seconds=5.0):
    """G64: fused subbband(pure) + subbband(noise) + make_single_IBM -> IRM + packed IBM bits (exact mode)."""
    from speech_enhancement_b200 import synth
    n = int(seconds * 16000)
    offsets = np.arange(n_utts + 1, dtype=np.int64) * n
    pure = np.empty(n_utts * n, np.int16)
    noise = np.empty(n_utts * n, np.int16)
    L = synth._load()

    def gen(i):
        L.se_synth_utterance(first_u + i, 16000, n, None, pure[i * n:].ctypes.data, noise[i * n:].ctypes.data)

    with ThreadPoolExecutor(max_workers=max(1, min(32, os.cpu_count() or 1))) as ex:
        list(ex.map(gen, range(n_utts)))
    nf = api.ibm_num_frames(n)
    fo = np.arange(n_utts + 1, dtype=np.int64) * nf
    d_p, d_n = dev(torch, pure), dev(torch, noise)
    d_irm = torch.empty(n_utts * nf * 64, dtype=torch.float32, device="cuda")
    d_bits = torch.empty(n_utts * nf, dtype=torch.int64, device="cuda")

    def step():
        ctx.ibm_raw(d_p.data_ptr(), d_n.data_ptr(), offsets, n_utts, d_irm.data_ptr(), d_bits.data_ptr(), fo,
                    mem=api.SE_MEM_DEVICE)

    step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    steps = min(args.steps, 2)
    ms = timed_steps(torch, dist, stream, step, 0, steps)
    k_ms, k_n = ctx.kernel_time(api.K_IBM)
    ctx.profiling(False)
    audio_s = n_utts * seconds
    hbm, _ = peaks()
    gbs = (2 * 2 * n_utts * n + n_utts * nf * (256 + 8)) / (k_ms / max(k_n, 1) * 1e-3) / 1e9
    return {"workload": "G64 subband x2 + make_single_IBM (exact), %d x %.0f s 16 kHz" % (n_utts, seconds),
            "xrt": audio_s * steps / (ms * 1e-3), "ms_per_step": ms / steps, "kernel_ms": k_ms / max(k_n, 1),
            "achieved_gbs": gbs, "frac": gbs / hbm}


def bench_subband(args, torch, api, ctx, dist, stream, first_u, n_utts=1000, seconds=5.0):
    """G64: subbband() materialised like the reference (64 int16 channel rows per utterance, exact mode)."""
    samples, offsets = make_corpus(first_u, n_utts, 16000, 0, 0, seconds=seconds)
    total = int(offsets[-1])
    d_in = dev(torch, samples)
    d_out = torch.empty(64 * total, dtype=torch.int16, device="cuda")

    def step():
        ctx.subband_raw(d_in.data_ptr(), offsets, n_utts, d_out.data_ptr(), api.SE_FLAG_EXACT, mem=api.SE_MEM_DEVICE)

    step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    steps = min(args.steps, 2)
    ms = timed_steps(torch, dist, stream, step, 0, steps)
    k_ms, k_n = ctx.kernel_time(api.K_SUBBAND)
    ctx.profiling(False)
    hbm, _ = peaks()
    gbs = (2 + 128) * total / (k_ms / max(k_n, 1) * 1e-3) / 1e9
    return {"workload": "G64 subbband (64 int16 rows, exact), %d x %.0f s 16 kHz" % (n_utts, seconds),
            "xrt": total / 16000.0 * steps / (ms * 1e-3), "ms_per_step": ms / steps, "kernel_ms": k_ms / max(k_n, 1),
            "achieved_gbs": gbs, "frac": gbs / hbm}


def bench_extract_wiener(args, torch, api, ctx, dist, stream, first_u, n_utts=2368, seconds=5.0):
    """extract_Wiener, the literal enhancement_extract_subband flow behind configs[1]: subbband() -> 64 channel signals
    -> func_Wiener over all 64 with one carried state (aurora_etsi_test.cpp:16-141), device-resident.  One utterance =
    one 64 x 5 s stream on one warp, so 2368 = 148 SMs x 16 warps utterances make exactly one wave."""
    samples, offsets = make_corpus(first_u, n_utts, 16000, 0, 0, seconds=seconds)
    total = int(offsets[-1])
    bo = np.zeros(n_utts + 1, np.int64)
    bo[1:] = np.cumsum(64 * (np.diff(offsets) // 160))
    d_in = dev(torch, samples)
    d_gains = torch.empty(int(bo[-1]) * 25, dtype=torch.float32, device="cuda")
    d_flags = torch.empty(int(bo[-1]) * 5, dtype=torch.int32, device="cuda")
    d_rows = torch.empty(n_utts, dtype=torch.int32, device="cuda")

    def step():
        api._check(ctx.L.se_extract_wiener_batch(ctx.h, d_in.data_ptr(), offsets.ctypes.data, n_utts, d_gains.data_ptr(),
                                                 d_flags.data_ptr(), d_rows.data_ptr(), bo.ctypes.data, api.SE_MEM_DEVICE),
                   "se_extract_wiener_batch")

    step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    steps = min(args.steps, 2)
    ms = timed_steps(torch, dist, stream, step, 0, steps)
    ks_ms, ks_n = ctx.kernel_time(api.K_SUBBAND)
    kn_ms, kn_n = ctx.kernel_time(api.K_NS_P16)
    ctx.profiling(False)
    hbm, _ = peaks()
    # compulsory bytes: int16 in, 25 gains + 5 flags per 160-sample block of each of the 64 channels out
    gbs = (2 * total + 120 * int(bo[-1])) / (ms / steps * 1e-3) / 1e9
    return {"workload": "extract_Wiener chain (subbband -> 64 x func_Wiener, carried state), %d x %.0f s 16 kHz" % (n_utts, seconds),
            "xrt": total / 16000.0 * steps / (ms * 1e-3), "ms_per_step": ms / steps,
            "kernel_ms": {"subband": ks_ms / max(ks_n, 1), "ns_p16": kn_ms / max(kn_n, 1)},
            "sub_band_audio_xrt": 64 * total / 16000.0 * steps / (ms * 1e-3), "achieved_gbs": gbs, "frac": gbs / hbm}


def resynth_inputs(first_u, n_utts):
    """configs[2]: n_utts x 10 s at 16 kHz + a random IBM (55 % of the T-F units on)."""
    samples, offsets = make_corpus(first_u, n_utts, 16000, 0, 0, seconds=10.0)
    nf = (160000 - 320) // 160 + 1                  # numFrame, resyth_64sub_IBM/cpp/main.cpp:110
    fo = np.arange(n_utts + 1, dtype=np.int64) * nf
    mask = (np.random.default_rng(first_u).random((n_utts * nf, 64)) > 0.45).astype(np.float32)
    return samples, offsets, mask, fo


def bench_resynth(args, torch, api, ctx, dist, stream, first_u, n_utts=1000, full=False, corpus=None):
    """G64: resyth_64sub_IBM on 10 s 16 kHz utterances (BASELINE configs[2] shape).  full=True adds the end-to-end leg
    through se_resynth_batch with pinned HOST buffers and returns the raw numbers for run_ours()."""
    samples, offsets, mask, fo = corpus or resynth_inputs(first_u, n_utts)
    total = int(offsets[-1])
    assert fo[1] == api.resynth_num_frames(160000, api.RESYNTH_IBM)
    d_in, d_mask = dev(torch, samples), dev(torch, mask)
    d_out = torch.empty(total, dtype=torch.int16, device="cuda")

    def step():
        ctx.resynth_raw(d_in.data_ptr(), offsets, n_utts, d_mask.data_ptr(), fo, api.RESYNTH_IBM, 0, d_out.data_ptr(),
                        mem=api.SE_MEM_DEVICE)

    w = min(args.warmup, 3 if full else 1)
    for _ in range(w):
        step()
    ctx.sync()
    ctx.reset_kernel_times()
    ctx.profiling(True)
    steps = args.steps if full else min(args.steps, 2)
    l0 = ctx.launch_count()
    with ClockSampler(dist.local) as clk:
        ms = timed_steps(torch, dist, stream, step, 0, steps)
    launches = ctx.launch_count() - l0
    f_ms, f_n = ctx.kernel_time(api.K_GT_FWD)
    b_ms, b_n = ctx.kernel_time(api.K_GT_BWD)
    ctx.profiling(False)
    audio_s = total / 16000.0
    hbm, _ = peaks()
    k = (f_ms + b_ms) / steps
    alg_bytes = int(total * G64_BYTES_PER_SAMPLE)
    gbs = alg_bytes / (k * 1e-3) / 1e9
    out = {"workload": "G64 resyth_64sub_IBM, %d x 10 s 16 kHz" % n_utts, "xrt": audio_s * steps / (ms * 1e-3),
           "ms_per_step": ms / steps, "kernel_ms": {"gt_fwd": f_ms / steps, "gt_bwd": b_ms / steps},
           "achieved_gbs": gbs, "frac": gbs / hbm}
    if not full:
        return out
    ref_out = d_out.cpu().numpy()
    h_in, h_mask = pinned(api, samples), pinned(api, mask)
    h_out = api.PinnedArray(total, np.int16)

    def step_e2e():
        ctx.resynth_raw(h_in.ptr, offsets, n_utts, h_mask.ptr, fo, api.RESYNTH_IBM, 0, h_out.ptr, mem=api.SE_MEM_HOST)

    solo_ms, e2e_ms = e2e_timed(torch, dist, step_e2e, steps)
    same = bool(np.array_equal(h_out.array, ref_out))
    h2d, d2h = samples.nbytes + mask.nbytes, h_out.array.nbytes
    for q in (h_in, h_mask, h_out):
        q.free()
    del d_in, d_mask, d_out
    torch.cuda.empty_cache()
    return dict(name=out["workload"], audio_s=audio_s, ms=ms, steps=steps, launches=launches, k_ms=k, alg_bytes=alg_bytes,
                kernel="gt_fwd_fast_kernel + gt_bwd_fast_kernel<IBM> (sum of the two launches)",
                kernel_split={"gt_fwd": f_ms / steps, "gt_bwd": b_ms / steps}, e2e_ms=e2e_ms, solo_ms=solo_ms, h2d=h2d, d2h=d2h,
                clocks=clk.summary(), e2e_matches_device=same, samples=samples, offsets=offsets, mask=mask, frame_offsets=fo)


# ------------------------------------------------------------------------------------------------
# the reference's own CPU code (oracle/_ref, or the plain-C port when _ref is absent)
# ------------------------------------------------------------------------------------------------
FS = {"p16": 16000, "p8": 8000, "resynth": 16000}


def cpu_leg(workload, samples, offsets, mask=None, frame_offsets=None, all_cores_s=4.0, one_core_s=3.0):
    """The reference's CPU implementation (oracle/_ref, else the port) on a BOUNDED sample of the workload's own batch:
    the first n utterances, n sized from a calibration pass so that the all-cores leg takes about all_cores_s seconds
    of wall clock and the 1-worker leg about one_core_s.  Returns the cpu_baseline object."""
    from oracle import pyoracle
    rb = pyoracle.RefBench(workload)
    cores = os.cpu_count() or 1
    nu = len(offsets) - 1
    fs = FS[workload]

    def run(n, workers):
        n = max(1, min(n, nu))
        sec, _ = rb.run(samples, offsets[:n + 1], workers, mask, None if frame_offsets is None else frame_offsets[:n + 1])
        return float(offsets[n]) / fs, sec, n

    a0, t0, n0 = run(cores, cores)                        # calibration (also warms the page cache / library)
    a_all, t_all, n_all = run(int(n0 * all_cores_s / max(t0, 1e-3)), cores)
    a_one, t_one, n_one = run(int(n_all * (one_core_s / max(t_all, 1e-3)) / cores), 1)
    return {"value": a_all / t_all, "unit": UNIT, "cores": cores, "kind": rb.kind,
            "sample": "first %d utterances of the workload (%.0f audio-s) on %d forked C workers, %.2f s wall"
                      % (n_all, a_all, cores, t_all),
            "one_thread": {"value": a_one / t_one, "unit": UNIT, "cores": 1,
                           "sample": "first %d utterances (%.0f audio-s), %.2f s wall" % (n_one, a_one, t_one)},
            "library": os.path.relpath(rb.lib_path, ROOT)}


METRIC = "audio-sec/sec (xRT) per B200 for NoiseSup+CompCeps & 64-sub IBM resynth, 1/2/4/8 GPU"
UNIT = "audio-s/s"


def workload_name(utts):
    return ("configs[1]: enhancement_extract_subband batch, %d synthetic 16 kHz noisy utterances (ragged 2-10 s), "
            "framed rfft+NoiseSup features (etsi_denoise_mapping_func_Wiener)" % utts)


P8_UTTS, RESYNTH_UTTS = 6000, 1000
OFFS = {"p16": 0, "p8": 500000, "resynth": 700000}      # first synthetic utterance id of each workload


def workload_label(w, args):
    return {"p16": workload_name(args.utts), "p8": "P8 etsi NoiseSup+CompCeps, %d x 10 s 8 kHz" % P8_UTTS,
            "resynth": "G64 resyth_64sub_IBM, %d x 10 s 16 kHz" % RESYNTH_UTTS}[w]


def workload_corpus(w, args, first_u, count=None):
    """(samples, offsets, mask, frame_offsets) of workload w (first `count` utterances; default = the whole batch)."""
    if w == "p16":
        return make_corpus(first_u + OFFS[w], count or args.utts, 16000, 2.0, 10.0) + (None, None)
    if w == "p8":
        return make_corpus(first_u + OFFS[w], count or P8_UTTS, 8000, 0, 0, seconds=10.0) + (None, None)
    return resynth_inputs(first_u + OFFS[w], count or RESYNTH_UTTS)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the chosen workload (oracle/_ref through
    oracle/ref_bench.c), all host cores, each step a bounded sample of the workload (its first n utterances, n sized
    so that a step takes about 3 s)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle
    w = args.workload
    rb = pyoracle.RefBench(w)
    cores = os.cpu_count() or 1
    fs = FS[w]
    total_utts = {"p16": args.utts, "p8": P8_UTTS, "resynth": RESYNTH_UTTS}[w]
    samples, offsets, mask, fo = workload_corpus(w, args, 0, min(total_utts, cores * {"p16": 96, "p8": 64, "resynth": 12}[w]))
    nu = len(offsets) - 1

    def run(n, workers):
        n = max(1, min(n, nu))
        sec, _ = rb.run(samples, offsets[:n + 1], workers, mask, None if fo is None else fo[:n + 1])
        return float(offsets[n]) / fs, sec, n

    a0, t0, n0 = run(cores, cores)                               # calibration
    n = max(cores, min(nu, int(n0 * 3.0 / max(t0, 1e-3))))
    for _ in range(min(args.warmup, 1)):
        run(n, cores)
    t_total, audio = 0.0, 0.0
    for _ in range(args.steps):
        a, t, _n = run(n, cores)
        t_total += t
        audio += a
    v = audio / t_total
    a1, t1, n1 = run(max(1, n // cores), 1)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_total / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_label(w, args), "sample": "first %d utterances per step" % n},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": rb.kind,
                             "sample": "first %d utterances of the workload per step (%.0f audio-s), %d steps, forked C workers"
                                       % (n, audio / args.steps, args.steps),
                             "one_thread": {"value": a1 / t1, "unit": UNIT, "cores": 1,
                                            "sample": "first %d utterances, %.2f s wall" % (n1, t1)},
                             "library": os.path.relpath(rb.lib_path, ROOT)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def bench_c1(args, torch, api, ctx, first_u):
    """configs[0]: ONE 10 s 8 kHz utterance through se_etsi_denoise_batch with host buffers (what the drop-in's
    etsi_denoise + CompCeps step does per call): median wall-clock latency of 15 calls."""
    x, off = make_corpus(first_u + OFFS["p8"], 1, 8000, 0, 0, seconds=10.0)
    den = np.zeros(len(x), np.int16)
    co = np.array([0, len(x) // 80 - 6], np.int64)
    ceps = np.zeros((int(co[1]), 14), np.float32)
    nc = np.zeros(1, np.int32)
    ts = []
    for _ in range(18):
        t0 = time.perf_counter()
        ctx.etsi_denoise_raw(x, off, 1, den, ceps, co, nc, mem=api.SE_MEM_HOST)
        ts.append(time.perf_counter() - t0)
    ms = float(np.median(ts[3:])) * 1e3
    return {"workload": "configs[0]: etsi NoiseSup+CompCeps on one 10 s 8 kHz utterance, per-call latency (host buffers)",
            "ms_per_call": ms, "value": 10.0 / (ms * 1e-3), "unit": UNIT,
            "note": "one utterance = one CTA of the pipelined kernel (csrc/se_ns_pipe.cuh: a block's work over six warps); "
                    "the recursion is serial in time, so this is a latency figure -- throughput lives in the batch calls"}


def assemble(w, r, args, dist, cpu):
    """value / e2e / roofline / cpu_baseline of one workload from its raw numbers (all ranks call this: collectives)."""
    steps = r.get("steps", args.steps)
    audio_all = dist.sum(r["audio_s"])
    e2e = audio_all * steps / (r["e2e_ms"] * 1e-3)
    d = {"workload": r["name"], "value": audio_all * steps / (r["ms"] * 1e-3), "unit": UNIT,
         "ms_per_step": r["ms"] / steps, "audio_s_per_gpu_per_step": r["audio_s"],
         "roofline": roofline(w, r["kernel"], r["alg_bytes"], r["k_ms"], r["clocks"].get("sm_mhz")),
         "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                 "ms_per_step": r["e2e_ms"] / steps, "matches_device_path": r["e2e_matches_device"],
                 "host_link_gbs_per_gpu": (r["h2d"] + r["d2h"]) * steps / (r["e2e_ms"] * 1e-3) / 1e9},
         "gpu_launches": r["launches"], "clocks": r["clocks"]}
    if "kernel_split" in r:
        d["roofline"]["kernel_ms_split"] = r["kernel_split"]
    solo = dist.max(r["solo_ms"] or 0.0)              # rank 0's solo run, known to every rank after the reduction
    if dist.on and solo > 0:
        e2e_solo = r["audio_s"] * steps / (solo * 1e-3)
        d["e2e"]["solo_one_gpu"] = {"value": e2e_solo, "unit": UNIT,
                                    "note": "rank 0 alone on the same box, the other ranks idle, same run"}
        d["e2e"]["e2e_efficiency"] = e2e / (args.gpus * e2e_solo)
    if "feat_ms" in r:
        d["e2e"]["features_only"] = {"value": audio_all * steps / (r["feat_ms"] * 1e-3), "unit": UNIT,
                                     "ms_per_step": r["feat_ms"] / steps, "d2h_bytes_per_step": r["feat_d2h"],
                                     "note": "same call with out_f32 = NULL: gains + flags only, which is all the "
                                             "reference's extract_Wiener driver keeps (aurora_etsi_test.cpp:66-133)"}
    if cpu:
        d["cpu_baseline"] = cpu
    return d


def run_ours(args):
    import torch
    from speech_enhancement_b200 import api
    local = int(os.environ.get("LOCAL_RANK", "0"))
    rank = int(os.environ.get("RANK", "0"))
    first_u = rank * 1000003                        # every rank gets its own utterances (weak scaling)
    order = [args.workload] + [w for w in ("p16", "p8", "resynth") if w != args.workload]

    # inputs of the three workloads, and (rank 0, N = 1 only) the CPU baselines on the same batches -- before any CUDA
    # context exists in this process: the CPU legs fork
    corpora, cpu = {}, {}
    if not args.profile_only:
        for w in order:
            corpora[w] = workload_corpus(w, args, first_u)
            if rank == 0 and args.gpus == 1 and not args.no_cpu:
                cpu[w] = cpu_leg(w, *corpora[w])

    # rank -> GPU: spread the ranks over the box (N = 2 -> GPUs 0 and 4 of 8, ...) unless --gpu-map dense
    n_vis = torch.cuda.device_count()
    stride = n_vis // args.gpus if (args.gpu_map == "stride" and args.gpus > 1 and n_vis >= args.gpus) else 1
    device = local * max(stride, 1)
    torch.cuda.set_device(device)
    dist = Dist(torch, args.gpus, device)
    stream = torch.cuda.Stream()
    ctx = api.Context(device)
    ctx.set_stream(stream.cuda_stream)
    if args.exact:
        ctx.set_exact(True)
    if args.pipe_mode >= 0:
        ctx.set_pipeline_lead(args.pipe_mode)
    if args.chunk_mb or args.chunk_utts:
        ctx.set_pipeline((args.chunk_mb or 512) << 20, args.chunk_utts or 1024)

    if args.profile_only:
        with torch.cuda.stream(stream):
            fn = {"p16": bench_p16, "p8": bench_p8, "resynth": bench_resynth, "ibm": bench_ibm, "subband": lambda *a: bench_subband(*a, n_utts=int(os.environ.get("SE_SUB_UTTS", "1000")), seconds=float(os.environ.get("SE_SUB_SEC", "5"))),
                  "p8feat": lambda *a: bench_p8(*a, features=True), "etsi16k": bench_etsi16k, "wiener": bench_extract_wiener}[args.profile_only]
            # p16: the same corpus as the default run's headline launch (profiles/traffic.json is keyed on its bytes)
            emit(fn(args, torch, api, ctx, dist, stream, first_u if args.profile_only == "p16" else first_u + 500000))
        ctx.close()
        dist.close()
        return
    res, extra = {}, None
    with torch.cuda.stream(stream):
        for w in order:
            c = corpora.pop(w)
            if w == "p16":
                r = bench_p16(args, torch, api, ctx, dist, stream, first_u, corpus=c[:2])
            elif w == "p8":
                r = bench_p8(args, torch, api, ctx, dist, stream, first_u + OFFS[w], n_utts=P8_UTTS, full=True, corpus=c[:2])
            else:
                r = bench_resynth(args, torch, api, ctx, dist, stream, first_u + OFFS[w], n_utts=RESYNTH_UTTS, full=True, corpus=c)
            for k in ("samples", "offsets", "mask", "frame_offsets"):
                r.pop(k, None)
            res[w] = assemble(w, r, args, dist, cpu.get(w))
            del c
        c1 = bench_c1(args, torch, api, ctx, first_u) if args.gpus == 1 else None
        if not args.no_extra and args.gpus == 1:
            extra = [bench_p8(args, torch, api, ctx, dist, stream, first_u + 500000, features=True),
                     bench_etsi16k(args, torch, api, ctx, dist, stream, first_u + 600000),
                     bench_ibm(args, torch, api, ctx, dist, stream, first_u + 800000),
                     bench_subband(args, torch, api, ctx, dist, stream, first_u + 900000),
                     bench_extract_wiener(args, torch, api, ctx, dist, stream, first_u + 950000)]
    head = res[args.workload]
    line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": head["workload"], "audio_s_per_gpu_per_step": head["audio_s_per_gpu_per_step"],
                       "l2": "inputs larger than L2 (%.2f GB per step)" % (head["e2e"]["h2d_bytes_per_step"] / 1e9),
                       "parallelism": "utterance-sharded, %d rank(s), no collective" % args.gpus,
                       "gpu_map": "rank k -> GPU %d*k of %d visible" % (max(stride, 1), n_vis)},
            "roofline": head["roofline"], "e2e": head["e2e"], "gpu_launches": head["gpu_launches"], "clocks": head["clocks"]}
    if "cpu_baseline" in head:
        line["cpu_baseline"] = head["cpu_baseline"]
    line["workloads"] = {w: res[w] for w in order[1:]}
    if c1:
        if "p8" in cpu:
            c1["cpu_baseline"] = dict(cpu["p8"]["one_thread"], kind=cpu["p8"]["kind"],
                                      note="reference etsi_denoise + CompCeps step on one core (batch of 10 s utterances)")
        line["workloads"]["c1"] = c1
    if extra:
        line["other_workloads"] = extra
    if rank == 0:
        emit(line)
    ctx.close()
    dist.close()


_JSON_OUT = None


def emit(obj):
    """The one JSON line goes to the process's real stdout; everything else printed to fd 1 by libraries (NCCL's
    version banner ignores NCCL_DEBUG_FILE) was redirected to stderr in __main__."""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(obj) + "\n")
    out.flush()


if __name__ == "__main__":
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
