"""BASELINE.json configs[4] on ONE B200: a 100 h synthetic noisy corpus (36 000 x 10 s) through the full path --
  (A) 8 kHz : ETSI NoiseSup + CompCeps                                   (se_etsi_denoise_batch)
  (B) 16 kHz: addnoise -> fused subbband x2 + make_single_IBM -> IBM-masked resynthesis of the mixture
             (se_addnoise_batch, se_ibm_batch, se_resynth_batch), chained on the device
in utterance chunks, host buffers in and out (pinned), corpus generated on the host cores while the GPU works on the
previous chunk is NOT attempted: generation and GPU time are reported separately.  Prints one JSON line.

Usage: python tools/c5_pipeline.py [hours=100] [chunk_utts=3000]
   or: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/c5_pipeline.py ...
       (one rank per GPU, contiguous utterance shards, no collective on the data path; a gloo all_gather collects the
       per-rank wall times and checksums at the end).  The checksum is a plain sum over per-utterance strided samples, so
       it is the same number for any sharding / chunking iff every utterance's outputs are bit-identical.
"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                             # noqa: E402  (device buffers for the chained G64 half)
from speech_enhancement_b200 import api, synth          # noqa: E402

HOURS = float(sys.argv[1]) if len(sys.argv) > 1 else 100.0
CHUNK = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
N_ALL = int(round(HOURS * 3600 / 10.0))
RANK, WORLD, LOCAL = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
U0, U1 = N_ALL * RANK // WORLD, N_ALL * (RANK + 1) // WORLD     # this rank's contiguous shard of the corpus
N = U1 - U0
T = max(1, min(32, (os.cpu_count() or 1) // WORLD))
torch.cuda.set_device(LOCAL)
ctx = api.Context(LOCAL)
L = synth._load()
gen_s = gpu_a = gpu_b = 0.0
rows_a = frames_b = 0
checksum = 0


def corpus16(first, count, n):
    pure = np.empty(count * n, np.int16)
    noise = np.empty(count * n, np.int16)

    def gen(i):
        L.se_synth_utterance(first + i, 16000, n, None, pure[i * n:].ctypes.data, noise[i * n:].ctypes.data)

    with ThreadPoolExecutor(T) as ex:
        list(ex.map(gen, range(count)))
    return pure, noise


for c0 in range(U0, U1, CHUNK):
    cnt = min(CHUNK, U1 - c0)
    # ---- (A) 8 kHz ETSI front end, host buffers through the C ABI
    t0 = time.perf_counter()
    s8, o8 = synth.corpus(200000 + c0, cnt, 8000, seconds=10.0, threads=T)
    gen_s += time.perf_counter() - t0
    t0 = time.perf_counter()
    r = ctx.etsi_denoise(s8, o8, ceps=True)
    gpu_a += time.perf_counter() - t0
    rows_a += int(r["n_ceps"].sum())
    checksum = (checksum + int(r["den_i16"].reshape(cnt, 80000)[:, ::997].astype(np.int64).sum())) & 0xffffffff
    del s8, r
    # ---- (B) 16 kHz: mixture -> masks -> resynthesis, chained on the device
    n = 160000
    t0 = time.perf_counter()
    pure, noise = corpus16(400000 + c0, cnt, n)
    gen_s += time.perf_counter() - t0
    offs = np.arange(cnt + 1, dtype=np.int64) * n
    nf = api.ibm_num_frames(n)
    fo = np.arange(cnt + 1, dtype=np.int64) * nf
    t0 = time.perf_counter()
    d_pure = torch.from_numpy(pure).cuda()
    d_noise = torch.from_numpy(noise).cuda()
    d_noisy = torch.empty_like(d_pure)
    d_irm = torch.empty(cnt * nf * 64, dtype=torch.float32, device="cuda")
    d_bits = torch.empty(cnt * nf, dtype=torch.int64, device="cuda")
    d_out = torch.empty_like(d_pure)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    api._check(ctx.L.se_addnoise_batch(ctx.h, d_pure.data_ptr(), d_noise.data_ptr(), d_noisy.data_ptr(),
                                       offs.ctypes.data, cnt, 0, api.SE_MEM_DEVICE), "se_addnoise_batch")
    ctx.ibm_raw(d_pure.data_ptr(), d_noise.data_ptr(), offs, cnt, d_irm.data_ptr(), d_bits.data_ptr(), fo,
                mem=api.SE_MEM_DEVICE)
    ctx.resynth_raw(d_noisy.data_ptr(), offs, cnt, d_irm.data_ptr(), fo, api.RESYNTH_IBM, 0, d_out.data_ptr(),
                    mem=api.SE_MEM_DEVICE)
    out = d_out.cpu().numpy()
    bits = d_bits.cpu().numpy()
    ctx.set_stream(None)
    gpu_b += time.perf_counter() - t0
    frames_b += cnt * nf
    checksum = (checksum + int(out.reshape(cnt, n)[:, ::997].astype(np.int64).sum())
                + int((bits.reshape(cnt, nf)[:, ::101] & 0xffff).sum())) & 0xffffffff
    del pure, noise, d_pure, d_noise, d_noisy, d_irm, d_bits, d_out, out, bits
    torch.cuda.empty_cache()

if WORLD > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
    mine = torch.tensor([gpu_a, gpu_b, gen_s, float(rows_a), float(frames_b), float(checksum)], dtype=torch.float64)
    allr = [torch.zeros_like(mine) for _ in range(WORLD)]
    dist.all_gather(allr, mine)
    dist.destroy_process_group()
    allr = torch.stack(allr).numpy()
    gpu_a, gpu_b, gen_s = float(allr[:, 0].max()), float(allr[:, 1].max()), float(allr[:, 2].max())
    rows_a, frames_b = int(allr[:, 3].sum()), int(allr[:, 4].sum())
    checksum = int(sum(int(v) for v in allr[:, 5])) & 0xffffffff
if RANK == 0:
    audio_s = N_ALL * 10.0
    print(json.dumps({
        "workload": "configs[4]: %.0f h synthetic corpus (%d x 10 s), 8 kHz ETSI NoiseSup+CompCeps and 16 kHz mix -> IBM/IRM -> "
                    "IBM-masked resynthesis, %d B200, chunks of %d utterances" % (HOURS, N_ALL, WORLD, CHUNK),
        "n_gpus": WORLD, "etsi_wall_s": gpu_a, "etsi_xrt": audio_s / gpu_a, "cepstral_frames": rows_a,
        "g64_wall_s": gpu_b, "g64_xrt": audio_s / gpu_b, "mask_frames": frames_b,
        "both_halves_xrt": audio_s / (gpu_a + gpu_b), "host_corpus_generation_s": gen_s, "host_threads_per_rank": T,
        "checksum": checksum,
        "note": "wall times (max over ranks) include every H2D/D2H copy (host arrays in, host arrays out); corpus "
                "generation is host work outside the measured path; the checksum does not depend on the sharding"}))
