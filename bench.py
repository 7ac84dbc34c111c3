#!/usr/bin/env python
"""Headline benchmark: predict images/sec, ResNet-50 224^2 bf16, batch 256 per B200 replica
(BASELINE.json `metric`, configs[2]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path (K1 preprocess -> K2 forward -> K3 softmax/top-k) over one batch
of 256 synthetic decoded uint8 HWC images per GPU.  The K-step loop is repeated in ROUNDS until at least
2 s of device time have been measured; the line reports the MEDIAN round (steps = K) and the clocks are
sampled through all rounds.
  value  : whole-job images/s with the uint8 inputs already resident in HBM, CUDA-event timed.
  e2e    : the same metric through the host-buffer C-ABI call b2p_predict_host — pinned host slab in,
           pinned (prob, index) out, H2D and D2H inside the timed region.
  e2e_engine : the same through the PRODUCT's dispatcher: rank 0 builds runtime.Engine over GPUs 0..N-1 in one
           process (what init() does), feeds it whole 256-image batches (pinned slabs, then JPEG files) and
           reports images/s and the per-GPU share (host clock: it is a host-dispatch number).
  roofline     : K2 (the tcgen05 conv/linear launches), algorithmic FLOPs / CUDA-event time vs the measured
                 BURST bf16 peak in MEASURED_PEAKS.json; the sustained peak is reported beside it.
  cpu_baseline : the oracle (reference predict contract filled in with torch CPU fp32) timed on the
                 host cores on a bounded sample of the same workload (rank 0, N=1 only).
  determinism  : every rank runs the same seed-1 batch once; the line says whether the logits are bit-identical
                 on every GPU.
Multi-GPU: one process per GPU (torchrun), independent replicas, no data-path collective; the only
torch.distributed traffic is the barrier and the max-over-ranks of the elapsed time ("weak" scaling).
`--impl reference` times the oracle CPU path alone, on the full batch per step (the reference itself
performs no arithmetic: src/model/model.py:23-31 returns a constant dict).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# NCCL writes its banner / debug lines to stdout by default; stdout must carry exactly one JSON line.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

ARCH = "resnet50"
BATCH = 256
TOPK = 5
IMG = 224
# conv/linear/attention-matmul 2*MACs per 224^2 image
FLOPS = {"resnet50": 8.1784e9, "resnet18": 3.6281e9, "vit_b_16": 35.128e9}
FLOP_PER_IMAGE = FLOPS[ARCH]
METRIC = "predict images/sec (ResNet-50 224^2 bf16), batch 256 per B200 replica"
POOL_SLABS = 6                       # 6 x 38.5 MB of inputs > 126 MB L2


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def make_batch(seed: int):
    from mlmicroservicetemplate_b200 import synth
    return synth.synthetic_batch(seed, BATCH, IMG, IMG)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "25", "-i", str(device)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


def workload_config(world: int, sub_batch: int = 0) -> dict:
    """One description of the workload, printed by BOTH arms so that their `config` compares equal."""
    per_img = IMG * IMG * 3
    slab_bytes = BATCH * ((per_img + 255) // 256 * 256)
    return {"workload": f"{ARCH}_224_b{BATCH}_predict", "arch": ARCH, "batch_per_gpu": BATCH,
            "global_batch": BATCH * world, "image": "224x224x3 uint8 HWC", "topk": TOPK,
            "weights": "torchvision random init, seed 0, BN folded, bf16",
            "parallelism": f"replicas x{world}, no collective", "sub_batch": sub_batch,
            "l2": f"inputs rotate over a pool of {POOL_SLABS} slabs = {POOL_SLABS * slab_bytes / 1e6:.0f} MB (> 126 MB L2)"}


def cpu_reference_run(steps: int, warmup: int, budget_s: float, sample_batch: int):
    """Oracle CPU predict path (fp32, all host threads): `steps` passes over `sample_batch` images of the batch."""
    import torch
    from oracle import model as om
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = om.build_model(ARCH)
    ims = make_batch(1)[:sample_batch]
    for _ in range(max(warmup, 0)):
        om.predict_arrays(ims, mode="direct", k=TOPK, model=model)
    times = []
    t_start = time.perf_counter()
    for i in range(max(steps, 1)):
        t0 = time.perf_counter()
        om.predict_arrays(ims, mode="direct", k=TOPK, model=model)
        times.append(time.perf_counter() - t0)
        if budget_s and time.perf_counter() - t_start > budget_s and len(times) >= 2:
            break
    best = min(times)
    return {"value": sample_batch / statistics.median(times), "best": sample_batch / best, "unit": "images/s",
            "cores": cores, "kind": "port",
            "sample": f"{len(times)} x {sample_batch} images of the {BATCH}-image batch, torch CPU fp32 "
                      f"(oracle/model.py, {ARCH}), {cores} threads, median of {len(times)}",
            "ms_per_step": 1e3 * statistics.median(times), "steps": len(times)}


def host_decode_run(budget_s: float = 3.0):
    """The reference's own decoder (Pillow, src/model/model.py:23) on the JPEG files of the workload, all cores."""
    import io
    from concurrent.futures import ThreadPoolExecutor
    import numpy as np
    from PIL import Image
    files = []
    for im in make_batch(1):
        buf = io.BytesIO()
        Image.fromarray(im).save(buf, "JPEG", quality=90)
        files.append(buf.getvalue())
    cores = os.cpu_count() or 1

    def one(f):
        return np.asarray(Image.open(io.BytesIO(f)).convert("RGB"))
    with ThreadPoolExecutor(cores) as ex:
        list(ex.map(one, files[:32]))
        t0 = time.perf_counter()
        n = 0
        while time.perf_counter() - t0 < budget_s:
            list(ex.map(one, files))
            n += len(files)
        dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": "images/s", "cores": cores, "what": "Pillow decode of the 224x224 q90 JPEG files"}


def run_reference_arm(args):
    """The reference's CPU implementation of the path (its contract filled in by the oracle), all host threads,
    EXACTLY --steps K timed steps after --warmup W, each step the full batch of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    K = args.steps if args.steps is not None else 10
    W = args.warmup if args.warmup is not None else 2
    r = cpu_reference_run(K, W, budget_s=0.0, sample_batch=BATCH)
    hd = host_decode_run()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "images/s", "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": W, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(max(world, args.gpus), args.sub_batch),
        "note": "reference predict() has no arithmetic; this is its contract filled in on CPU (oracle), rank 0 only",
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_decode": hd,
        "e2e_from_jpeg": {"value": 1.0 / (1.0 / r["value"] + 1.0 / hd["value"]), "unit": "images/s",
                          "what": "Pillow decode then the CPU predict, both on all host cores, back to back"},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def k2_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum over the K2 launches of one step, from the committed ncu pass."""
    p = os.path.join(ROOT, "profiles", "k2_dram_traffic.json")
    try:
        d = json.load(open(p))
        e = d.get(f"{ARCH}_b{BATCH}")
        if e:
            return e["bytes"], e["source"]
    except (OSError, ValueError):
        pass
    return None, None


def engine_leg(n_gpus: int, prog, host_slabs, slab_bytes, geoms, files, budget_s: float):
    """The product's own multi-GPU path: ONE process, runtime.Engine over GPUs 0..n-1 (what init() builds), whole
    256-image batches handed to the least-loaded replica.  Host-clock throughput, per-GPU share of the batches."""
    import collections
    from mlmicroservicetemplate_b200.runtime import Engine, GpuReplicaWorker, JpegFile, PackedBatch
    workers = [GpuReplicaWorker(d, prog, BATCH, TOPK, "direct") for d in range(n_gpus)]
    eng = Engine(workers, max_batch=BATCH, max_wait_us=1000)
    out = {"gpus": n_gpus, "process": "one python process, one dispatcher thread + one worker thread per GPU"}

    def drive(make_item, label):
        depth = 3 * n_gpus
        pend = collections.deque()
        for i in range(2 * n_gpus):                      # warm-up: every replica sees a batch twice
            pend.append(eng.submit_batch(make_item(i)))
        while pend:
            pend.popleft().result()
        eng.dispatch_log.clear()
        done = 0
        i = 0
        checksum = 0.0
        t0 = time.perf_counter()
        while True:
            while len(pend) < depth:
                pend.append(eng.submit_batch(make_item(i)))
                i += 1
            probs, idx, errors = pend.popleft().result()
            assert not errors
            checksum += float(probs[0, 0])
            done += 1
            if time.perf_counter() - t0 > budget_s:
                break
        while pend:
            pend.popleft().result()
            done += 1
        dt = time.perf_counter() - t0
        share = collections.Counter(w for w, _ in eng.dispatch_log)
        out[label] = {"value": done * BATCH / dt, "unit": "images/s", "batches": done, "seconds": dt,
                      "batch_histogram": {str(BATCH): done},
                      "per_gpu_share": {str(g): share.get(g, 0) / max(1, sum(share.values())) for g in range(n_gpus)},
                      "checksum": checksum}

    packed = [PackedBatch(hs, slab_bytes, geoms) for hs in host_slabs]
    drive(lambda i: packed[i % len(packed)], "from_pinned_slabs")
    jbatches = [[JpegFile(f, IMG, IMG) for f in files[s:] + files[:s]] for s in range(3)]
    drive(lambda i: jbatches[i % len(jbatches)], "from_jpeg_files")
    eng.close()
    return out


def main():
    global ARCH, BATCH, FLOP_PER_IMAGE, METRIC
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps per round [100; reference arm 10]")
    ap.add_argument("--warmup", type=int, default=None, help="untimed steps [10, at least 3; reference arm 2]")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sub-batch", type=int, default=int(os.environ.get("B2P_SUB_BATCH", "0")))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-engine", action="store_true", help="skip the e2e_engine leg")
    ap.add_argument("--min-seconds", type=float, default=2.0, help="device time to accumulate over the rounds")
    ap.add_argument("--arch", default=ARCH, choices=sorted(FLOPS), help="non-default values give a secondary line")
    ap.add_argument("--batch", type=int, default=BATCH)
    args = ap.parse_args()
    if (args.arch, args.batch) != (ARCH, BATCH):
        ARCH, BATCH, FLOP_PER_IMAGE = args.arch, args.batch, FLOPS[args.arch]
        METRIC = f"predict images/sec ({ARCH} 224^2 bf16), batch {BATCH} per B200 replica"

    if args.impl == "reference":
        return run_reference_arm(args)
    K = args.steps if args.steps is not None else 100
    W = max(args.warmup if args.warmup is not None else 10, 3)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun the way the driver does
        os.execvp(sys.executable, [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                   f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
                                   "--master-port", "29511", os.path.abspath(__file__), *sys.argv[1:]])
    import hashlib
    import numpy as np
    import torch
    from mlmicroservicetemplate_b200 import backend as B, program as P
    from mlmicroservicetemplate_b200.ranks import Ranks
    from mlmicroservicetemplate_b200.runtime import seeded_model

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the predict path has no CPU fallback")
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    ranks = Ranks.from_env(backend="nccl", device=torch.device("cuda", local_rank))
    rank = ranks.rank

    model = seeded_model(ARCH, 0)
    prog = P.build_program(model, ARCH)
    rep = B.Replica(local_rank, prog, max_batch=BATCH, sub_batch=args.sub_batch)
    stream = torch.cuda.Stream()

    # ---- synthetic inputs: a pool of distinct uint8 slabs (> L2 in total), pinned on the host and on the device
    per_img = IMG * IMG * 3
    stride = (per_img + 255) // 256 * 256
    slab_bytes = BATCH * stride
    geoms = [(i * stride, IMG, IMG, IMG, IMG, 0, 0) for i in range(BATCH)]
    descs = B.make_images(geoms)
    host_slabs, dev_slabs = [], []

    def slab_of(images):
        a = np.zeros(slab_bytes, np.uint8)
        for i, im in enumerate(images):
            a[i * stride:i * stride + per_img] = im.reshape(-1)
        return a
    base = make_batch(1 + rank)
    base_np = slab_of(base)
    for s in range(POOL_SLABS):
        hs = torch.empty(slab_bytes, dtype=torch.uint8, pin_memory=True)
        hs.copy_(torch.from_numpy(np.roll(base_np.reshape(BATCH, stride), s, axis=0).reshape(-1)))
        host_slabs.append(hs)
        dev_slabs.append(hs.cuda(non_blocking=False))
    x = torch.empty(BATCH, 3, IMG, IMG, dtype=torch.bfloat16, device="cuda")
    logits = torch.empty(BATCH, prog.out_elems, dtype=torch.float32, device="cuda")
    probs = torch.empty(BATCH, TOPK, dtype=torch.float32, device="cuda")
    idx = torch.empty(BATCH, TOPK, dtype=torch.int32, device="cuda")
    probs_h = torch.empty(BATCH, TOPK, dtype=torch.float32, pin_memory=True)
    idx_h = torch.empty(BATCH, TOPK, dtype=torch.int32, pin_memory=True)

    barrier, max_over_ranks = ranks.barrier, ranks.max_over_ranks

    # ---------------------------------------------------------------- the same batch on every GPU
    common = torch.from_numpy(base_np if rank == 0 else slab_of(make_batch(1))).cuda()
    rep.preprocess(common, descs, BATCH, x, stream=stream)
    rep.forward(x, BATCH, logits, stream=stream)
    stream.synchronize()
    digest = hashlib.sha256(logits.cpu().numpy().tobytes()).digest()
    determinism = {"logits_sha256_equal_across_ranks": ranks.same_on_all_ranks(digest), "ranks": world,
                   "logits_sha256_rank0": digest.hex()[:16],
                   "what": "every rank ran the seed-1 batch of 256 images through K1+K2 on its own GPU"}
    del common

    def step_resident(i):
        # the chained device-resident entry point: K1 (crop-only images go straight into the stem's layout),
        # K2, K3 on one stream; the library records CUDA events around each stage (b2p_stats)
        rep.predict(dev_slabs[i % POOL_SLABS], descs, BATCH, TOPK, probs, idx, stream=stream)

    def timed_rounds(step, min_seconds, with_stage_times=False, max_rounds=60):
        """Rounds of exactly K steps, each bracketed by a barrier + synchronize; returns the per-round times
        (max over ranks, ms) and, optionally, the mean per-stage CUDA-event times of each round's calls."""
        rounds, stages = [], []
        total = 0.0
        it = 0
        while True:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if with_stage_times:
                rep.stats()                   # start a fresh averaging window
            barrier()
            e0.record(stream)
            for i in range(K):
                step(it)
                it += 1
            e1.record(stream)
            barrier()
            ms = max_over_ranks(e0.elapsed_time(e1))
            rounds.append(ms)
            if with_stage_times:
                st = rep.stats()
                stages.append((st["ms_preprocess"], st["ms_forward"], st["ms_epilogue"]))
            total += ms
            if (total >= min_seconds * 1e3 and len(rounds) >= 3) or len(rounds) >= max_rounds:
                return rounds, stages

    # ---------------------------------------------------------------- device-resident throughput (`value`)
    for i in range(W):
        step_resident(i)
    launches0 = rep.stats()["kernel_launches"]
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    res_rounds, stages = timed_rounds(step_resident, args.min_seconds, with_stage_times=True)
    clocks = sampler.stop() if sampler else None
    launches_per_round = (rep.stats()["kernel_launches"] - launches0) // len(res_rounds)
    total_ms = statistics.median(res_rounds)
    k1_ms = statistics.fmean(s[0] for s in stages)
    k2_ms = statistics.fmean(s[1] for s in stages)
    k3_ms = statistics.fmean(s[2] for s in stages)
    value = world * BATCH * K / (total_ms * 1e-3)

    # ---------------------------------------------------------------- end to end through host buffers (`e2e`)
    def step_host(i):
        rep.predict_host(host_slabs[i % POOL_SLABS], slab_bytes, descs, BATCH, TOPK, probs_h, idx_h, stream=stream)
    for i in range(W):
        step_host(i)
    e2e_rounds, _ = timed_rounds(step_host, args.min_seconds / 2)
    e2e_ms = statistics.median(e2e_rounds)
    e2e_value = world * BATCH * K / (e2e_ms * 1e-3)
    stage = rep.stats()
    ms_h2d = max_over_ranks(stage["ms_h2d"])
    checksum = float(probs_h.sum().item())          # the D2H result is really read on the host

    # ---------------------------------------------------------------- end to end from JPEG FILES (`e2e_from_jpeg`)
    # The step before the path: the batch arrives as JPEG files in host memory, is decoded on the
    # GPU (b2p_jpeg_decode: Huffman + IDCT + upsample/colour kernels) into the device slab, then K1-K3 run on it.
    import io as _io
    from PIL import Image as _Image
    files = []
    for im in base:
        buf = _io.BytesIO()
        _Image.fromarray(im).save(buf, "JPEG", quality=90)
        files.append(buf.getvalue())
    dec = B.JpegDecoder(local_rank)
    jslab = torch.empty(slab_bytes, dtype=torch.uint8, device="cuda")
    jstatus = torch.zeros(BATCH, dtype=torch.int32, device="cuda")
    joffs = [i * stride for i in range(BATCH)]
    rotated = [files[s:] + files[:s] for s in range(POOL_SLABS)]

    def step_jpeg(i):
        dec.decode(rotated[i % POOL_SLABS], jslab, slab_bytes, joffs, status=jstatus, stream=stream)
        rep.predict(jslab, descs, BATCH, TOPK, probs, idx, stream=stream)
        with torch.cuda.stream(stream):
            probs_h.copy_(probs, non_blocking=True)
            idx_h.copy_(idx, non_blocking=True)

    for i in range(W):
        step_jpeg(i)
    stream.synchronize()
    ref_px = np.asarray(_Image.open(_io.BytesIO(rotated[(W - 1) % POOL_SLABS][0])).convert("RGB"))
    jpeg_exact = bool(np.array_equal(jslab[:per_img].cpu().numpy().reshape(IMG, IMG, 3), ref_px)) and int(jstatus.abs().sum().item()) == 0
    jpeg_rounds, _ = timed_rounds(step_jpeg, args.min_seconds / 2)
    jpeg_ms = statistics.median(jpeg_rounds)
    jpeg_value = world * BATCH * K / (jpeg_ms * 1e-3)
    jst = dec.stats()
    jpeg_file_bytes = sum(len(f) for f in files)
    jpeg_checksum = float(probs_h.sum().item())
    dec.close()

    # per-batch latency (submit -> results on the host), synchronising every step: 200 batches
    lat = []
    for i in range(200):
        t0 = time.perf_counter()
        rep.predict_host(host_slabs[i % POOL_SLABS], slab_bytes, descs, BATCH, TOPK, probs_h, idx_h, stream=stream)
        stream.synchronize()
        lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()
    barrier()

    # ---------------------------------------------------------------- the product's dispatcher over all GPUs
    engine = None
    if not args.no_engine:
        if rank == 0:
            try:
                engine = engine_leg(world, prog, host_slabs, slab_bytes, geoms, files, budget_s=1.5)
            except Exception as e:                      # reported, never fatal for the headline line
                engine = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.set_device(local_rank)
        ranks.host_barrier()        # the other ranks wait here with their GPUs idle

    peaks = load_peaks()
    achieved_tflops = FLOP_PER_IMAGE * BATCH / (k2_ms * 1e-3) / 1e12
    traffic, traffic_src = k2_traffic()
    roofline = {
        "kernel": "K2 forward: every tcgen05 launch of one forward (conv_gemm / stem / attention kernels)",
        "bound": "tensor", "achieved": achieved_tflops, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
        "frac": achieved_tflops / peaks["bf16_tflops"],
        "frac_of_sustained_peak": achieved_tflops / peaks["bf16_tflops_sustained"],
        "peak_sustained": peaks["bf16_tflops_sustained"],
        "peak_source": peaks["source"] + " (burst figure, SURVEY 8(d); the sustained one is beside it)",
        "flop_per_launch_group": FLOP_PER_IMAGE * BATCH, "ms": k2_ms,
        "traffic": traffic, "traffic_source": traffic_src,
    }

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_run(steps=50, warmup=1, budget_s=15.0, sample_batch=min(32, BATCH))
        cpu_baseline = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        timed_s = (sum(res_rounds) + sum(e2e_rounds) + sum(jpeg_rounds)) * 1e-3
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": workload_config(world, args.sub_batch),
            "rounds": {"n": len(res_rounds), "steps_each": K, "ms_median": total_ms, "ms_min": min(res_rounds),
                       "ms_max": max(res_rounds), "timed_region_s_all_legs": timed_s,
                       "what": "value / ms_per_step are the MEDIAN round of K steps"},
            "stages_ms": {"k1_preprocess": k1_ms, "k2_forward": k2_ms, "k3_softmax_topk": k3_ms},
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": slab_bytes + BATCH * 32,
                    "d2h_bytes_per_step": BATCH * TOPK * 8, "ms_per_step": e2e_ms / K, "rounds": len(e2e_rounds),
                    "ms_h2d": ms_h2d, "from_jpeg_value": jpeg_value,
                    "stages_ms_last_call": {k: stage[k] for k in ("ms_h2d", "ms_preprocess", "ms_forward", "ms_epilogue", "ms_d2h")},
                    "api": "b2p_predict_host (pinned host slab in, pinned probs/idx out)", "checksum": checksum},
            "e2e_from_jpeg": {"value": jpeg_value, "unit": "images/s", "ms_per_step": jpeg_ms / K,
                              "h2d_bytes_per_step": jpeg_file_bytes, "d2h_bytes_per_step": BATCH * TOPK * 8,
                              "decode_ms_last_call": {"copy": jst.ms_copy, "huffman": jst.ms_huffman, "idct": jst.ms_idct,
                                                      "colour": jst.ms_colour},
                              "decoded_pixels_bit_exact_vs_pillow": jpeg_exact, "checksum": jpeg_checksum,
                              "api": "b2p_jpeg_decode (JPEG files in host memory -> device slab) + b2p_predict + D2H",
                              "files": f"{BATCH} x {IMG}x{IMG} quality 90 4:2:0, {jpeg_file_bytes / BATCH:.0f} B each"},
            "e2e_engine": engine,
            "determinism": determinism,
            "latency_ms": {"p50": lat[len(lat) // 2], "p99": lat[int(len(lat) * 0.99) - 1], "max": lat[-1], "n": len(lat),
                           "what": f"one {BATCH}-image batch, host slab -> host results, per GPU, synchronised every batch"},
            "gpu_launches": int(launches_per_round),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
        }
        print(json.dumps(line))
    rep.close()
    ranks.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
