#!/usr/bin/env python
"""Headline benchmark: predict images/sec, ResNet-50 224^2 bf16, batch 256 per B200 replica
(BASELINE.json `metric`, configs[2]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path (K1 preprocess -> K2 forward -> K3 softmax/top-k) over one batch
of 256 synthetic decoded uint8 HWC images per GPU.
  value  : whole-job images/s with the uint8 inputs already resident in HBM, CUDA-event timed.
  e2e    : the same metric through the host-buffer C-ABI call b2p_predict_host — pinned host slab in,
           pinned (prob, index) out, H2D and D2H inside the timed region.
  roofline     : K2 (the tcgen05 conv/linear launches), algorithmic FLOPs / CUDA-event time vs the
                 measured bf16 peak in MEASURED_PEAKS.json.
  cpu_baseline : the oracle (reference predict contract filled in with torch CPU fp32) timed on the
                 host cores on a bounded sample of the same workload (rank 0, N=1 only).
Multi-GPU: one process per GPU (torchrun), independent replicas, no data-path collective; the only
torch.distributed traffic is the barrier and the max-over-ranks of the elapsed time ("weak" scaling).
`--impl reference` times the oracle CPU path alone (the reference itself performs no arithmetic:
src/model/model.py:23-31 returns a constant dict).
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
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(device)],
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


def cpu_reference_run(steps: int, warmup: int, budget_s: float, sample_batch: int):
    """Oracle CPU predict path (fp32, all host threads) on a bounded sample of the workload."""
    import torch
    from oracle import model as om
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = om.build_model(ARCH)
    ims = make_batch(1)[:sample_batch]
    for _ in range(max(1, min(warmup, 1))):
        om.predict_arrays(ims[: max(4, sample_batch // 8)], mode="direct", k=TOPK, model=model)
    times = []
    t_start = time.perf_counter()
    for i in range(max(steps, 1)):
        t0 = time.perf_counter()
        om.predict_arrays(ims, mode="direct", k=TOPK, model=model)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_start > budget_s and len(times) >= 2:
            break
    best = min(times)
    return {"value": sample_batch / statistics.median(times), "best": sample_batch / best, "unit": "images/s",
            "cores": cores, "kind": "port",
            "sample": f"{len(times)} x {sample_batch} images of the 256-image batch, torch CPU fp32 "
                      f"(oracle/model.py), {cores} threads, median of {len(times)}",
            "ms_per_step": 1e3 * statistics.median(times), "steps": len(times)}


def host_decode_run(budget_s: float = 3.0):
    """The reference's own decoder (Pillow, src/model/model.py:23) on the 256 JPEG files of the workload, all cores."""
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
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    r = cpu_reference_run(args.steps, args.warmup, budget_s=120.0, sample_batch=32)
    hd = host_decode_run()
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "images/s", "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": 1, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "resnet50_224_b256_predict", "batch_per_gpu": BATCH, "image": "224x224x3 uint8",
                   "note": "reference predict() has no arithmetic; this is its contract filled in on CPU (oracle)"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host_decode": hd,
        "e2e_from_jpeg": {"value": 1.0 / (1.0 / r["value"] + 1.0 / hd["value"]), "unit": "images/s",
                          "what": "Pillow decode then the CPU predict, both on all host cores, back to back"},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def main():
    global ARCH, BATCH, FLOP_PER_IMAGE, METRIC
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sub-batch", type=int, default=int(os.environ.get("B2P_SUB_BATCH", "0")))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--arch", default=ARCH, choices=sorted(FLOPS), help="non-default values give a secondary line")
    ap.add_argument("--batch", type=int, default=BATCH)
    args = ap.parse_args()
    if (args.arch, args.batch) != (ARCH, BATCH):
        ARCH, BATCH, FLOP_PER_IMAGE = args.arch, args.batch, FLOPS[args.arch]
        METRIC = f"predict images/sec ({ARCH} 224^2 bf16), batch {BATCH} per B200 replica"
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        return run_reference_arm(args)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun the way the driver does
        os.execvp(sys.executable, [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                   f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
                                   "--master-port", "29511", os.path.abspath(__file__), *sys.argv[1:]])
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
    base = make_batch(1 + rank)
    base_np = np.zeros(slab_bytes, np.uint8)
    for i, im in enumerate(base):
        base_np[i * stride:i * stride + per_img] = im.reshape(-1)
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

    K, W = args.steps, args.warmup

    def step_resident(i, evs=None):
        slab = dev_slabs[i % POOL_SLABS]
        if evs: evs[0].record(stream)
        rep.preprocess(slab, descs, BATCH, x, stream=stream)
        if evs: evs[1].record(stream)
        rep.forward(x, BATCH, logits, stream=stream)
        if evs: evs[2].record(stream)
        rep.softmax_topk(logits, BATCH, prog.out_elems, TOPK, probs, idx, stream=stream)
        if evs: evs[3].record(stream)

    # ---------------------------------------------------------------- device-resident throughput (`value`)
    for i in range(W):
        step_resident(i)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
    launches0 = rep.stats()["kernel_launches"]
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    for i in range(K):
        step_resident(i, evs[i])
    barrier()
    clocks = sampler.stop() if sampler else None
    launches_per_run = rep.stats()["kernel_launches"] - launches0
    total_ms = max_over_ranks(evs[0][0].elapsed_time(evs[-1][3]))
    k1_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / K
    k2_ms = sum(e[1].elapsed_time(e[2]) for e in evs) / K
    k3_ms = sum(e[2].elapsed_time(e[3]) for e in evs) / K
    value = world * BATCH * K / (total_ms * 1e-3)

    # ---------------------------------------------------------------- end to end through host buffers (`e2e`)
    for i in range(W):
        rep.predict_host(host_slabs[i % POOL_SLABS], slab_bytes, descs, BATCH, TOPK, probs_h, idx_h, stream=stream)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(K):
        rep.predict_host(host_slabs[i % POOL_SLABS], slab_bytes, descs, BATCH, TOPK, probs_h, idx_h, stream=stream)
    e1.record(stream)
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = world * BATCH * K / (e2e_ms * 1e-3)
    stage = rep.stats()
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
    barrier()
    j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    j0.record(stream)
    for i in range(K):
        step_jpeg(i)
    j1.record(stream)
    barrier()
    jpeg_ms = max_over_ranks(j0.elapsed_time(j1))
    jpeg_value = world * BATCH * K / (jpeg_ms * 1e-3)
    jst = dec.stats()
    jpeg_file_bytes = sum(len(f) for f in files)
    jpeg_checksum = float(probs_h.sum().item())
    dec.close()

    # per-batch latency (submit -> results on the host), synchronising every step
    lat = []
    for i in range(min(K, 30)):
        t0 = time.perf_counter()
        rep.predict_host(host_slabs[i % POOL_SLABS], slab_bytes, descs, BATCH, TOPK, probs_h, idx_h, stream=stream)
        stream.synchronize()
        lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()

    peaks = load_peaks()
    achieved_tflops = FLOP_PER_IMAGE * BATCH / (k2_ms * 1e-3) / 1e12
    roofline = {
        "kernel": "K2 forward: every tcgen05 launch of one forward (conv_gemm / stem / attention kernels)",
        "bound": "tensor", "achieved": achieved_tflops, "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
        "frac": achieved_tflops / peaks["bf16_tflops_sustained"],
        "frac_of_burst_peak": achieved_tflops / peaks["bf16_tflops"], "peak_burst": peaks["bf16_tflops"],
        "peak_source": peaks["source"] + " (sustained: kernel timed inside a long step)",
        "flop_per_launch_group": FLOP_PER_IMAGE * BATCH, "ms": k2_ms,
        # dram__bytes_read.sum + dram__bytes_write.sum over the K2 launches of one 256-image ResNet-50 step, from the
        # ncu pass committed as profiles/r01j_launches_dram_raw_ncu.csv (bytes per launch group, like `achieved`)
        "traffic": 8.42e9 if (ARCH, BATCH) == ("resnet50", 256) else None,
        "traffic_source": "profiles/r01j_launches_dram_raw_ncu.csv",
    }

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_run(steps=50, warmup=1, budget_s=15.0, sample_batch=32)
        cpu_baseline = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"{ARCH}_224_b{BATCH}_predict", "arch": ARCH, "batch_per_gpu": BATCH,
                       "global_batch": BATCH * world, "image": "224x224x3 uint8 HWC", "topk": TOPK,
                       "weights": "torchvision random init, seed 0, BN folded, bf16",
                       "parallelism": f"replicas x{world}, no collective", "sub_batch": args.sub_batch,
                       "l2": f"inputs rotate over a pool of {POOL_SLABS} slabs = {POOL_SLABS * slab_bytes / 1e6:.0f} MB (> 126 MB L2)"},
            "stages_ms": {"k1_preprocess": k1_ms, "k2_forward": k2_ms, "k3_softmax_topk": k3_ms},
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": slab_bytes + BATCH * 32,
                    "d2h_bytes_per_step": BATCH * TOPK * 8, "ms_per_step": e2e_ms / K,
                    "stages_ms_last_call": {k: stage[k] for k in ("ms_h2d", "ms_preprocess", "ms_forward", "ms_epilogue", "ms_d2h")},
                    "api": "b2p_predict_host (pinned host slab in, pinned probs/idx out)", "checksum": checksum},
            "e2e_from_jpeg": {"value": jpeg_value, "unit": "images/s", "ms_per_step": jpeg_ms / K,
                              "h2d_bytes_per_step": jpeg_file_bytes, "d2h_bytes_per_step": BATCH * TOPK * 8,
                              "decode_ms_last_call": {"copy": jst.ms_copy, "huffman": jst.ms_huffman, "idct": jst.ms_idct,
                                                      "colour": jst.ms_colour},
                              "decoded_pixels_bit_exact_vs_pillow": jpeg_exact, "checksum": jpeg_checksum,
                              "api": "b2p_jpeg_decode (JPEG files in host memory -> device slab) + b2p_predict + D2H",
                              "files": f"{BATCH} x {IMG}x{IMG} quality 90 4:2:0, {jpeg_file_bytes / BATCH:.0f} B each"},
            "latency_ms": {"p50": lat[len(lat) // 2], "p99": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "n": len(lat),
                           "what": "one 256-image batch, host slab -> host results, per GPU"},
            "gpu_launches": int(launches_per_run),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
        }
        print(json.dumps(line))
    rep.close()
    ranks.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
