"""Decode throughput of the GPU JPEG decoder (b2p_jpeg_decode) next to the reference's decoder (Pillow) on the host cores.

    python benchmarks/jpeg_decode.py [--size 224] [--batch 256] [--quality 90] [--subsampling 2] [--restart-blocks 0]

GPU figure: files resident in host memory -> uint8 RGB HWC images resident in HBM, i.e. header parsing on the calling
thread + one host->device copy of the entropy-coded bytes + the three kernels, timed with CUDA events around the call
(max over --iters iterations is not used: the mean of the timed iterations is reported, after warm-up).
Host figure: `np.asarray(Image.open(f).convert('RGB'))` per file, on one core and on a thread pool over all cores
(Pillow releases the GIL while decoding).  Prints one JSON line.
"""
import argparse
import io
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def picture(h, w, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    a = np.stack([127 + 100 * np.sin(xx / 17.0 + yy / 31.0 + seed), 127 + 90 * np.cos(xx / 9.0 - yy / 13.0),
                  127 + 120 * np.sin(xx * yy / 3000.0 + seed)], -1) + rng.normal(0, 6, (h, w, 3))
    return np.clip(a, 0, 255).astype(np.uint8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, nargs="+", default=[224])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--quality", type=int, default=90)
    ap.add_argument("--subsampling", type=int, default=2)
    ap.add_argument("--restart-blocks", type=int, default=0)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--distinct", type=int, default=32)
    ap.add_argument("--no-pillow", action="store_true", help="skip the host-decoder timing (kernel sweeps)")
    args = ap.parse_args()
    import torch
    from PIL import Image, ImageFile
    from mlmicroservicetemplate_b200 import backend as B
    ImageFile.MAXBLOCK = 1 << 26
    h = args.size[0]
    w = args.size[1] if len(args.size) > 1 else h
    kw = dict(quality=args.quality, subsampling=args.subsampling)
    if args.restart_blocks:
        kw["restart_marker_blocks"] = args.restart_blocks
    distinct = []
    for s in range(min(args.distinct, args.batch)):
        b = io.BytesIO()
        Image.fromarray(picture(h, w, s)).save(b, "JPEG", **kw)
        distinct.append(b.getvalue())
    files = [distinct[i % len(distinct)] for i in range(args.batch)]
    file_bytes = sum(len(f) for f in files)

    # ---- host: Pillow
    def pil(f):
        return np.asarray(Image.open(io.BytesIO(f)).convert("RGB"))
    t0 = time.perf_counter()
    n1 = 0
    budget = 0.05 if args.no_pillow else 2.0
    while time.perf_counter() - t0 < budget:
        pil(files[n1 % len(files)])
        n1 += 1
    one_core = n1 / (time.perf_counter() - t0)
    cores = os.cpu_count() or 1
    with ThreadPoolExecutor(cores) as ex:
        list(ex.map(pil, files))
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 1.5 * budget:
            list(ex.map(pil, files))
            reps += 1
        all_cores = reps * len(files) / (time.perf_counter() - t0)

    # ---- GPU
    torch.cuda.set_device(0)
    dec = B.JpegDecoder(0)
    per = (h * w * 3 + 255) // 256 * 256
    offsets = [i * per for i in range(args.batch)]
    slab = torch.zeros(per * args.batch, dtype=torch.uint8, device="cuda")
    status = torch.zeros(args.batch, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(args.warmup):
        dec.decode(files, slab, slab.numel(), offsets, status=status)
    torch.cuda.synchronize()
    ref = pil(files[0])
    got = slab[:h * w * 3].cpu().numpy().reshape(h, w, 3)
    exact = bool(np.array_equal(ref, got)) and int(status.abs().sum().item()) == 0
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.iters)]
    host_ms = []
    wall0 = time.perf_counter()
    for a, b in ev:
        flush.zero_()                                   # evict the previous iteration's data from L2
        a.record()
        t0 = time.perf_counter()
        dec.decode(files, slab, slab.numel(), offsets, status=status)
        host_ms.append((time.perf_counter() - t0) * 1e3)
        b.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - wall0
    gpu_ms = [a.elapsed_time(b) for a, b in ev]
    ms = float(np.mean(gpu_ms))
    st = dec.stats()
    out = {
        "metric": "JPEG decode images/sec (host files -> RGB uint8 HWC in HBM)",
        "workload": f"{args.batch} x {h}x{w} q{args.quality} subsampling {args.subsampling} restart_blocks {args.restart_blocks}",
        "gpu_images_per_s": args.batch / (ms / 1e3), "gpu_ms_per_batch": ms, "gpu_ms_min": float(np.min(gpu_ms)),
        "host_call_ms": float(np.mean(host_ms)),
        "kernel_ms": {"copy": st.ms_copy, "huffman": st.ms_huffman, "idct": st.ms_idct, "colour": st.ms_colour}, "wall_images_per_s": args.batch * args.iters / wall,
        "compressed_MB_per_s": file_bytes / (ms / 1e3) / 1e6, "pixels_GB_per_s": args.batch * h * w * 3 / (ms / 1e3) / 1e9,
        "bit_exact_vs_pillow": exact, "file_bytes_per_image": file_bytes / args.batch,
        "pillow_images_per_s_1core": one_core, "pillow_images_per_s_all_cores": all_cores, "cores": cores,
        "iters": args.iters, "warmup": args.warmup,
    }
    print(json.dumps(out))


if __name__ == "__main__":
    main()
