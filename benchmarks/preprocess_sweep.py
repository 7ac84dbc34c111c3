#!/usr/bin/env python
"""BASELINE.json configs[3]: preprocess-only sweep, synthetic 320-1920 px images -> 224^2 normalised
bf16, achieved HBM GB/s vs the measured peak.

    python benchmarks/preprocess_sweep.py [--batch 64] [--iters 20] [--out profiles/r01_k1_sweep.json]

Per size and mode (direct = resize straight to 224^2; preset = resize-256 + centre-crop-224):
  bytes  = algorithmic bytes of SURVEY §8(d): n * (3*H*W [window for preset] + 301,056)
  GB/s   = bytes / CUDA-event time of b2p_preprocess on device-resident uint8 HWC input
  parity = max |fp32 output - torch CPU| on two images of the batch (gate 1e-6) and bf16 ulps (gate 1)
Inputs rotate over several slabs so that consecutive iterations do not re-read L2-resident data.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SIZES = [(320, 320), (480, 480), (640, 640), (800, 800), (1024, 1024), (1280, 1280), (1600, 1600), (1920, 1920),
         (480, 640), (720, 1280), (1080, 1920)]


def _tap_window(in_size, out_size, i):
    """(first source index, count) of the antialias triangle filter for output index i."""
    scale = in_size / out_size
    support = scale if scale >= 1 else 1.0
    center = scale * (i + 0.5)
    lo = max(int(center - support + 0.5), 0)
    return lo, min(int(center + support + 0.5), in_size) - lo


def preset_window_bytes(h, w):
    """Source bytes that feed the 224^2 crop in preset mode (window extent on each axis)."""
    from mlmicroservicetemplate_b200.runtime import preset_geometry
    rh, rw, top, left = preset_geometry(h, w)
    y0 = _tap_window(h, rh, top)[0]
    y1 = sum(_tap_window(h, rh, top + 223))
    x0 = _tap_window(w, rw, left)[0]
    x1 = sum(_tap_window(w, rw, left + 223))
    return 3 * (y1 - y0) * (x1 - x0)


def torch_cpu_preprocess(img, res_h, res_w, top, left):
    """What torch's CPU path gives for stage P (fp32, antialiased bilinear, no re-quantisation): the parity figure
    this sweep prints next to each timing.  The gating parity tests live in tests/test_gpu_preprocess.py."""
    import numpy as np
    import torch
    import torch.nn.functional as F
    x = torch.from_numpy(np.ascontiguousarray(img)).permute(2, 0, 1)[None].float()
    if (x.shape[2], x.shape[3]) != (res_h, res_w):
        x = F.interpolate(x, size=(res_h, res_w), mode="bilinear", align_corners=False, antialias=True)
    x = x[:, :, top:top + 224, left:left + 224]
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    return (x / 255.0 - mean) / std


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="")
    ap.add_argument("--sizes", default="", help="e.g. 320x320,1080x1920 (default: the full config-4 list)")
    ap.add_argument("--modes", default="direct,preset")
    args = ap.parse_args()
    global SIZES
    if args.sizes:
        SIZES = [tuple(int(v) for v in t.split("x")) for t in args.sizes.split(",")]

    import numpy as np
    import torch
    from mlmicroservicetemplate_b200 import backend as B, program as P
    from mlmicroservicetemplate_b200 import synth as images
    from mlmicroservicetemplate_b200.runtime import image_geometry

    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
    peak = peaks["hbm_gbs"]
    n = args.batch
    p = P.Program(name="k1_only")
    p.in_elems = 8; p.out_elems = 8
    p.ops.append(P._mk_op(B.OP_AVGPOOL, B.SLOT_INPUT, B.SLOT_OUTPUT, in_hwc=(1, 1, 8), out_hwc=(1, 1, 8)))
    rep = B.Replica(0, p, max_batch=n)
    out16 = torch.empty(n, 3, 224, 224, dtype=torch.bfloat16, device="cuda")
    out32 = torch.empty(2, 3, 224, 224, dtype=torch.float32, device="cuda")
    results = []
    for (h, w) in SIZES:
        per = h * w * 3
        stride = (per + 255) // 256 * 256
        base = [images.synthetic_image(2 * 100003 + i, h, w) for i in range(4)]      # image seed 2
        slab_np = np.zeros(n * stride, np.uint8)
        for i in range(n):
            slab_np[i * stride:i * stride + per] = np.roll(base[i % 4].reshape(-1), i * 3)   # distinct content per slot
        for i in range(2):
            slab_np[i * stride:i * stride + per] = base[i].reshape(-1)
        n_slabs = max(2, int(300e6 // (n * stride)) + 1)          # > 126 MB L2 in total
        slabs = [torch.from_numpy(slab_np).cuda() for _ in range(n_slabs)]
        for mode in args.modes.split(","):
            geom = image_geometry(h, w, mode)
            descs = B.make_images([(i * stride, h, w, *geom) for i in range(n)])
            # parity on the first two images
            rep.preprocess(slabs[0], B.make_images([(i * stride, h, w, *geom) for i in range(2)]), 2, out32, out_f32=True)
            rep.preprocess(slabs[0], descs, n, out16)
            torch.cuda.synchronize()
            ref = torch.cat([torch_cpu_preprocess(base[i], *geom) for i in range(2)], 0)
            err = (out32.cpu() - ref).abs().max().item()
            def key(t):
                i16 = t.view(torch.int16).to(torch.int32)
                return torch.where(i16 < 0, -(i16 & 0x7FFF), i16)
            ulps = (key(out16[:2].cpu()) - key(ref.to(torch.bfloat16))).abs().max().item()
            for i in range(3):
                rep.preprocess(slabs[i % n_slabs], descs, n, out16)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(args.iters):
                rep.preprocess(slabs[i % n_slabs], descs, n, out16)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.iters
            in_bytes = 3 * h * w if mode == "direct" else preset_window_bytes(h, w)
            total = n * (in_bytes + 301056)
            gbs = total / (ms * 1e-3) / 1e9
            r = {"h": h, "w": w, "mode": mode, "batch": n, "ms": ms, "images_per_s": n / (ms * 1e-3),
                 "algorithmic_bytes": total, "gb_per_s": gbs, "frac_of_measured_hbm_peak": gbs / peak,
                 "fp32_max_err": err, "bf16_ulps": int(ulps)}
            results.append(r)
            print(json.dumps(r), flush=True)
            assert err <= 1e-6 and ulps <= 1, (h, w, mode, err, ulps)
        del slabs
        torch.cuda.empty_cache()
    rep.close()
    if args.out:
        json.dump({"peak_hbm_gbs": peak, "results": results}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
