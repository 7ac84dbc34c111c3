#!/usr/bin/env python
"""Synchronous latency of one b2p_predict_host call (pinned host slab in, top-5 on the host out) at small batch sizes:
what one micro-batch of the serving path costs on an otherwise idle GPU.

    python benchmarks/small_batch_latency.py [--arch resnet50] [--batches 1 4 8 16 32 64 128 256]"""
import argparse
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="resnet50")
    ap.add_argument("--batches", type=int, nargs="+", default=[1, 4, 8, 16, 32, 64, 128, 256])
    ap.add_argument("--calls", type=int, default=60)
    args = ap.parse_args()
    import numpy as np
    import torch
    from mlmicroservicetemplate_b200 import backend as B, program as P, synth
    from mlmicroservicetemplate_b200.runtime import seeded_model
    torch.cuda.set_device(0)
    prog = P.build_program(seeded_model(args.arch, 0), args.arch)
    nmax = max(args.batches)
    rep = B.Replica(0, prog, max_batch=nmax)
    stride = (224 * 224 * 3 + 255) // 256 * 256
    base = synth.synthetic_batch(1, min(nmax, 32), 224, 224)
    slab = torch.empty(nmax * stride, dtype=torch.uint8).pin_memory()
    sl = slab.numpy()
    for i in range(nmax):
        sl[i * stride:i * stride + 224 * 224 * 3] = base[i % len(base)].reshape(-1)
    probs = torch.empty(nmax, 5).pin_memory()
    idx = torch.empty(nmax, 5, dtype=torch.int32).pin_memory()
    stream = torch.cuda.Stream()
    out = {"arch": args.arch, "what": "b2p_predict_host, one call at a time, wall clock around call + stream sync", "batches": {}}
    for n in args.batches:
        descs = B.make_images([(i * stride, 224, 224, 224, 224, 0, 0) for i in range(n)])
        ts = []
        for c in range(args.calls + 10):
            t0 = time.perf_counter()
            rep.predict_host(slab, n * stride, descs, n, 5, probs, idx, stream=stream)
            stream.synchronize()
            if c >= 10:
                ts.append((time.perf_counter() - t0) * 1e3)
        st = rep.stats()
        out["batches"][str(n)] = {"p50_ms": round(statistics.median(ts), 4), "min_ms": round(min(ts), 4),
                                  "images_per_s_at_p50": round(n / statistics.median(ts) * 1e3),
                                  "k2_ms": round(st["ms_forward"], 4), "h2d_ms": round(st["ms_h2d"], 4)}
    rep.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
