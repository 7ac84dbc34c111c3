#!/usr/bin/env python
"""One replica, a few chained predict steps and nothing else: the command the ncu launch lists under profiles/ are
taken from (the per-kernel view of one step).

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python benchmarks/step_profile.py [--arch resnet50] [--batch 256] [--steps 3]

Without ncu it prints the mean CUDA-event time per step (a number printed UNDER ncu is not a benchmark value)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="resnet50")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import numpy as np
    import torch
    from mlmicroservicetemplate_b200 import backend as B, program as P, synth
    from mlmicroservicetemplate_b200.runtime import seeded_model
    torch.cuda.set_device(0)
    prog = P.build_program(seeded_model(args.arch, 0), args.arch)
    rep = B.Replica(0, prog, max_batch=args.batch)
    n = args.batch
    stride = (224 * 224 * 3 + 255) // 256 * 256
    base = synth.synthetic_batch(1, min(n, 32), 224, 224)
    slab = np.zeros(n * stride, np.uint8)
    for i in range(n):
        slab[i * stride:i * stride + 224 * 224 * 3] = base[i % len(base)].reshape(-1)
    slab_d = torch.from_numpy(slab).cuda()
    descs = B.make_images([(i * stride, 224, 224, 224, 224, 0, 0) for i in range(n)])
    probs = torch.empty(n, 5, device="cuda")
    idx = torch.empty(n, 5, dtype=torch.int32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rep.predict(slab_d, descs, n, 5, probs, idx)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.steps):
        rep.predict(slab_d, descs, n, 5, probs, idx)
    e1.record()
    torch.cuda.synchronize()
    print(json.dumps({"arch": args.arch, "batch": n, "ms_per_step": e0.elapsed_time(e1) / args.steps, "stats": rep.stats()}))
    rep.close()


if __name__ == "__main__":
    main()
