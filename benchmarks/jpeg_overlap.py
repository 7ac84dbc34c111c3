"""JPEG files -> top-k with the decode of batch i+1 running on a second stream while batch i goes through K1-K3.

    python benchmarks/jpeg_overlap.py [--arch resnet50] [--batch 256] [--steps 60]

Two device slabs alternate; events order "decode(i) -> predict(i)" and "predict(i) has read slab -> decode(i+2) may
overwrite it".  Prints one JSON line with the serial (one stream) and the pipelined (two streams) images/s.
"""
import argparse
import io
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="resnet50")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=6)
    ap.add_argument("--size", type=int, nargs=2, default=[224, 224], help="H W of the uploads (e.g. 1080 1920: photos)")
    ap.add_argument("--distinct", type=int, default=64)
    args = ap.parse_args()
    import numpy as np
    import torch
    from PIL import Image
    from mlmicroservicetemplate_b200 import backend as B, program as P, synth
    from mlmicroservicetemplate_b200.runtime import image_geometry, seeded_model
    torch.cuda.set_device(0)
    n, k = args.batch, 5
    H, W = args.size
    files = []
    for im in synth.synthetic_batch(1, min(n, args.distinct), H, W):
        buf = io.BytesIO()
        Image.fromarray(im).save(buf, "JPEG", quality=90)
        files.append(buf.getvalue())
    files = [files[i % len(files)] for i in range(n)]
    rot = [files[s:] + files[:s] for s in range(4)]
    stride = (H * W * 3 + 255) // 256 * 256
    offs = [i * stride for i in range(n)]
    geom = image_geometry(H, W, "direct" if (H, W) == (224, 224) else "preset")
    descs = B.make_images([(o, H, W, *geom) for o in offs])
    rep = B.Replica(0, P.build_program(seeded_model(args.arch, 0), args.arch), max_batch=n)
    dec = B.JpegDecoder(0)
    slabs = [torch.empty(n * stride, dtype=torch.uint8, device="cuda") for _ in range(2)]
    status = torch.zeros(n, dtype=torch.int32, device="cuda")
    probs = torch.empty(n, k, device="cuda")
    idx = torch.empty(n, k, dtype=torch.int32, device="cuda")
    probs_h = torch.empty(n, k, pin_memory=True)
    idx_h = torch.empty(n, k, dtype=torch.int32, pin_memory=True)
    s_dec, s_run = torch.cuda.Stream(), torch.cuda.Stream()

    def serial(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s_run)
        for i in range(steps):
            dec.decode(rot[i % 4], slabs[0], slabs[0].numel(), offs, status=status, stream=s_run)
            rep.predict(slabs[0], descs, n, k, probs, idx, stream=s_run)
            with torch.cuda.stream(s_run):
                probs_h.copy_(probs, non_blocking=True)
                idx_h.copy_(idx, non_blocking=True)
        e1.record(s_run)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    def pipelined(steps):
        decoded = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s_run)
        s_dec.wait_stream(s_run)
        for i in range(steps + 1):
            b = i & 1
            if i < steps:
                if i >= 2:
                    s_dec.wait_event(consumed[b])          # predict(i-2) has finished reading this slab
                dec.decode(rot[i % 4], slabs[b], slabs[b].numel(), offs, status=status, stream=s_dec)
                decoded[b].record(s_dec)
            if i >= 1:
                pb = (i - 1) & 1
                s_run.wait_event(decoded[pb])
                rep.predict(slabs[pb], descs, n, k, probs, idx, stream=s_run)
                consumed[pb].record(s_run)
                with torch.cuda.stream(s_run):
                    probs_h.copy_(probs, non_blocking=True)
                    idx_h.copy_(idx, non_blocking=True)
        e1.record(s_run)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    serial(args.warmup)
    ms_serial = serial(args.steps)
    ref = probs_h.clone()
    pipelined(args.warmup)
    ms_pipe = pipelined(args.steps)
    same = bool(torch.equal(ref, probs_h)) if (args.steps % 4) == 0 else None
    print(json.dumps({"arch": args.arch, "batch": n, "upload": f"{H}x{W} JPEG q90 4:2:0", "serial_ms": ms_serial, "serial_images_per_s": n / ms_serial * 1e3,
                      "pipelined_ms": ms_pipe, "pipelined_images_per_s": n / ms_pipe * 1e3,
                      "same_result_as_serial": same, "file_bytes": sum(len(f) for f in files) // n}))
    dec.close()
    rep.close()


if __name__ == "__main__":
    main()
