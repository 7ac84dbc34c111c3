#!/usr/bin/env python
"""BASELINE.json configs[4]: a classifier (default ViT-B/16) behind the unchanged HTTP contract with
micro-batched concurrent /predict clients.

    python benchmarks/serve_concurrent.py [--arch vit_b_16] [--clients 64] [--requests 512]
                                          [--max-batch 64] [--waits-us 500,2000,8000] [--gpus 0,1,...]

Starts the product server in-process (FastAPI TestClient = the real ASGI app, plugin and engine; no
sockets), posts 224x224 JPEGs (image seeds 100+i) from `clients` threads and reports, per MAX_WAIT_US:
request p50/p99 latency, achieved batch-size histogram, requests/s, and where the time goes.  Host JPEG
decode stays on the CPU by design and bounds this number; the device-only rate for
the same model is bench.py --arch <arch> --batch <max-batch>.
"""
from __future__ import annotations

import argparse
import collections
import io
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="vit_b_16")
    ap.add_argument("--clients", type=int, default=64)
    ap.add_argument("--requests", type=int, default=512)
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--waits-us", default="500,2000,8000")
    ap.add_argument("--gpus", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    from PIL import Image
    from mlmicroservicetemplate_b200 import synth as images
    payloads = []
    for i in range(64):
        buf = io.BytesIO()
        Image.fromarray(images.synthetic_image(100 + i, 224, 224)).save(buf, format="JPEG", quality=90)
        payloads.append(buf.getvalue())

    results = []
    for wait_us in [int(w) for w in args.waits_us.split(",")]:
        os.environ.update(MODEL=args.arch, REGISTER="0", MAX_BATCH=str(args.max_batch), MAX_WAIT_US=str(wait_us),
                          TOPK="5", PREPROCESS="preset", SEED="0")
        if args.gpus:
            os.environ["GPUS"] = args.gpus
        import concurrent.futures as cf
        from fastapi.testclient import TestClient
        from mlmicroservicetemplate_b200.src.model import model as plugin
        from mlmicroservicetemplate_b200.src.server import dependency, main as srv
        plugin.shutdown()
        dependency.pool = cf.ThreadPoolExecutor(4)
        srv.dep.pool = dependency.pool
        srv._init_submitted = False
        dependency.model_settings.ready_to_predict = False
        with TestClient(srv.app, raise_server_exceptions=False) as client:
            t0 = time.time()
            while not dependency.model_settings.ready_to_predict and time.time() - t0 < 300:
                time.sleep(0.05)
            assert dependency.model_settings.ready_to_predict
            for i in range(8):     # warm-up
                client.post("/predict", files={"image_file": ("w.jpg", payloads[i], "image/jpeg")})
            engine = plugin._state["engine"]
            engine.batch_sizes.clear()
            lat = []
            lock = threading.Lock()
            counter = iter(range(args.requests))

            def worker():
                while True:
                    with lock:
                        i = next(counter, None)
                    if i is None:
                        return
                    t = time.perf_counter()
                    r = client.post("/predict", files={"image_file": (f"{i}.jpg", payloads[i % len(payloads)], "image/jpeg")})
                    dt = time.perf_counter() - t
                    assert r.status_code == 200, r.text
                    with lock:
                        lat.append(dt)

            threads = [threading.Thread(target=worker) for _ in range(args.clients)]
            t_start = time.perf_counter()
            [t.start() for t in threads]
            [t.join() for t in threads]
            wall = time.perf_counter() - t_start
            lat.sort()
            hist = collections.Counter(engine.batch_sizes)
            n_gpus = len(engine.workers)
            r = {"arch": args.arch, "gpus": n_gpus, "clients": args.clients, "requests": args.requests,
                 "max_batch": args.max_batch, "max_wait_us": wait_us, "requests_per_s": args.requests / wall,
                 "latency_ms": {"p50": 1e3 * lat[len(lat) // 2], "p99": 1e3 * lat[int(len(lat) * 0.99) - 1], "max": 1e3 * lat[-1]},
                 "batches": len(engine.batch_sizes), "mean_batch": sum(engine.batch_sizes) / max(1, len(engine.batch_sizes)),
                 "batch_histogram": dict(sorted(hist.items()))}
            results.append(r)
            print(json.dumps(r), flush=True)
        plugin.shutdown()
    if args.out:
        json.dump(results, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
