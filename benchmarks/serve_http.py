#!/usr/bin/env python
"""SURVEY §8(f) rank 2 / BASELINE.json configs[4] over REAL sockets: the multi-process deployment
(src/server/serve.py: N uvicorn front-end processes + one GPU dispatcher process, uploads handed over through shared
memory) under concurrent /predict clients.

    python benchmarks/serve_http.py --arch resnet50 --workers 1,2,4,8 --clients 64 --seconds 6 [--gpus 0,1,...]

Per worker count: requests/s, request p50/p99, and the dispatcher's view of GPU use.  Clients are separate processes
(keep-alive http.client connections, pre-encoded multipart bodies of 224x224 q90 JPEGs, image seeds 100+i), so their
cost does not sit in the servers' interpreters; they do share the box's host cores with them.
"""
from __future__ import annotations

import argparse
import http.client
import io
import json
import multiprocessing as mp
import os
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def multipart(jpeg: bytes, name: str) -> tuple:
    boundary = "b2pboundary7d1f"
    body = (f"--{boundary}\r\nContent-Disposition: form-data; name=\"image_file\"; filename=\"{name}\"\r\n"
            f"Content-Type: image/jpeg\r\n\r\n").encode() + jpeg + f"\r\n--{boundary}--\r\n".encode()
    return body, f"multipart/form-data; boundary={boundary}"


def client_proc(port: int, threads: int, seconds: float, bodies, ctype: str, out_q) -> None:
    import threading
    lat, errors = [], [0]
    lock = threading.Lock()
    t_end = time.perf_counter() + seconds

    def run(tid):
        conn = http.client.HTTPConnection("127.0.0.1", port, timeout=60)
        mine = []
        i = tid
        while time.perf_counter() < t_end:
            body = bodies[i % len(bodies)]
            i += 1
            t0 = time.perf_counter()
            try:
                conn.request("POST", "/predict", body=body, headers={"Content-Type": ctype})
                r = conn.getresponse()
                data = r.read()
                if r.status != 200 or b'"success"' not in data:
                    errors[0] += 1
            except (OSError, http.client.HTTPException):
                errors[0] += 1
                conn.close()
                conn = http.client.HTTPConnection("127.0.0.1", port, timeout=60)
                continue
            mine.append(time.perf_counter() - t0)
        with lock:
            lat.extend(mine)
        conn.close()

    ts = [threading.Thread(target=run, args=(t,)) for t in range(threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    out_q.put((lat, errors[0]))


def free_port() -> int:
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def wait_ready(port: int, proc, timeout_s: float = 600.0) -> None:
    t0 = time.time()
    while time.time() - t0 < timeout_s:
        if proc.poll() is not None:
            raise RuntimeError(f"server exited with code {proc.returncode}")
        try:
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=5)
            c.request("GET", "/status")
            r = c.getresponse(); r.read(); c.close()
            if r.status == 200:
                return
        except OSError:
            pass
        time.sleep(0.25)
    raise RuntimeError("server did not become ready")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="resnet50")
    ap.add_argument("--workers", default="1,2,4,8")
    ap.add_argument("--clients", type=int, default=64, help="concurrent connections in total")
    ap.add_argument("--client-procs", type=int, default=4)
    ap.add_argument("--seconds", type=float, default=6.0)
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--max-wait-us", type=int, default=2000)
    ap.add_argument("--gpus", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    from PIL import Image
    from mlmicroservicetemplate_b200 import synth
    bodies, ctype = [], None
    for i in range(64):
        buf = io.BytesIO()
        Image.fromarray(synth.synthetic_image(100 + i, 224, 224)).save(buf, format="JPEG", quality=90)
        body, ctype = multipart(buf.getvalue(), f"{i}.jpg")
        bodies.append(body)

    results = []
    for nw in [int(v) for v in args.workers.split(",")]:
        port = free_port()
        env = dict(os.environ, MODEL=args.arch, REGISTER="0", MAX_BATCH=str(args.max_batch),
                   MAX_WAIT_US=str(args.max_wait_us), TOPK="5", PREPROCESS="preset", SEED="0", DECODE="gpu",
                   PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
        if args.gpus:
            env["GPUS"] = args.gpus
        stats_path = os.path.join(tempfile.gettempdir(), f"b2p-dispatch-stats-{os.getpid()}-{nw}.json")
        env["B2P_DISPATCH_STATS"] = stats_path
        srv = subprocess.Popen([sys.executable, "-m", "mlmicroservicetemplate_b200.src.server.serve", "--workers", str(nw),
                                "--host", "127.0.0.1", "--port", str(port)], env=env, cwd=ROOT,
                               start_new_session=True, stdout=subprocess.DEVNULL)
        try:
            wait_ready(port, srv)
            # every front-end worker initialises lazily on its first requests: warm them all
            q = mp.Queue()
            warm = mp.Process(target=client_proc, args=(port, 4 * nw, 1.5, bodies, ctype, q)); warm.start(); q.get(); warm.join()
            procs, per = [], max(1, args.clients // args.client_procs)
            t0 = time.perf_counter()
            for _ in range(args.client_procs):
                p = mp.Process(target=client_proc, args=(port, per, args.seconds, bodies, ctype, q))
                p.start(); procs.append(p)
            lat, errs = [], 0
            for _ in procs:
                l, e = q.get()
                lat += l; errs += e
            [p.join() for p in procs]
            wall = time.perf_counter() - t0
            lat.sort()
            r = {"arch": args.arch, "http_workers": nw, "gpus": args.gpus or "all", "clients": per * args.client_procs,
                 "client_procs": args.client_procs, "seconds": wall, "requests": len(lat), "errors": errs,
                 "requests_per_s": len(lat) / wall, "host_cores": os.cpu_count(),
                 "latency_ms": {"p50": 1e3 * lat[len(lat) // 2], "p99": 1e3 * lat[int(len(lat) * 0.99) - 1],
                                "max": 1e3 * lat[-1]} if lat else None,
                 "max_batch": args.max_batch, "max_wait_us": args.max_wait_us,
                 "transport": "real HTTP on 127.0.0.1, keep-alive; uploads -> dispatcher through shared memory"}
            results.append(r)
        finally:
            # the launcher, its uvicorn workers and the dispatcher form one process group: stop them all
            import signal
            srv.terminate()
            try:
                srv.wait(timeout=20)
            except subprocess.TimeoutExpired:
                pass
            try:
                os.killpg(srv.pid, signal.SIGKILL)
            except (ProcessLookupError, PermissionError):
                pass
            if results and results[-1].get("http_workers") == nw:
                try:
                    results[-1]["dispatcher"] = json.load(open(stats_path))     # includes the warm-up requests
                except (OSError, ValueError):
                    results[-1]["dispatcher"] = None
                print(json.dumps(results[-1]), flush=True)
    if args.out:
        json.dump(results, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
