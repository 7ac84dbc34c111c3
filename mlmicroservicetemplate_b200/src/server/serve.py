"""Multi-process deployment of the unchanged HTTP contract:

    python -m mlmicroservicetemplate_b200.src.server.serve --workers 8 --port 8000

starts ONE dispatcher process that owns the GPU replicas (the model plugin's init(), then ingest.DispatcherServer on
a Unix socket) and N uvicorn worker processes running src/server/main.py as it is; the workers' plugin connects to the
dispatcher (ENGINE_SOCKET) instead of building replicas.  The reference runs a single uvicorn process
(docker-compose.yml:6, `uvicorn src.server.main:app`); with --workers 1 and no dispatcher this package does the same.

    python -m mlmicroservicetemplate_b200.src.server.serve --dispatcher-only --socket /tmp/b2p.sock
runs just the dispatcher (what the launcher spawns).
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import tempfile
import threading
import time


def run_dispatcher(socket_path: str) -> int:
    from ...ingest import DispatcherServer
    from ..model import model as plugin
    os.environ.pop("ENGINE_SOCKET", None)
    plugin.init()
    engine = plugin._state["engine"]
    if engine is None:
        print("dispatcher: MODEL=template needs no dispatcher", file=sys.stderr)
        return 2
    names = plugin._state["classes"]
    info = {"classes": list(names), "n_classes": len(names), "decode": plugin._state["decode"],
            "gpus": len(engine.workers), "pid": os.getpid()}
    server = DispatcherServer(engine, socket_path, info)
    stop = threading.Event()
    for sig in (signal.SIGTERM, signal.SIGINT):
        signal.signal(sig, lambda *_: stop.set())
    print(f"dispatcher ready on {socket_path}: {info['gpus']} GPU replica(s), {info['n_classes']} classes", flush=True)
    stop.wait()
    server.close()
    stats_path = os.getenv("B2P_DISPATCH_STATS")
    if stats_path:
        # what the micro-batcher achieved: batch-size histogram and each GPU's share of the images
        import collections
        import json
        hist = collections.Counter(engine.batch_sizes)
        share = collections.Counter()
        for w, n in engine.dispatch_log:
            share[w] += n
        total = max(1, sum(share.values()))
        with open(stats_path, "w") as f:
            json.dump({"requests": server.requests, "batches": len(engine.batch_sizes),
                       "mean_batch": sum(engine.batch_sizes) / max(1, len(engine.batch_sizes)),
                       "batch_histogram": {str(k): v for k, v in sorted(hist.items())},
                       "per_gpu_share": {str(g): share.get(g, 0) / total for g in range(info["gpus"])}}, f)
    plugin.shutdown()
    return 0


def wait_for_socket(path: str, proc: subprocess.Popen, timeout_s: float = 600.0) -> None:
    from multiprocessing.connection import Client
    from ...ingest import AUTHKEY
    t0 = time.time()
    while time.time() - t0 < timeout_s:
        if proc.poll() is not None:
            raise RuntimeError(f"the dispatcher exited with code {proc.returncode} during start-up")
        if os.path.exists(path):
            try:
                Client(path, family="AF_UNIX", authkey=AUTHKEY).close()
                return
            except (OSError, EOFError):
                pass
        time.sleep(0.1)
    raise RuntimeError("the dispatcher did not come up in time")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=int(os.getenv("HTTP_WORKERS", "4")))
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=int(os.getenv("PORT", "8000")))
    ap.add_argument("--socket", default="")
    ap.add_argument("--dispatcher-only", action="store_true")
    ap.add_argument("--log-level", default="warning")
    args = ap.parse_args(argv)
    sock = args.socket or os.path.join(tempfile.gettempdir(), f"b2p-dispatch-{os.getpid()}.sock")
    if args.dispatcher_only:
        return run_dispatcher(sock)
    import uvicorn
    disp = subprocess.Popen([sys.executable, "-m", "mlmicroservicetemplate_b200.src.server.serve", "--dispatcher-only",
                             "--socket", sock])
    try:
        wait_for_socket(sock, disp)
        os.environ["ENGINE_SOCKET"] = sock
        os.environ.setdefault("REGISTER", "0" if args.workers > 1 else "1")   # one heartbeat is enough
        uvicorn.run("mlmicroservicetemplate_b200.src.server.main:app", host=args.host, port=args.port,
                    workers=args.workers, log_level=args.log_level)
    finally:
        disp.terminate()
        try:
            disp.wait(timeout=30)
        except subprocess.TimeoutExpired:
            disp.kill()
    return 0


if __name__ == "__main__":
    sys.exit(main())
