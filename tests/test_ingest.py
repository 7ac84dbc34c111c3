"""SURVEY §8(f) rank 2 on CPU: front-end processes hand uploads to the dispatcher process through shared memory
(ingest.py).  The dispatcher here wraps runtime.Engine around a recording stand-in for the GPU worker (test-only; the
product has no CPU worker); the front end runs in a SEPARATE process so the bytes really cross a process boundary."""
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import textwrap

import numpy as np

from mlmicroservicetemplate_b200.ingest import SLOT_BYTES, DispatcherServer, HostView, RemoteEngine, ShmRing
from mlmicroservicetemplate_b200.runtime import Engine, JpegFile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class EchoWorker:
    """run_batch stand-in: class id = first payload byte + payload length % 251, read IN PLACE from wherever the
    dispatcher put the view (shared memory for ring payloads)."""

    def __init__(self):
        self.batches = []
        self.saw_host_view = 0

    def run_batch(self, images):
        self.batches.append(len(images))
        idx = []
        for im in images:
            if isinstance(im, JpegFile):
                if isinstance(im.data, HostView):
                    self.saw_host_view += 1
                    first, n = ctypes.string_at(im.data.address, 1)[0], im.data.length
                else:
                    first, n = im.data[0], len(im.data)
            else:
                first, n = int(im.reshape(-1)[0]), im.size
            idx.append([first, n % 251])
        return np.ones((len(images), 2), np.float32), np.array(idx, np.int32)


CLIENT = textwrap.dedent("""
    import json, sys
    import numpy as np
    sys.path.insert(0, %r)
    from mlmicroservicetemplate_b200.ingest import RemoteEngine
    from mlmicroservicetemplate_b200.runtime import JpegFile
    eng = RemoteEngine(sys.argv[1], ring_bytes=int(sys.argv[2]))
    futs, want = [], []
    for i in range(40):
        if i %% 2:
            data = bytes([i]) + b"x" * (1000 + 37 * i)
            futs.append(eng.submit(JpegFile(data, 8, 8)))
            want.append([i, len(data) %% 251])
        else:
            im = np.full((16 + i, 24, 3), i, np.uint8)
            futs.append(eng.submit(im))
            want.append([i, im.size %% 251])
    got = [[int(v) for v in f.result(timeout=30)[1]] for f in futs]
    big = eng.submit(JpegFile(bytes([7]) * (3 << 20), 8, 8)).result(timeout=30)[1]     # larger than a small ring
    print("RESULT " + json.dumps({"ok": got == want, "info": eng.info, "big": [int(v) for v in big]}))
    eng.close()
""" % ROOT)


def _run_client(sock, ring_bytes):
    p = subprocess.run([sys.executable, "-c", CLIENT, sock, str(ring_bytes)], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr[-2000:]
    return json.loads([l for l in p.stdout.splitlines() if l.startswith("RESULT ")][0][7:])


def test_requests_cross_the_process_boundary_through_shared_memory():
    w = EchoWorker()
    eng = Engine([w], max_batch=8, max_wait_us=2000)
    sock = os.path.join(tempfile.mkdtemp(), "d.sock")
    srv = DispatcherServer(eng, sock, {"classes": ["a", "b"], "n_classes": 2, "decode": "gpu", "gpus": 1})
    try:
        r = _run_client(sock, 64 << 20)
        assert r["ok"], "every request must come back with its own payload's answer"
        assert r["info"]["n_classes"] == 2 and r["info"]["decode"] == "gpu"
        assert r["big"] == [7, (3 << 20) % 251]
        assert w.saw_host_view >= 20, "JPEG payloads must be read in place from the shared ring"
        assert max(w.batches) > 1, "requests of one front end are micro-batched in the dispatcher"
        # a ring too small for the big payload: the bytes ride on the socket instead, same answers
        r2 = _run_client(sock, 4 * SLOT_BYTES)
        assert r2["ok"] and r2["big"] == [7, (3 << 20) % 251]
    finally:
        srv.close()
        eng.close()


def test_two_front_ends_share_one_dispatcher():
    w = EchoWorker()
    eng = Engine([w], max_batch=16, max_wait_us=5000)
    sock = os.path.join(tempfile.mkdtemp(), "d.sock")
    srv = DispatcherServer(eng, sock, {"classes": ["a"], "n_classes": 1, "decode": "host", "gpus": 1})
    try:
        procs = [subprocess.Popen([sys.executable, "-c", CLIENT, sock, str(32 << 20)], stdout=subprocess.PIPE,
                                  stderr=subprocess.PIPE, text=True) for _ in range(2)]
        for p in procs:
            out, err = p.communicate(timeout=120)
            assert p.returncode == 0, err[-2000:]
            assert json.loads([l for l in out.splitlines() if l.startswith("RESULT ")][0][7:])["ok"]
        assert sum(w.batches) == 2 * 41
    finally:
        srv.close()
        eng.close()


def test_ring_allocator_first_fit_and_reuse():
    ring = ShmRing(8 * SLOT_BYTES)
    try:
        a = ring.alloc(SLOT_BYTES)             # 1 granule
        b = ring.alloc(3 * SLOT_BYTES - 5)     # 3 granules
        c = ring.alloc(4 * SLOT_BYTES)         # 4 granules -> full
        assert a == (0, 1) and b == (1, 3) and c == (4, 4)
        assert ring.alloc(1) is None
        ring.release(*b)
        assert ring.alloc(4 * SLOT_BYTES) is None, "no run of four free granules"
        assert ring.alloc(2 * SLOT_BYTES) == (1, 2)
        assert ring.alloc(SLOT_BYTES) == (3, 1)
        ring.release(*a)
        assert ring.alloc(10) == (0, 1)
    finally:
        ring.close()


def test_remote_engine_in_process_round_trip_and_dispatcher_loss():
    w = EchoWorker()
    eng = Engine([w], max_batch=4, max_wait_us=500)
    sock = os.path.join(tempfile.mkdtemp(), "d.sock")
    srv = DispatcherServer(eng, sock, {"classes": ["a"], "n_classes": 1, "decode": "gpu", "gpus": 1})
    remote = RemoteEngine(sock, ring_bytes=8 * SLOT_BYTES)
    try:
        im = np.full((10, 10, 3), 9, np.uint8)
        probs, idx = remote.submit(im).result(timeout=10)
        assert [int(v) for v in idx] == [9, im.size % 251]
        assert all(remote.ring.free), "the granules of an answered request go back to the ring"
    finally:
        remote.close()
        srv.close()
        eng.close()


def test_http_front_end_process_uses_the_dispatcher(tmp_path):
    """A real uvicorn front-end process (src/server/main.py unchanged, ENGINE_SOCKET set) in front of a dispatcher:
    POST /predict answers in the reference's envelope (reference src/server/main.py:137-140) with the dispatcher's
    class names, and the JPEG bytes reached the dispatcher in place (HostView)."""
    import http.client
    import io
    import socket
    import time
    from PIL import Image
    sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
    from serve_http import multipart
    w = EchoWorker()
    eng = Engine([w], max_batch=4, max_wait_us=500)
    sock = str(tmp_path / "d.sock")
    names = [f"c{i}" for i in range(256)]
    srv = DispatcherServer(eng, sock, {"classes": names, "n_classes": 256, "decode": "gpu", "gpus": 1})
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, ENGINE_SOCKET=sock, MODEL="resnet50", REGISTER="0", PYTHONPATH=ROOT)
    proc = subprocess.Popen([sys.executable, "-m", "uvicorn", "mlmicroservicetemplate_b200.src.server.main:app",
                             "--host", "127.0.0.1", "--port", str(port), "--log-level", "warning"], env=env, cwd=ROOT)
    try:
        t0 = time.time()
        status = None
        while time.time() - t0 < 60:
            try:
                c = http.client.HTTPConnection("127.0.0.1", port, timeout=5)
                c.request("GET", "/status")
                r = c.getresponse(); r.read(); status = r.status; c.close()
                if status == 200:
                    break
            except OSError:
                pass
            time.sleep(0.2)
        assert status == 200
        buf = io.BytesIO()
        Image.fromarray(np.full((32, 32, 3), 120, np.uint8)).save(buf, format="JPEG", quality=90)
        jpeg = buf.getvalue()
        body, ctype = multipart(jpeg, "x.jpg")
        c = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
        c.request("POST", "/predict", body=body, headers={"Content-Type": ctype})
        r = c.getresponse()
        out = json.loads(r.read())
        assert r.status == 200 and out["status"] == "success"
        assert out["result"]["classes"] == names
        assert set(out["result"]["result"]) == {names[jpeg[0]], names[len(jpeg) % 251]}
        assert w.saw_host_view == 1
    finally:
        proc.terminate()
        proc.wait(timeout=30)
        srv.close()
        eng.close()
