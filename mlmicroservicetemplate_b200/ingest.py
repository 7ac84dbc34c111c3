"""HTTP ingest scaling: several front-end server processes, ONE process that owns the GPUs.

The reference serves everything from one uvicorn process (docker-compose.yml:6) and one event loop
(src/server/main.py:119-135).  Multipart parsing and routing cost about a millisecond of Python per request, so a
single process tops out near 1.2 k requests/s while one B200 replica classifies ~80 k images/s.  Here

    N front-end processes (uvicorn workers running src/server/main.py unchanged)
        parse the upload, check the JPEG header, copy the file bytes (or, for non-JPEG uploads, the decoded pixels)
        ONCE into their own shared-memory ring and send a 60-byte message;
    1 dispatcher process (this module's DispatcherServer around runtime.Engine)
        maps every ring, builds zero-copy views of the payloads, micro-batches ACROSS all front ends and drives one
        replica per GPU; results (k probabilities + k class ids) travel back over the same socket.

No pixel or file byte crosses a pipe: the socket carries offsets.  The GPU JPEG decoder reads the file bytes straight
out of the shared ring (HostView), so an upload is copied exactly once on the host between the HTTP body and HBM.

The front end's handle on all this is RemoteEngine, which has the duck type of runtime.Engine that the model plugin
uses (submit -> Future, close), so src/model/model.py is the same code in both deployments.
"""
from __future__ import annotations

import os
import threading
from concurrent.futures import Future
from dataclasses import dataclass
from multiprocessing import shared_memory
from multiprocessing.connection import Client, Listener
from typing import Dict, List, Optional, Tuple

import numpy as np

AUTHKEY = b"b2p-ingest"
SLOT_BYTES = 256 << 10          # allocation granule of a ring
KIND_JPEG, KIND_PIXELS = 1, 2


@dataclass
class HostView:
    """`length` bytes at host address `address` that somebody else keeps alive (a slice of a shared-memory ring).
    backend.JpegDecoder.decode accepts it in place of a bytes object."""
    address: int
    length: int

    def __len__(self) -> int:
        return self.length


class ShmRing:
    """A shared-memory segment handed out in runs of SLOT_BYTES granules (first fit; payloads are short-lived)."""

    def __init__(self, nbytes: int):
        self.slots = max(4, nbytes // SLOT_BYTES)
        self.shm = shared_memory.SharedMemory(create=True, size=self.slots * SLOT_BYTES)
        self.free = np.ones(self.slots, dtype=bool)
        self.lock = threading.Lock()
        self.cursor = 0

    @property
    def name(self) -> str:
        return self.shm.name

    def alloc(self, nbytes: int) -> Optional[Tuple[int, int]]:
        """-> (first granule, number of granules) or None when the ring is full."""
        need = max(1, -(-nbytes // SLOT_BYTES))
        with self.lock:
            n = self.slots
            start = self.cursor
            for _ in range(2):                      # from the cursor to the end, then from the beginning
                i = start
                while i + need <= n:
                    run = self.free[i:i + need]
                    if run.all():
                        self.free[i:i + need] = False
                        self.cursor = (i + need) % n
                        return i, need
                    i += int(np.argmin(run)) + 1    # jump past the first busy granule of the window
                if start == 0:
                    break
                start = 0
        return None

    def release(self, first: int, count: int) -> None:
        with self.lock:
            self.free[first:first + count] = True

    def close(self) -> None:
        try:
            self.shm.close()
            self.shm.unlink()
        except FileNotFoundError:
            pass


class RemoteEngine:
    """Front-end side: the Engine duck type (submit / close) in front of a dispatcher process."""

    def __init__(self, socket_path: str, ring_bytes: int = 256 << 20):
        self.conn = Client(socket_path, family="AF_UNIX", authkey=AUTHKEY)
        self.ring = ShmRing(ring_bytes)
        self.send_lock = threading.Lock()
        self.pending: Dict[int, Tuple[Future, Optional[Tuple[int, int]]]] = {}
        self.pending_lock = threading.Lock()
        self.next_id = 0
        self._closed = False
        self.conn.send(("hello", self.ring.name, self.ring.slots * SLOT_BYTES, os.getpid()))
        tag, info = self.conn.recv()
        if tag != "welcome":
            raise RuntimeError(f"dispatcher refused the connection: {info}")
        self.info = info                      # {"classes": [...], "n_classes": int, "decode": "gpu"|"host", "gpus": int}
        self.classes = info.get("classes")
        self.workers = [None] * int(info.get("gpus", 1))
        self.reader = threading.Thread(target=self._read_loop, name="b2p-ingest-reader", daemon=True)
        self.reader.start()

    def submit(self, image) -> Future:
        """image: runtime.JpegFile (file bytes, decoded on the GPU) or a decoded uint8 HWC array."""
        if self._closed:
            raise RuntimeError("engine is shut down")
        fut: Future = Future()
        if isinstance(image, np.ndarray):
            kind, h, w = KIND_PIXELS, int(image.shape[0]), int(image.shape[1])
            payload = memoryview(np.ascontiguousarray(image)).cast("B")
        else:
            kind, h, w = KIND_JPEG, int(image.height), int(image.width)
            payload = memoryview(image.data)
        nbytes = len(payload)
        grant = self.ring.alloc(nbytes)
        with self.pending_lock:
            rid = self.next_id
            self.next_id += 1
            self.pending[rid] = (fut, grant)
        if grant is not None:
            off = grant[0] * SLOT_BYTES
            self.ring.shm.buf[off:off + nbytes] = payload       # the one host copy of this upload
            msg = ("req", rid, kind, h, w, off, nbytes, None)
        else:
            msg = ("req", rid, kind, h, w, -1, nbytes, bytes(payload))   # ring full: the bytes ride on the socket
        with self.send_lock:
            self.conn.send(msg)
        return fut

    def _read_loop(self) -> None:
        try:
            while True:
                msg = self.conn.recv()
                if msg[0] != "res":
                    continue
                _, rid, ok, a, b = msg
                with self.pending_lock:
                    fut, grant = self.pending.pop(rid, (None, None))
                if grant is not None:
                    self.ring.release(*grant)
                if fut is None or fut.done():
                    continue
                if ok:
                    fut.set_result((a, b))
                else:
                    fut.set_exception(RuntimeError(a))
        except (EOFError, OSError):
            pass
        finally:
            with self.pending_lock:
                left = list(self.pending.values())
                self.pending.clear()
            for fut, _ in left:
                if not fut.done():
                    fut.set_exception(RuntimeError("the GPU dispatcher went away"))

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            with self.send_lock:
                self.conn.send(("bye",))
        except (OSError, ValueError):
            pass
        try:
            self.conn.close()
        except OSError:
            pass
        self.ring.close()


class DispatcherServer:
    """Dispatcher side: accepts front-end connections and feeds their requests to ONE runtime.Engine."""

    def __init__(self, engine, socket_path: str, info: dict):
        self.engine = engine
        self.info = dict(info)
        self.socket_path = socket_path
        if os.path.exists(socket_path):
            os.unlink(socket_path)
        self.listener = Listener(socket_path, family="AF_UNIX", authkey=AUTHKEY)
        self.threads: List[threading.Thread] = []
        self.stopping = False
        self.requests = 0
        self.accept_thread = threading.Thread(target=self._accept_loop, name="b2p-dispatch-accept", daemon=True)
        self.accept_thread.start()

    def _accept_loop(self) -> None:
        while not self.stopping:
            try:
                conn = self.listener.accept()
            except (OSError, EOFError):
                break
            t = threading.Thread(target=self._serve, args=(conn,), name="b2p-dispatch-conn", daemon=True)
            t.start()
            self.threads.append(t)

    def _serve(self, conn) -> None:
        from .runtime import JpegFile
        shm = None
        send_lock = threading.Lock()

        def reply(rid, fut):
            try:
                probs, idx = fut.result()
                msg = ("res", rid, True, np.asarray(probs), np.asarray(idx))
            except BaseException as e:
                msg = ("res", rid, False, f"{type(e).__name__}: {e}", None)
            try:
                with send_lock:
                    conn.send(msg)
            except (OSError, ValueError):
                pass

        try:
            hello = conn.recv()
            if hello[0] != "hello":
                conn.send(("error", "expected hello"))
                return
            _, name, size, pid = hello
            shm = shared_memory.SharedMemory(name=name)
            # the segment belongs to the front end: keep this process's resource tracker from unlinking it at exit
            try:
                from multiprocessing import resource_tracker
                resource_tracker.unregister(shm._name, "shared_memory")
            except Exception:
                pass
            base = np.frombuffer(shm.buf, dtype=np.uint8)
            base_addr = base.ctypes.data
            conn.send(("welcome", self.info))
            while True:
                msg = conn.recv()
                if msg[0] == "bye":
                    break
                _, rid, kind, h, w, off, nbytes, inline = msg
                self.requests += 1
                if kind == KIND_JPEG:
                    data = inline if off < 0 else HostView(base_addr + off, nbytes)
                    image = JpegFile(data, h, w)
                else:
                    flat = np.frombuffer(inline, dtype=np.uint8) if off < 0 else base[off:off + nbytes]
                    image = flat.reshape(h, w, 3)
                fut = self.engine.submit(image)
                fut.add_done_callback(lambda f, rid=rid: reply(rid, f))
        except (EOFError, OSError):
            pass
        finally:
            try:
                conn.close()
            except OSError:
                pass
            if shm is not None:
                base = None
                try:
                    shm.close()
                except BufferError:
                    # requests still in flight hold views into the segment: leave the mapping to them (it is
                    # unmapped when the last view dies) and only drop this object's handles
                    try:
                        os.close(shm._fd)
                    except OSError:
                        pass
                    shm._fd = -1
                    shm._buf = None
                    shm._mmap = None

    def close(self) -> None:
        self.stopping = True
        try:
            self.listener.close()
        except OSError:
            pass
        if os.path.exists(self.socket_path):
            try:
                os.unlink(self.socket_path)
            except OSError:
                pass
