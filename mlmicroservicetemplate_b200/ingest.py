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
import pickle
import socket
import struct
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass
from multiprocessing import shared_memory
from multiprocessing.connection import Client, Listener
from typing import Dict, List, Optional, Tuple

import numpy as np

AUTHKEY = b"b2p-ingest"
SLOT_BYTES = 256 << 10          # allocation granule of a ring
KIND_JPEG, KIND_PIXELS = 1, 2
SEND_TIMEOUT_S = 2.0            # dispatcher -> front end; a front end that does not read for this long is dropped


@dataclass
class HostView:
    """`length` bytes at host address `address` that somebody else keeps alive (a slice of a shared-memory ring).
    backend.JpegDecoder.decode accepts it in place of a bytes object."""
    address: int
    length: int

    def __len__(self) -> int:
        return self.length


class ShmRing:
    """A shared-memory segment handed out in runs of SLOT_BYTES granules (first fit; payloads are short-lived).
    The free map is a bytearray (1 = free), so finding a run of n granules is one C-speed substring search."""

    def __init__(self, nbytes: int):
        self.slots = max(4, nbytes // SLOT_BYTES)
        self.shm = shared_memory.SharedMemory(create=True, size=self.slots * SLOT_BYTES)
        self.free = bytearray(b"\x01" * self.slots)
        self.lock = threading.Lock()
        self.cursor = 0

    @property
    def name(self) -> str:
        return self.shm.name

    def alloc(self, nbytes: int) -> Optional[Tuple[int, int]]:
        """-> (first granule, number of granules) or None when the ring is full."""
        need = max(1, -(-nbytes // SLOT_BYTES))
        run = b"\x01" * need
        with self.lock:
            i = self.free.find(run, self.cursor)        # from the cursor to the end, then from the beginning
            if i < 0 and self.cursor:
                i = self.free.find(run)
            if i < 0:
                return None
            self.free[i:i + need] = bytes(need)
            self.cursor = (i + need) % self.slots
            return i, need

    def release(self, first: int, count: int) -> None:
        with self.lock:
            self.free[first:first + count] = b"\x01" * count

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
                # one message carries every answer the dispatcher had ready for this front end: (id, ok, k, probs
                # bytes (float32) | error text, class-id bytes (int32))
                for rid, ok, k, a, b in msg[1]:
                    with self.pending_lock:
                        fut, grant = self.pending.pop(rid, (None, None))
                    if grant is not None:
                        self.ring.release(*grant)
                    if fut is None or fut.done():
                        continue
                    if ok:
                        fut.set_result((np.frombuffer(a, np.float32, k), np.frombuffer(b, np.int32, k)))
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


class _FrontEnd:
    """One connected front-end process: its connection, its mapped ring and the answers waiting to be sent."""
    __slots__ = ("conn", "shm", "base", "base_addr", "outbox", "rx", "sock")

    def __init__(self, conn, shm, base):
        self.conn = conn
        # answers go out through a socket object over the same descriptor so that a send can TIME OUT: one sender
        # thread serves every front end, and a front end that stopped reading must not stall the others for ever
        self.sock = socket.socket(fileno=os.dup(conn.fileno()))
        self.sock.settimeout(SEND_TIMEOUT_S)
        self.shm = shm
        self.base = base
        self.base_addr = base.ctypes.data
        self.outbox: List[tuple] = []
        self.rx = bytearray()             # bytes read from the socket that do not yet make a whole message


class DispatcherServer:
    """Dispatcher side: accepts front-end connections and feeds their requests to ONE runtime.Engine.

    Two I/O threads serve ALL front ends: one receives (a select over every connection, each drained while it has
    messages), one sends.  A thread per connection looked natural and was the first version, but the interpreter lock
    turned it into a convoy: measured with a stand-in engine on 8 cores, 18 k req/s with one front end fell to 12 k with
    four.  Answers are coalesced: the engine resolves a whole micro-batch in one go, each done-callback only appends to
    its front end's outbox, and the sender then ships everything that piled up for a front end as ONE message."""

    def __init__(self, engine, socket_path: str, info: dict):
        self.engine = engine
        self.info = dict(info)
        self.socket_path = socket_path
        if os.path.exists(socket_path):
            os.unlink(socket_path)
        self.listener = Listener(socket_path, family="AF_UNIX", authkey=AUTHKEY)
        self.stopping = False
        self.requests = 0
        self.fronts: Dict[object, _FrontEnd] = {}        # connection -> front end
        self.fronts_lock = threading.Lock()
        self.out_lock = threading.Lock()
        self.out_wake = threading.Event()
        self.out_dirty: List[_FrontEnd] = []
        self.threads = [threading.Thread(target=self._accept_loop, name="b2p-dispatch-accept", daemon=True),
                        threading.Thread(target=self._recv_loop, name="b2p-dispatch-recv", daemon=True),
                        threading.Thread(target=self._send_loop, name="b2p-dispatch-send", daemon=True)]
        for t in self.threads:
            t.start()

    # ---- accept: handshake (hello -> map the ring -> welcome), then the connection joins the receive set
    def _accept_loop(self) -> None:
        while not self.stopping:
            try:
                conn = self.listener.accept()
            except (OSError, EOFError):
                break
            try:
                hello = conn.recv()
                if hello[0] != "hello":
                    conn.send(("error", "expected hello"))
                    conn.close()
                    continue
                _, name, size, pid = hello
                shm = shared_memory.SharedMemory(name=name)
                # the segment belongs to the front end: keep this process's resource tracker from unlinking it at exit
                try:
                    from multiprocessing import resource_tracker
                    resource_tracker.unregister(shm._name, "shared_memory")
                except Exception:
                    pass
                fe = _FrontEnd(conn, shm, np.frombuffer(shm.buf, dtype=np.uint8))
                conn.send(("welcome", self.info))
            except (EOFError, OSError, ValueError):
                try:
                    conn.close()
                except OSError:
                    pass
                continue
            with self.fronts_lock:
                self.fronts[conn] = fe

    # ---- receive
    def _recv_loop(self) -> None:
        from multiprocessing.connection import wait
        from .runtime import JpegFile
        while not self.stopping:
            with self.fronts_lock:
                conns = list(self.fronts)
            if not conns:
                time.sleep(0.005)
                continue
            try:
                ready = wait(conns, timeout=0.05)
            except (OSError, ValueError):
                # a connection was closed between the snapshot and the wait: drop the dead ones, read nothing (a read
                # on a healthy connection that has no data would block every front end)
                ready = []
                for c in conns:
                    try:
                        dead = c.closed or c.fileno() < 0
                    except (OSError, ValueError):
                        dead = True
                    if dead:
                        self._drop(c)
            for conn in ready:
                fe = self.fronts.get(conn)
                if fe is None:
                    continue
                try:
                    # one read takes everything the front end has sent so far; the connection's framing is a 4-byte
                    # big-endian length followed by a pickle (multiprocessing.connection), parsed here in bulk
                    try:
                        chunk = os.read(conn.fileno(), 1 << 18)
                    except BlockingIOError:          # (the send timeout makes the descriptor non-blocking)
                        continue
                    if not chunk:
                        raise EOFError
                    rx = fe.rx
                    rx += chunk
                    pos, end = 0, len(rx)
                    while end - pos >= 4:
                        n = int.from_bytes(rx[pos:pos + 4], "big", signed=True)
                        if n < 0:
                            raise ValueError("oversized message")
                        if end - pos - 4 < n:
                            break
                        msg = pickle.loads(memoryview(rx)[pos + 4:pos + 4 + n])
                        pos += 4 + n
                        if msg[0] == "bye":
                            raise EOFError
                        _, rid, kind, h, w, off, nbytes, inline = msg
                        self.requests += 1
                        if kind == KIND_JPEG:
                            data = inline if off < 0 else HostView(fe.base_addr + off, nbytes)
                            image = JpegFile(data, h, w)
                        else:
                            flat = np.frombuffer(inline, dtype=np.uint8) if off < 0 else fe.base[off:off + nbytes]
                            image = flat.reshape(h, w, 3)
                        fut = self.engine.submit(image)
                        fut.add_done_callback(lambda f, rid=rid, fe=fe: self._reply(fe, rid, f))
                    del rx[:pos]
                except (EOFError, OSError, ValueError, pickle.UnpicklingError):
                    self._drop(conn)

    def _reply(self, fe: _FrontEnd, rid: int, fut) -> None:
        try:
            probs, idx = fut.result()
            p32 = np.ascontiguousarray(probs, dtype=np.float32)
            item = (rid, True, int(p32.size), p32.tobytes(), np.ascontiguousarray(idx, dtype=np.int32).tobytes())
        except BaseException as e:
            item = (rid, False, 0, f"{type(e).__name__}: {e}", None)
        with self.out_lock:
            if not fe.outbox:
                self.out_dirty.append(fe)
            fe.outbox.append(item)
        self.out_wake.set()

    # ---- send
    def _send_loop(self) -> None:
        while True:
            self.out_wake.wait(0.1)
            self.out_wake.clear()
            with self.out_lock:
                dirty, self.out_dirty = self.out_dirty, []
                work = [(fe, fe.outbox) for fe in dirty]
                for fe in dirty:
                    fe.outbox = []
            for fe, items in work:
                try:
                    payload = pickle.dumps(("res", items), protocol=pickle.HIGHEST_PROTOCOL)
                    fe.sock.sendall(struct.pack("!i", len(payload)) + payload)     # multiprocessing.connection framing
                except (OSError, ValueError):        # gone, or not reading (socket.timeout is an OSError)
                    self._drop(fe.conn)
            if self.stopping and not work:
                return

    def _drop(self, conn) -> None:
        with self.fronts_lock:
            fe = self.fronts.pop(conn, None)
        if fe is None:
            return
        for closer in (conn.close, fe.sock.close):
            try:
                closer()
            except OSError:
                pass
        fe.base = None
        try:
            fe.shm.close()
        except BufferError:
            # requests still in flight hold views into the segment: leave the mapping to them (it is unmapped when
            # the last view dies) and only drop this object's handles
            try:
                os.close(fe.shm._fd)
            except OSError:
                pass
            fe.shm._fd = -1
            fe.shm._buf = None
            fe.shm._mmap = None

    def close(self) -> None:
        self.stopping = True
        self.out_wake.set()
        try:
            self.listener.close()
        except OSError:
            pass
        for conn in list(self.fronts):
            self._drop(conn)
        if os.path.exists(self.socket_path):
            try:
                os.unlink(self.socket_path)
            except OSError:
                pass
