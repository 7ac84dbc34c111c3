"""Host runtime under the model plugin: micro-batching queue, per-GPU replica workers and the
round-robin dispatcher.

The reference serves one request at a time on the event loop (src/server/main.py:135 calls
predict() synchronously).  Here predict() enqueues a decoded uint8 HWC image and waits on a future;
a batcher thread groups requests (up to MAX_BATCH, or whatever arrived within MAX_WAIT_US of the
first one) and hands whole batches to one worker thread per GPU — to the replica with the fewest images
in flight, in rotation while they are level.  A worker copies the batch into one of its TWO pinned host
slabs and makes ONE asynchronous call into the C ABI (b2p_predict_host): H2D copy, K1 preprocess, K2
forward, K3 softmax/top-k and the D2H copy of k (prob, index) pairs per image; it submits batch i+1
before it waits for batch i, so the host-side packing of one batch runs under the GPU work of the other.
Uploads that are baseline JPEGs can skip the host decoder altogether (JpegFile): their bytes are
decoded by b2p_jpeg_decode straight into the device slab and the device-resident chain (b2p_predict)
runs on it.

The path shards by request: replicas are independent, no tensor crosses GPUs, so there is no
collective anywhere.

Workers are duck-typed (`run_batch(images, geoms) -> (probs, idx)`) so the queueing logic is tested on
CPU with a recording stand-in (tests/test_runtime.py); the only product implementation is
GpuReplicaWorker, which has no CPU fallback.
"""
from __future__ import annotations

import queue
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np


def preset_geometry(h: int, w: int, resize: int = 256, crop: int = 224) -> Tuple[int, int, int, int]:
    """torchvision ImageClassification preset (SP/torchvision/transforms/_presets.py:58-65): short side ->
    256 with the long side int(256*long/short) (functional.py:368-370), centre crop with
    int(round((size-224)/2)) (functional.py:592-593).  Returns (res_h, res_w, top, left)."""
    if w <= h:
        res_w, res_h = resize, int(resize * h / w)
    else:
        res_h, res_w = resize, int(resize * w / h)
    return res_h, res_w, int(round((res_h - crop) / 2.0)), int(round((res_w - crop) / 2.0))


def image_geometry(h: int, w: int, mode: str) -> Tuple[int, int, int, int]:
    """'direct': resize straight to 224x224; 'preset': resize-256 + centre-crop-224."""
    if mode == "preset":
        return preset_geometry(h, w)
    if mode == "direct":
        return 224, 224, 0, 0
    raise ValueError(f"unknown preprocess mode {mode!r}")


@dataclass
class JpegFile:
    """An upload the GPU decoder covers (b2p_jpeg_inspect said yes): the file bytes travel to the device as they
    are and the pixels first exist in HBM.  Everything else reaches the engine as decoded uint8 HWC pixels."""
    data: bytes
    height: int
    width: int


@dataclass
class PackedBatch:
    """A whole batch that is already packed the way the C ABI wants it: `slab` is a pinned uint8 torch tensor holding
    n decoded HWC images back to back, `geoms` their (offset, h, w, res_h, res_w, top, left) tuples.  Producers that
    own pinned memory (decode threads, the ingest processes' shared slab, bench.py) hand such batches to
    Engine.submit_batch(): no pixel is copied again on the host."""
    slab: object
    nbytes: int
    geoms: Sequence[Tuple[int, int, int, int, int, int, int]]
    descs: object = None               # b2p_image array, built once by the first replica that sees the batch

    def __len__(self) -> int:
        return len(self.geoms)


class CorruptImage(ValueError):
    """The entropy-coded data of one upload was damaged (the GPU decoder's status word was non-zero)."""


@dataclass
class _Request:
    image: object                      # np.ndarray (uint8 HWC) or JpegFile
    future: Future
    t_enqueue: float = field(default_factory=time.perf_counter)


@dataclass
class _BatchRequest:
    images: object                     # list of images / JpegFiles, or a PackedBatch
    future: Future
    n: int


class GpuReplicaWorker:
    """One replica: a b2p handle on one GPU, its stream and two pinned host slabs (+ result buffers), used
    alternately so that batch i+1 is packed and submitted while batch i is still on the GPU."""

    DEPTH = 2

    def __init__(self, device: int, program, max_batch: int, topk: int, mode: str, sub_batch: int = 0,
                 slab_bytes: int = 64 << 20):
        import torch
        from . import backend as B
        if not torch.cuda.is_available():
            raise RuntimeError("GpuReplicaWorker needs a CUDA device; there is no CPU implementation of this path")
        self.torch = torch
        self.B = B
        self.device = device
        self.max_batch = max_batch
        self.topk = topk
        self.mode = mode
        with torch.cuda.device(device):
            self.replica = B.Replica(device, program, max_batch=max_batch, sub_batch=sub_batch)
            self.stream = torch.cuda.Stream(device=device)
            self.slabs = [torch.empty(slab_bytes, dtype=torch.uint8, pin_memory=True) for _ in range(self.DEPTH)]
            self.probs = [torch.empty((max_batch, topk), dtype=torch.float32, pin_memory=True) for _ in range(self.DEPTH)]
            self.idx = [torch.empty((max_batch, topk), dtype=torch.int32, pin_memory=True) for _ in range(self.DEPTH)]
            self.done = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.slabs_np = [t.numpy() for t in self.slabs]
        self.slot = 0
        # GPU JPEG decode (created on first use): decoder object, device slabs, device result buffers
        self.jpeg = None
        self.dev_slab = [None] * self.DEPTH
        self.dev_status = None
        self.dev_probs = None
        self.dev_idx = None
        self.status_host = None

    def _ensure_jpeg(self, slot: int, nbytes: int) -> None:
        torch = self.torch
        dev = f"cuda:{self.device}"
        if self.jpeg is None:
            self.jpeg = self.B.JpegDecoder(self.device)
            self.dev_status = [torch.zeros(self.max_batch, dtype=torch.int32, device=dev) for _ in range(self.DEPTH)]
            self.dev_probs = [torch.empty((self.max_batch, self.topk), dtype=torch.float32, device=dev) for _ in range(self.DEPTH)]
            self.dev_idx = [torch.empty((self.max_batch, self.topk), dtype=torch.int32, device=dev) for _ in range(self.DEPTH)]
            self.status_host = [torch.zeros(self.max_batch, dtype=torch.int32, pin_memory=True) for _ in range(self.DEPTH)]
        if self.dev_slab[slot] is None or nbytes > self.dev_slab[slot].numel():
            self.stream.synchronize()
            self.dev_slab[slot] = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=dev)

    def _ensure_slab(self, slot: int, nbytes: int) -> None:
        if nbytes > self.slabs[slot].numel():
            self.done[slot].synchronize()        # the previous batch of this slot has left the slab
            self.slabs[slot] = self.torch.empty(int(nbytes * 1.25), dtype=self.torch.uint8, pin_memory=True)
            self.slabs_np[slot] = self.slabs[slot].numpy()

    def submit_batch(self, images: Sequence[np.ndarray]):
        """Packs the batch into the next pinned slab and queues the whole chain on the stream; returns a ticket for
        collect().  At most DEPTH tickets may be outstanding (the slab of a ticket is reused DEPTH submits later)."""
        n = len(images)
        if n == 0 or n > self.max_batch:
            raise ValueError(f"batch of {n} outside 1..{self.max_batch}")
        if isinstance(images, PackedBatch):
            # already in pinned memory in slab layout: one asynchronous call, nothing to pack
            if images.descs is None:
                images.descs = self.B.make_images(list(images.geoms))
            s = self.slot
            self.slot = (s + 1) % self.DEPTH
            self.done[s].synchronize()
            with self.torch.cuda.device(self.device):
                self.replica.predict_host(images.slab, images.nbytes, images.descs, n, self.topk, self.probs[s],
                                          self.idx[s], stream=self.stream)
                self.done[s].record(self.stream)
            return (s, n, ())
        offs, shapes, total = [], [], 0
        for im in images:
            if isinstance(im, JpegFile):
                h, w = im.height, im.width
            else:
                if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] != 3:
                    raise ValueError("images must be uint8 HWC with 3 channels")
                h, w = im.shape[0], im.shape[1]
            offs.append(total)
            shapes.append((h, w))
            total += (h * w * 3 + 255) // 256 * 256
        geoms = [(o, h, w, *image_geometry(h, w, self.mode)) for o, (h, w) in zip(offs, shapes)]
        descs = self.B.make_images(geoms)
        jpeg_items = [i for i, im in enumerate(images) if isinstance(im, JpegFile)]
        host_items = [i for i, im in enumerate(images) if not isinstance(im, JpegFile)]
        s = self.slot
        self.slot = (s + 1) % self.DEPTH
        self.done[s].synchronize()               # slot s's previous batch (DEPTH submits ago) is complete
        if host_items:
            self._ensure_slab(s, total)
            slab_np = self.slabs_np[s]
            for i in host_items:
                slab_np[offs[i]:offs[i] + images[i].size] = images[i].reshape(-1)
        torch = self.torch
        if not jpeg_items:
            with torch.cuda.device(self.device):
                self.replica.predict_host(self.slabs[s], total, descs, n, self.topk, self.probs[s], self.idx[s],
                                          stream=self.stream)
                self.done[s].record(self.stream)
            return (s, n, ())
        # At least one file is decoded on the GPU: the batch slab lives in HBM.  Host-decoded images are copied
        # into it, JPEG files are decoded into it (K-J1..J3), then the device-resident predict chain runs.
        with torch.cuda.device(self.device), torch.cuda.stream(self.stream):
            self._ensure_jpeg(s, total)
            dev_slab = self.dev_slab[s]
            if host_items:
                lo = min(offs[i] for i in host_items)
                hi = max(offs[i] + shapes[i][0] * shapes[i][1] * 3 for i in host_items)
                dev_slab[lo:hi].copy_(self.slabs[s][lo:hi], non_blocking=True)
            self.jpeg.decode([images[i].data for i in jpeg_items], dev_slab, dev_slab.numel(),
                             [offs[i] for i in jpeg_items], status=self.dev_status[s], stream=self.stream)
            self.replica.predict(dev_slab, descs, n, self.topk, self.dev_probs[s], self.dev_idx[s], stream=self.stream)
            self.probs[s][:n].copy_(self.dev_probs[s][:n], non_blocking=True)
            self.idx[s][:n].copy_(self.dev_idx[s][:n], non_blocking=True)
            self.status_host[s][:len(jpeg_items)].copy_(self.dev_status[s][:len(jpeg_items)], non_blocking=True)
            self.done[s].record(self.stream)
        return (s, n, tuple(jpeg_items))

    def collect(self, ticket):
        """Waits for a submitted batch; returns (probs[n,k], idx[n,k], {image index: exception})."""
        s, n, jpeg_items = ticket
        self.done[s].synchronize()
        errors = {}
        if jpeg_items:
            st = self.status_host[s][:len(jpeg_items)].numpy()
            for j, i in enumerate(jpeg_items):
                if st[j]:
                    errors[i] = CorruptImage(f"JPEG entropy-coded data is damaged (decoder status {int(st[j])})")
        return self.probs[s][:n].numpy().copy(), self.idx[s][:n].numpy().copy(), errors

    def run_batch(self, images: Sequence[np.ndarray]):
        """Synchronous form: one batch, results on return."""
        return self.collect(self.submit_batch(images))

    def close(self) -> None:
        for e in self.done:
            e.synchronize()
        if self.jpeg is not None:
            self.jpeg.close()
        self.replica.close()


class Engine:
    """Micro-batching queue + least-loaded dispatch over replica workers.

    Workers are duck-typed.  The minimum is `run_batch(images) -> (probs, idx[, errors])`; a worker that also has
    `submit_batch(images) -> ticket` / `collect(ticket)` (GpuReplicaWorker) is driven with up to `worker.DEPTH`
    batches in flight."""

    def __init__(self, workers: Sequence, max_batch: int = 64, max_wait_us: int = 2000):
        if not workers:
            raise ValueError("at least one replica worker is required")
        self.workers = list(workers)
        self.max_batch = int(max_batch)
        self.max_wait_s = max_wait_us * 1e-6
        self._inbox: "queue.Queue[Optional[_Request]]" = queue.Queue()
        self._worker_q: List["queue.Queue[Optional[List[_Request]]]"] = [queue.Queue(maxsize=4) for _ in self.workers]
        self._rr = 0
        self._closed = False
        self._load = [0] * len(self.workers)       # images handed to a replica and not yet answered
        self._load_lock = threading.Lock()
        self.batch_sizes: List[int] = []          # achieved batch-size histogram source
        self.dispatch_log: List[Tuple[int, int]] = []   # (worker index, batch size)
        self._threads = [threading.Thread(target=self._batcher, name="b2p-batcher", daemon=True)]
        self._threads += [threading.Thread(target=self._worker_loop, args=(i,), name=f"b2p-gpu{i}", daemon=True)
                          for i in range(len(self.workers))]
        for t in self._threads:
            t.start()

    # ---- client side
    def submit(self, image) -> Future:
        if self._closed:
            raise RuntimeError("engine is shut down")
        fut: Future = Future()
        self._inbox.put(_Request(image, fut))
        return fut

    def submit_batch(self, images) -> Future:
        """A whole batch (a list of images / JpegFiles, or a PackedBatch) that bypasses the micro-batcher and goes
        straight to the least-loaded replica.  The future resolves to (probs[n,k], idx[n,k], {index: exception})."""
        if self._closed:
            raise RuntimeError("engine is shut down")
        fut: Future = Future()
        n = len(images)
        w = self._pick_worker(n)
        self.dispatch_log.append((w, n))
        self._worker_q[w].put(_BatchRequest(images, fut, n))
        return fut

    def predict_many(self, images: Sequence[np.ndarray]):
        futs = [self.submit(im) for im in images]
        return [f.result() for f in futs]

    def _pick_worker(self, n: int) -> int:
        """The replica with the fewest images in flight; while they are level the choice simply rotates, so an idle
        engine hands whole batches round-robin (north_star) and one slow GPU no longer stalls dispatch to the rest."""
        nw = len(self.workers)
        with self._load_lock:
            w = min(range(nw), key=lambda i: (self._load[(self._rr + i) % nw], i))
            w = (self._rr + w) % nw
            self._load[w] += n
        self._rr = (w + 1) % nw
        return w

    # ---- batcher thread
    def _batcher(self) -> None:
        while True:
            first = self._inbox.get()
            if first is None:
                break
            batch = [first]
            deadline = time.perf_counter() + self.max_wait_s
            stop = False
            while len(batch) < self.max_batch:
                timeout = deadline - time.perf_counter()
                try:
                    nxt = self._inbox.get(timeout=max(timeout, 0)) if timeout > 0 else self._inbox.get_nowait()
                except queue.Empty:
                    break
                if nxt is None:
                    stop = True
                    break
                batch.append(nxt)
            w = self._pick_worker(len(batch))
            self.batch_sizes.append(len(batch))
            self.dispatch_log.append((w, len(batch)))
            self._worker_q[w].put(batch)
            if stop:
                break
        for q in self._worker_q:
            q.put(None)

    # ---- one thread per replica
    def _answer(self, i: int, batch, out=None, exc: Optional[BaseException] = None) -> None:
        if isinstance(batch, _BatchRequest):
            try:
                if not batch.future.done():
                    if exc is not None:
                        batch.future.set_exception(exc)
                    else:
                        batch.future.set_result((out[0], out[1], out[2] if len(out) > 2 else {}))
            finally:
                with self._load_lock:
                    self._load[i] -= batch.n
            return
        try:
            if exc is None:
                probs, idx = out[0], out[1]
                errors = out[2] if len(out) > 2 else {}      # per-image failures do not fail the whole batch
            for j, r in enumerate(batch):
                if r.future.done():
                    continue
                if exc is not None:
                    r.future.set_exception(exc)      # propagate to every waiter: predict() re-raises -> HTTP 500
                elif j in errors:
                    r.future.set_exception(errors[j])
                else:
                    r.future.set_result((probs[j], idx[j]))
        finally:
            with self._load_lock:
                self._load[i] -= len(batch)

    def _worker_loop(self, i: int) -> None:
        worker = self.workers[i]
        pipelined = hasattr(worker, "submit_batch") and hasattr(worker, "collect")
        depth = int(getattr(worker, "DEPTH", 1)) if pipelined else 1
        inflight: List[Tuple[List[_Request], object]] = []       # (batch, ticket), oldest first
        q = self._worker_q[i]
        stopping = False

        def finish_oldest():
            batch, ticket = inflight.pop(0)
            try:
                self._answer(i, batch, out=worker.collect(ticket))
            except BaseException as e:
                self._answer(i, batch, exc=e)

        while not (stopping and not inflight):
            batch = None
            if not stopping:
                try:
                    # with work on the GPU only look at the queue; with none, sleep on it
                    batch = q.get_nowait() if inflight else q.get()
                    if batch is None:
                        stopping = True
                except queue.Empty:
                    batch = None
            if isinstance(batch, _BatchRequest):
                if not batch.future.set_running_or_notify_cancel():
                    with self._load_lock:
                        self._load[i] -= batch.n
                else:
                    try:
                        if pipelined:
                            inflight.append((batch, worker.submit_batch(batch.images)))
                        else:
                            self._answer(i, batch, out=worker.run_batch(batch.images))
                    except BaseException as e:
                        self._answer(i, batch, exc=e)
                if len(inflight) < depth and not stopping:
                    continue
            elif batch:
                n_before = len(batch)
                # A waiter may have given up while its request was queued (client disconnect, timeout, shutdown:
                # asyncio.wrap_future cancels the future).  Mark the rest running so they can no longer be cancelled
                # under us, and drop the cancelled ones instead of letting set_result() raise for the whole batch.
                batch = [r for r in batch if r.future.set_running_or_notify_cancel()]
                if n_before != len(batch):
                    with self._load_lock:
                        self._load[i] -= n_before - len(batch)
                if batch:
                    try:
                        if pipelined:
                            inflight.append((batch, worker.submit_batch([r.image for r in batch])))
                        else:
                            self._answer(i, batch, out=worker.run_batch([r.image for r in batch]))
                    except BaseException as e:
                        self._answer(i, batch, exc=e)
                if len(inflight) < depth and not stopping:
                    continue                      # room for one more before waiting on the oldest
            if inflight:
                finish_oldest()

    def close(self) -> None:
        """Drains the queue (queued requests are still served), stops the threads, frees the replicas."""
        if self._closed:
            return
        self._closed = True
        self._inbox.put(None)
        for t in self._threads:
            t.join(timeout=30)
        for w in self.workers:
            close = getattr(w, "close", None)
            if close:
                close()


def seeded_model(arch: str, seed: int):
    """Random-init torchvision classifier, fully determined by (arch, seed).  torchvision zero-initialises
    the ViT head (SP/torchvision/models/vision_transformer.py:264-266), which would make every logit 0;
    it is re-drawn from N(0, 0.02) with a fixed generator so that a random-init ViT has a usable output."""
    import torch
    import torchvision
    torch.manual_seed(seed)
    model = getattr(torchvision.models, arch)(weights=None).eval()
    if arch.startswith("vit"):
        g = torch.Generator().manual_seed(123)
        with torch.no_grad():
            model.heads.head.weight.normal_(0.0, 0.02, generator=g)
    return model


def build_gpu_engine(arch: str, gpus: Optional[Sequence[int]], max_batch: int, max_wait_us: int, topk: int,
                     mode: str, seed: int, sub_batch: int = 0, state_dict: Optional[dict] = None,
                     packed_path: Optional[str] = None):
    """What init() does: a packed program file, or a (seeded random-init / checkpointed) torchvision model turned
    into a layer program, -> one replica per GPU.  Returns (engine, number of classes); the class names of a packed
    file are left on `engine.classes`."""
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device visible: the B200 predict path has no CPU fallback")
    classes = None
    if packed_path:
        from .packed import load_program
        prog, classes = load_program(packed_path)
    else:
        import torchvision  # noqa: F401
        from . import program as P
        model = seeded_model(arch, seed)
        if state_dict is not None:
            model.load_state_dict(state_dict)
        prog = P.build_program(model, arch)
    devs = list(gpus) if gpus else list(range(torch.cuda.device_count()))
    workers = [GpuReplicaWorker(d, prog, max_batch, topk, mode, sub_batch) for d in devs]
    engine = Engine(workers, max_batch=max_batch, max_wait_us=max_wait_us)
    engine.classes = classes
    return engine, prog.out_elems
