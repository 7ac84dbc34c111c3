"""Host runtime under the model plugin: micro-batching queue, per-GPU replica workers and the
round-robin dispatcher.

The reference serves one request at a time on the event loop (src/server/main.py:135 calls
predict() synchronously).  Here predict() enqueues a decoded uint8 HWC image and waits on a future;
a batcher thread groups requests (up to MAX_BATCH, or whatever arrived within MAX_WAIT_US of the
first one) and hands whole batches round-robin to one worker thread per GPU.  A worker copies the
batch into its pinned host slab and makes ONE call into the C ABI (b2p_predict_host): H2D copy,
K1 preprocess, K2 forward, K3 softmax/top-k and the D2H copy of k (prob, index) pairs per image.

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
class _Request:
    image: np.ndarray
    future: Future
    t_enqueue: float = field(default_factory=time.perf_counter)


class GpuReplicaWorker:
    """One replica: a b2p handle on one GPU, its stream and a pinned host slab."""

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
            self.slab = torch.empty(slab_bytes, dtype=torch.uint8, pin_memory=True)
            self.probs = torch.empty((max_batch, topk), dtype=torch.float32, pin_memory=True)
            self.idx = torch.empty((max_batch, topk), dtype=torch.int32, pin_memory=True)
        self.slab_np = self.slab.numpy()

    def _ensure_slab(self, nbytes: int) -> None:
        if nbytes > self.slab.numel():
            self.stream.synchronize()
            self.slab = self.torch.empty(int(nbytes * 1.25), dtype=self.torch.uint8, pin_memory=True)
            self.slab_np = self.slab.numpy()

    def run_batch(self, images: Sequence[np.ndarray]):
        n = len(images)
        if n == 0 or n > self.max_batch:
            raise ValueError(f"batch of {n} outside 1..{self.max_batch}")
        offs, total = [], 0
        for im in images:
            if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] != 3:
                raise ValueError("images must be uint8 HWC with 3 channels")
            offs.append(total)
            total += (im.size + 255) // 256 * 256
        self._ensure_slab(total)
        geoms = []
        for im, o in zip(images, offs):
            self.slab_np[o:o + im.size] = im.reshape(-1)
            geoms.append((o, im.shape[0], im.shape[1], *image_geometry(im.shape[0], im.shape[1], self.mode)))
        descs = self.B.make_images(geoms)
        with self.torch.cuda.device(self.device):
            self.replica.predict_host(self.slab, total, descs, n, self.topk, self.probs, self.idx, stream=self.stream)
            self.stream.synchronize()
        return self.probs[:n].numpy().copy(), self.idx[:n].numpy().copy()

    def close(self) -> None:
        self.replica.close()


class Engine:
    """Micro-batching queue + round-robin dispatch over replica workers."""

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
        self.batch_sizes: List[int] = []          # achieved batch-size histogram source
        self.dispatch_log: List[Tuple[int, int]] = []   # (worker index, batch size)
        self._threads = [threading.Thread(target=self._batcher, name="b2p-batcher", daemon=True)]
        self._threads += [threading.Thread(target=self._worker_loop, args=(i,), name=f"b2p-gpu{i}", daemon=True)
                          for i in range(len(self.workers))]
        for t in self._threads:
            t.start()

    # ---- client side
    def submit(self, image: np.ndarray) -> Future:
        if self._closed:
            raise RuntimeError("engine is shut down")
        fut: Future = Future()
        self._inbox.put(_Request(image, fut))
        return fut

    def predict_many(self, images: Sequence[np.ndarray]):
        futs = [self.submit(im) for im in images]
        return [f.result() for f in futs]

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
            w = self._rr
            self._rr = (self._rr + 1) % len(self.workers)
            self.batch_sizes.append(len(batch))
            self.dispatch_log.append((w, len(batch)))
            self._worker_q[w].put(batch)
            if stop:
                break
        for q in self._worker_q:
            q.put(None)

    # ---- one thread per replica
    def _worker_loop(self, i: int) -> None:
        worker = self.workers[i]
        while True:
            batch = self._worker_q[i].get()
            if batch is None:
                break
            try:
                probs, idx = worker.run_batch([r.image for r in batch])
                for j, r in enumerate(batch):
                    r.future.set_result((probs[j], idx[j]))
            except BaseException as e:          # propagate to every waiter: predict() re-raises -> HTTP 500
                for r in batch:
                    if not r.future.done():
                        r.future.set_exception(e)

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
                     mode: str, seed: int, sub_batch: int = 0, state_dict: Optional[dict] = None):
    """What init() does: random-init (seeded) torchvision model -> layer program -> one replica per GPU."""
    import torch
    import torchvision
    from . import program as P
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device visible: the B200 predict path has no CPU fallback")
    model = seeded_model(arch, seed)
    if state_dict is not None:
        model.load_state_dict(state_dict)
    prog = P.build_program(model, arch)
    devs = list(gpus) if gpus else list(range(torch.cuda.device_count()))
    workers = [GpuReplicaWorker(d, prog, max_batch, topk, mode, sub_batch) for d in devs]
    return Engine(workers, max_batch=max_batch, max_wait_us=max_wait_us), prog.out_elems
