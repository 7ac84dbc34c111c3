"""Model plugin — the drop-in for CodyRichter/MLMicroserviceTemplate src/model/model.py.

Same two module-level functions, same signatures, same result-dict shape:
    init() -> None                                             (reference model.py:6-13)
    predict(image_file: UploadFile = File(...)) -> dict        (reference model.py:16-31)
returning {'classes': [every class name], 'result': {class_name: value}}.

What changed underneath: init() builds one model replica per visible B200 through the C ABI
(libb200predict.so) plus the micro-batching engine; predict() decodes the upload on the host with
Pillow (exactly the reference's `Image.open(image_file.file)`, now followed by .convert('RGB')),
enqueues the uint8 HWC pixels and blocks on a future, so concurrent callers are batched.
`predict_async` is the awaitable twin the server uses so the event loop is never blocked.

Environment knobs (defaults in brackets):
    MODEL [resnet50]    resnet18 | resnet50 | vit_b_16 | template   (`template` = the reference's constant-dict
                        example plugin, no GPU involved — BASELINE.json configs[0])
    GPUS  [all]         comma-separated device indices
    MAX_BATCH [64]  MAX_WAIT_US [2000]  TOPK [5]  SEED [0]  PREPROCESS [preset] (preset | direct)
    SUB_BATCH [0]       images per L2-resident pass of the layer program (0 = whole batch)
    WEIGHTS             path of a packed program (`python -m mlmicroservicetemplate_b200.packed`): folded bf16 weights in
                        kernel layout + class names, loaded as-is (no torchvision at start-up); MODEL is then ignored
    CHECKPOINT          path of a torch.save'd torchvision state_dict for MODEL (default: seeded random init)
    LABELS              text file with one class name per line (default: positional names)
    ENGINE_SOCKET       path of a GPU dispatcher's Unix socket (src/server/serve.py starts both): this process then
                        builds no replica; uploads go to the dispatcher through a shared-memory ring of INGEST_SHM_MB [64]
    DECODE [gpu]        gpu: uploads that are baseline JPEGs (b2p_jpeg_inspect) are decoded on the GPU, bit-exact
                        with Pillow; everything else (PNG, progressive / CMYK JPEG, ...) is decoded by Pillow
                        as in the reference.  host: every upload goes through Pillow.
"""
from __future__ import annotations

import asyncio
import os
import threading
from typing import List, Optional

import numpy as np
from fastapi import File, UploadFile
from PIL import Image

_lock = threading.Lock()
_state = {"engine": None, "classes": None, "arch": None, "decode_pool": None, "decode": "host"}

TEMPLATE_RESULT = {
    'classes': ['isGreen', 'isRed'],
    'result': {'isGreen': 0, 'isRed': 1},
}


def class_names(n: int) -> List[str]:
    # Random-init weights have no label file; names are positional (checkpoint + labels: SURVEY §8f row 3).
    return [f"class_{i:04d}" for i in range(n)]


def init() -> None:
    """Runs once at start-up (reference src/server/main.py:77-84 calls it on a pool thread)."""
    with _lock:
        if _state["engine"] is not None or _state["arch"] == "template":
            return
        arch = os.getenv("MODEL", "resnet50")
        _state["arch"] = arch
        if arch == "template":
            return
        from concurrent.futures import ThreadPoolExecutor
        if os.getenv("ENGINE_SOCKET"):
            # Front-end process of a multi-process deployment (ingest.py): the replicas live in
            # the dispatcher process; this process parses uploads and hands their bytes over through shared memory.
            from ...ingest import RemoteEngine
            engine = RemoteEngine(os.environ["ENGINE_SOCKET"], ring_bytes=int(os.getenv("INGEST_SHM_MB", "64")) << 20)
            _state["engine"] = engine
            _state["decode"] = engine.info.get("decode", "gpu")
            _state["classes"] = list(engine.info["classes"])
            _state["decode_pool"] = ThreadPoolExecutor(max_workers=int(os.getenv("DECODE_THREADS", "4")))
            return
        from ...runtime import build_gpu_engine
        gpus = os.getenv("GPUS")
        devs = [int(g) for g in gpus.split(",")] if gpus else None
        state_dict = None
        if os.getenv("CHECKPOINT") and not os.getenv("WEIGHTS"):
            import torch
            sd = torch.load(os.environ["CHECKPOINT"], map_location="cpu", weights_only=True)
            state_dict = sd.get("state_dict", sd) if isinstance(sd, dict) else sd
        engine, n_classes = build_gpu_engine(
            arch=arch, gpus=devs, max_batch=int(os.getenv("MAX_BATCH", "64")),
            max_wait_us=int(os.getenv("MAX_WAIT_US", "2000")), topk=int(os.getenv("TOPK", "5")),
            mode=os.getenv("PREPROCESS", "preset"), seed=int(os.getenv("SEED", "0")),
            sub_batch=int(os.getenv("SUB_BATCH", "0")), state_dict=state_dict, packed_path=os.getenv("WEIGHTS"))
        _state["engine"] = engine
        _state["decode"] = os.getenv("DECODE", "gpu")
        names = getattr(engine, "classes", None)
        if os.getenv("LABELS"):
            with open(os.environ["LABELS"], encoding="utf-8") as f:
                names = [ln.rstrip("\n") for ln in f if ln.strip()]
        if names is not None and len(names) != n_classes:
            raise RuntimeError(f"{len(names)} class names for a model with {n_classes} outputs")
        _state["classes"] = list(names) if names is not None else class_names(n_classes)
        _state["decode_pool"] = ThreadPoolExecutor(max_workers=int(os.getenv("DECODE_THREADS", str(os.cpu_count() or 4))))


def shutdown() -> None:
    """Drains the batch queue and destroys the replicas (the server's shutdown hook calls this)."""
    with _lock:
        eng, _state["engine"] = _state["engine"], None
        pool, _state["decode_pool"] = _state["decode_pool"], None
        _state["arch"] = None
        _state["decode"] = "host"
    if eng is not None:
        eng.close()
    if pool is not None:
        pool.shutdown(wait=False)


def _decode(fileobj):
    """Stage D.  With DECODE=gpu a baseline JPEG is only header-checked here and handed on as bytes (the GPU
    decodes it); any other PIL-readable file is decoded on the host to uint8 HWC RGB, as the reference would."""
    if _state.get("decode") == "gpu":
        from ... import backend
        from ...runtime import JpegFile
        data = fileobj.read()
        info = backend.jpeg_inspect(data)
        if info is not None:
            return JpegFile(data, info.height, info.width)
        import io
        fileobj = io.BytesIO(data)
    image = Image.open(fileobj)
    return np.asarray(image.convert('RGB'))


def _to_result(probs, idx) -> dict:
    names = _state["classes"]
    return {
        'classes': names,
        'result': {names[int(i)]: float(p) for p, i in zip(probs, idx)},
    }


def _engine():
    eng = _state["engine"]
    if eng is None:
        raise RuntimeError("model is not initialised: call init() first")
    return eng


def predict(image_file: UploadFile = File(...)):
    """
    Interface method between model and server. This signature must not be
    changed and your model must be able to predict given a file-like object
    with the image as an input.
    """
    if _state["arch"] == "template":
        Image.open(image_file.file)              # what the reference template does (model.py:23)
        return {'classes': list(TEMPLATE_RESULT['classes']), 'result': dict(TEMPLATE_RESULT['result'])}
    arr = _decode(image_file.file)
    probs, idx = _engine().submit(arr).result()
    return _to_result(probs, idx)


async def predict_async(image_file: UploadFile = File(...)):
    """Awaitable twin of predict(): decode on a worker thread, await the batched result."""
    if _state["arch"] == "template":
        return predict(image_file)
    loop = asyncio.get_running_loop()
    arr = await loop.run_in_executor(_state["decode_pool"], _decode, image_file.file)
    probs, idx = await asyncio.wrap_future(_engine().submit(arr))
    return _to_result(probs, idx)


def _dumps(obj) -> bytes:
    # exactly what starlette's JSONResponse.render does, so the bytes on the wire do not change
    import json
    return json.dumps(obj, ensure_ascii=False, allow_nan=False, indent=None, separators=(",", ":")).encode("utf-8")


async def predict_async_json(image_file: UploadFile = File(...)) -> bytes:
    """predict_async() already rendered as the JSON text of its result dict.  The 'classes' list (every class name,
    identical in every answer) is rendered once per model instead of being walked by the framework's encoder on every
    request — that walk was the largest per-request cost on the event loop."""
    if _state["arch"] == "template":
        return _dumps(predict(image_file))
    loop = asyncio.get_running_loop()
    arr = await loop.run_in_executor(_state["decode_pool"], _decode, image_file.file)
    probs, idx = await asyncio.wrap_future(_engine().submit(arr))
    names = _state["classes"]
    cached = _state.get("classes_json")
    if cached is None or cached[0] is not names:
        cached = (names, _dumps(names))
        _state["classes_json"] = cached
    result = {names[int(i)]: float(p) for p, i in zip(probs, idx)}
    return b'{"classes":' + cached[1] + b',"result":' + _dumps(result) + b'}'
