"""ctypes binding of libb200predict.so (the C ABI in include/b200predict.h).

This module is deliberately thin: it mirrors the structs, loads the in-tree shared library and
raises if it is missing or a call fails.  There is no Python/torch fallback for any entry point —
a missing library or a non-sm_100 device is an error, never a slow path.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional, Sequence

PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("B2P_LIB", PKG / "lib" / "libb200predict.so"))   # B2P_LIB: A/B-test another build

ABI_VERSION = 3
SLOT_INPUT, SLOT_OUTPUT, SLOT_NONE = 0, -2, -1
OP_CONV, OP_LINEAR, OP_MAXPOOL, OP_AVGPOOL, OP_STEM_PACK, OP_NCHW_TO_NHWC, OP_STEM_CONV = 1, 2, 3, 4, 5, 6, 7
OP_PATCHIFY, OP_TOKENS, OP_LAYERNORM, OP_ATTENTION, OP_CONV_PAIR, OP_BOTTLENECK = 8, 9, 10, 11, 12, 13
ACT_NONE, ACT_RELU, ACT_GELU = 0, 1, 2
OUT_HW = 224
MAX_TOPK = 8

ERRORS = {0: "OK", -1: "B2P_E_ARG", -2: "B2P_E_DEVICE", -3: "B2P_E_CUDA", -4: "B2P_E_NOMEM", -5: "B2P_E_LIMIT", -6: "B2P_E_UNSUPPORTED"}


class B2PError(RuntimeError):
    def __init__(self, code: int, detail: str):
        super().__init__(f"{ERRORS.get(code, code)}: {detail}")
        self.code = code
        self.detail = detail


class Op(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("src", C.c_int32), ("dst", C.c_int32), ("res", C.c_int32),
        ("in_h", C.c_int32), ("in_w", C.c_int32), ("in_c", C.c_int32),
        ("out_h", C.c_int32), ("out_w", C.c_int32), ("out_c", C.c_int32),
        ("kh", C.c_int32), ("kw", C.c_int32), ("stride", C.c_int32),
        ("pad_lo", C.c_int32), ("pad_hi", C.c_int32), ("relu", C.c_int32),
        ("weight", C.c_int32), ("bias", C.c_int32), ("reserved", C.c_int32 * 4),
        ("src2", C.c_int32), ("src2_c", C.c_int32), ("src2_stride", C.c_int32), ("reserved2", C.c_int32),
    ]


class ModelDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("max_batch", C.c_int32), ("sub_batch", C.c_int32),
        ("n_ops", C.c_int32), ("ops", C.POINTER(Op)),
        ("n_slots", C.c_int32), ("slot_elems", C.POINTER(C.c_int64)),
        ("n_blobs", C.c_int32), ("blob_ptrs", C.POINTER(C.c_void_p)), ("blob_bytes", C.POINTER(C.c_int64)),
        ("in_elems_per_image", C.c_int64), ("out_elems_per_image", C.c_int64),
    ]


class Image(C.Structure):
    _fields_ = [
        ("offset", C.c_int64), ("h", C.c_int32), ("w", C.c_int32),
        ("res_h", C.c_int32), ("res_w", C.c_int32), ("top", C.c_int32), ("left", C.c_int32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("ms_h2d", C.c_float), ("ms_preprocess", C.c_float), ("ms_forward", C.c_float),
        ("ms_epilogue", C.c_float), ("ms_d2h", C.c_float),
        ("calls", C.c_int64), ("kernel_launches", C.c_int64),
    ]


class JpegInfo(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("components", C.c_int32),
        ("h_samp", C.c_int32), ("v_samp", C.c_int32), ("restart_interval", C.c_int32),
        ("segments", C.c_int32), ("reserved", C.c_int32), ("scan_bytes", C.c_int64),
    ]


class JpegStats(C.Structure):
    _fields_ = [("ms_copy", C.c_float), ("ms_huffman", C.c_float), ("ms_idct", C.c_float), ("ms_colour", C.c_float),
                ("calls", C.c_int64)]


E_UNSUPPORTED = -6

# every symbol include/b200predict.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("b2p_abi_version", C.c_int32, []),
    ("b2p_build_info", C.c_char_p, []),
    ("b2p_create", C.c_int32, [C.c_int32, C.POINTER(ModelDesc), C.POINTER(C.c_void_p)]),
    ("b2p_destroy", C.c_int32, [C.c_void_p]),
    ("b2p_last_error", C.c_char_p, [C.c_void_p]),
    ("b2p_stats_get", C.c_int32, [C.c_void_p, C.POINTER(Stats)]),
    ("b2p_preprocess", C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(Image), C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]),
    ("b2p_forward", C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    ("b2p_softmax_topk", C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("b2p_predict", C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(Image), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("b2p_predict_host", C.c_int32, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(Image), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("b2p_last_logits", C.c_void_p, [C.c_void_p]),
    ("b2p_head_static", C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p,
                                    C.c_int32, C.c_void_p, C.c_void_p]),
    ("b2p_jpeg_inspect", C.c_int32, [C.c_char_p, C.c_int64, C.POINTER(JpegInfo)]),
    ("b2p_jpeg_create", C.c_int32, [C.c_int32, C.POINTER(C.c_void_p)]),
    ("b2p_jpeg_destroy", C.c_int32, [C.c_void_p]),
    ("b2p_jpeg_last_error", C.c_char_p, [C.c_void_p]),
    ("b2p_jpeg_stats_get", C.c_int32, [C.c_void_p, C.POINTER(JpegStats)]),
    ("b2p_jpeg_decode", C.c_int32, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.c_int32, C.c_void_p,
                                    C.c_int64, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]),
]

HEAD_STATIC_MAX_IMAGES = 256
HEAD_STATIC_MAX_CHANNELS = 2048

_lib: Optional[C.CDLL] = None


def load_library(path: Optional[os.PathLike] = None) -> C.CDLL:
    """Loads libb200predict.so and binds every declared symbol.  Raises if anything is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise RuntimeError(
            f"{p} is missing: build it with `python -m mlmicroservicetemplate_b200.build` "
            "(there is no fallback implementation of the predict path)")
    lib = C.CDLL(str(p))
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.b2p_abi_version() != ABI_VERSION:
        raise RuntimeError(f"ABI mismatch: library {lib.b2p_abi_version()} != binding {ABI_VERSION}")
    if path is None:
        _lib = lib
    return lib


def make_images(geoms: Sequence[tuple]) -> "C.Array[Image]":
    """geoms: (offset, h, w, res_h, res_w, top, left) per image."""
    arr = (Image * len(geoms))()
    for i, g in enumerate(geoms):
        arr[i] = Image(*[int(v) for v in g])
    return arr


def head_static(feat, n: int, hw: int, c: int, weight, weight_rows: int, bias, classes: int, logits, stream=None) -> None:
    """Allocation-free classifier head (average pool + linear) on the current device; see the header.
    All tensor arguments are device tensors (or raw device addresses) owned by the caller."""
    lib = load_library()
    rc = lib.b2p_head_static(Replica._ptr(feat), n, hw, c, Replica._ptr(weight), weight_rows, Replica._ptr(bias),
                             classes, Replica._ptr(logits), Replica._stream(stream))
    if rc != 0:
        raise B2PError(rc, (lib.b2p_last_error(None) or b"").decode())


def jpeg_inspect(data: bytes) -> Optional[JpegInfo]:
    """Header walk of one file on the host (no device needed).  Returns the geometry when the GPU decoder covers
    the file, None when it is a fine image for the host decoder only (progressive, CMYK, not a JPEG at all ...)."""
    lib = load_library()
    info = JpegInfo()
    rc = lib.b2p_jpeg_inspect(data, len(data), C.byref(info))
    return info if rc == 0 else None


class JpegDecoder:
    """GPU baseline-JPEG decoder bound to one device (b2p_jpeg_* in the header).  One thread, one stream."""

    def __init__(self, device: int):
        self.lib = load_library()
        self.device = int(device)
        h = C.c_void_p()
        rc = self.lib.b2p_jpeg_create(self.device, C.byref(h))
        if rc != 0:
            raise B2PError(rc, (self.lib.b2p_jpeg_last_error(None) or b"").decode())
        self.h = h

    def decode(self, files: Sequence[bytes], slab, slab_bytes: int, offsets: Sequence[int], status=None, stream=None) -> None:
        """files -> uint8 HWC RGB images at slab + offsets[i] (device memory).  Asynchronous on `stream`."""
        n = len(files)
        # a file is a bytes object, or anything with .address / .length (ingest.HostView: bytes that live in a
        # shared-memory ring and are read in place); the library has consumed the host bytes when it returns
        arr = (C.c_void_p * max(n, 1))()
        keep = []
        for i, f in enumerate(files):
            if hasattr(f, "address"):
                arr[i] = int(f.address)
            else:
                b = f if isinstance(f, bytes) else bytes(f)
                keep.append(b)
                arr[i] = C.cast(C.c_char_p(b), C.c_void_p).value
        lens = (C.c_int64 * max(n, 1))(*[len(f) for f in files])
        offs = (C.c_int64 * max(n, 1))(*[int(o) for o in offsets])
        rc = self.lib.b2p_jpeg_decode(self.h, arr, lens, n, Replica._ptr(slab), int(slab_bytes), offs,
                                      Replica._ptr(status) if status is not None else None, Replica._stream(stream))
        if rc != 0:
            raise B2PError(rc, (self.lib.b2p_jpeg_last_error(self.h) or b"").decode())

    def stats(self) -> JpegStats:
        st = JpegStats()
        rc = self.lib.b2p_jpeg_stats_get(self.h, C.byref(st))
        if rc != 0:
            raise B2PError(rc, "b2p_jpeg_stats_get")
        return st

    def close(self) -> None:
        if self.h:
            self.lib.b2p_jpeg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Replica:
    """One model replica on one GPU (what init() creates per device)."""

    def __init__(self, device: int, program, max_batch: int, sub_batch: int = 0):
        self.lib = load_library()
        self.device = int(device)
        self.program = program
        self.max_batch = int(max_batch)
        ops = (Op * len(program.ops))(*program.ops)
        slot_elems = (C.c_int64 * len(program.slot_elems))(*program.slot_elems)
        nb = len(program.blobs)
        blob_ptrs = (C.c_void_p * max(nb, 1))(*[b.data_ptr() for b in program.blobs])
        blob_bytes = (C.c_int64 * max(nb, 1))(*[b.numel() * b.element_size() for b in program.blobs])
        desc = ModelDesc(ABI_VERSION, self.max_batch, int(sub_batch), len(program.ops), ops,
                         len(program.slot_elems), slot_elems, nb, blob_ptrs, blob_bytes,
                         int(program.in_elems), int(program.out_elems))
        h = C.c_void_p()
        rc = self.lib.b2p_create(self.device, C.byref(desc), C.byref(h))
        if rc != 0:
            raise B2PError(rc, (self.lib.b2p_last_error(None) or b"").decode())
        self.h = h
        self.classes = int(program.out_elems)

    # -- helpers
    def _check(self, rc: int) -> None:
        if rc != 0:
            raise B2PError(rc, (self.lib.b2p_last_error(self.h) or b"").decode())

    @staticmethod
    def _stream(stream) -> C.c_void_p:
        if stream is None:
            import torch
            stream = torch.cuda.current_stream()
        return C.c_void_p(int(getattr(stream, "cuda_stream", stream)))

    def close(self) -> None:
        if getattr(self, "h", None):
            self.lib.b2p_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- entry points (device pointers are ints / objects with data_ptr())
    @staticmethod
    def _ptr(t) -> C.c_void_p:
        return C.c_void_p(int(t.data_ptr() if hasattr(t, "data_ptr") else t))

    def preprocess(self, slab, images, n: int, out, out_f32: bool = False, stream=None) -> None:
        self._check(self.lib.b2p_preprocess(self.h, self._ptr(slab), images, n, self._ptr(out), int(out_f32),
                                            self._stream(stream)))

    def forward(self, x, n: int, out, stream=None) -> None:
        self._check(self.lib.b2p_forward(self.h, self._ptr(x), n, self._ptr(out), self._stream(stream)))

    def softmax_topk(self, logits, n: int, classes: int, k: int, probs, idx, stream=None) -> None:
        self._check(self.lib.b2p_softmax_topk(self.h, self._ptr(logits), n, classes, k, self._ptr(probs),
                                              self._ptr(idx), self._stream(stream)))

    def predict(self, slab, images, n: int, k: int, probs, idx, stream=None) -> None:
        self._check(self.lib.b2p_predict(self.h, self._ptr(slab), images, n, k, self._ptr(probs), self._ptr(idx),
                                         self._stream(stream)))

    def predict_host(self, slab_host, slab_bytes: int, images, n: int, k: int, probs_host, idx_host,
                     stream=None) -> None:
        self._check(self.lib.b2p_predict_host(self.h, self._ptr(slab_host), int(slab_bytes), images, n, k,
                                              self._ptr(probs_host), self._ptr(idx_host), self._stream(stream)))

    def last_logits_ptr(self) -> int:
        return int(self.lib.b2p_last_logits(self.h) or 0)

    def stats(self) -> dict:
        s = Stats()
        self._check(self.lib.b2p_stats_get(self.h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in Stats._fields_}
