"""On-disk form of a layer program: what `init()` loads instead of re-deriving weights.

The reference's `init()` docstring (src/model/model.py:7-10; README.md:103-104) says the function exists to
"fetch / check model files".  A `.b2pw` file is exactly what `b2p_create` consumes, so start-up is one read: the
flat op list (`b2p_op` structs as laid out in include/b200predict.h), the slot sizes, and every weight blob
already BN-folded, bf16-rounded, padded and in the kernels' native layouts — plus the class names for the result dict.

    python -m mlmicroservicetemplate_b200.packed --arch resnet50 [--state-dict ckpt.pt] [--labels labels.txt] --out resnet50.b2pw

File layout (little endian):
    magic  "B2PW"            4 B
    version                  u32   (1)
    abi_version              u32   (must equal the library's B2P_ABI_VERSION: the op struct is part of the ABI)
    header_len               u64
    header                   JSON, utf-8: name, in_elems, out_elems, flops_per_image, slot_elems, n_ops,
                                          blobs: [{dtype, shape, offset, nbytes}], classes: [...] | null, sha256 of the payload
    ops                      n_ops x sizeof(b2p_op) raw bytes
    payload                  blobs, each 256-byte aligned (offsets relative to the payload start)
"""
from __future__ import annotations

import ctypes as C
import hashlib
import json
import os
import struct
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import backend as B
from .program import Program

MAGIC = b"B2PW"
VERSION = 1
_ALIGN = 256
_DTYPES = {"bfloat16": torch.bfloat16, "float32": torch.float32, "float16": torch.float16, "int32": torch.int32,
           "uint8": torch.uint8, "int16": torch.int16}


def _blob_bytes(t: torch.Tensor) -> bytes:
    t = t.detach().contiguous().cpu()
    return t.view(torch.uint8).reshape(-1).numpy().tobytes() if t.numel() else b""


def save_program(prog: Program, path: str, classes: Optional[Sequence[str]] = None) -> dict:
    """Writes `prog` to `path` and returns the header that was written."""
    blobs_meta, payload, pos = [], [], 0
    for t in prog.blobs:
        raw = _blob_bytes(t)
        pad = (-pos) % _ALIGN
        payload.append(b"\0" * pad)
        pos += pad
        blobs_meta.append({"dtype": str(t.dtype).replace("torch.", ""), "shape": list(t.shape), "offset": pos,
                           "nbytes": len(raw)})
        payload.append(raw)
        pos += len(raw)
    payload_bytes = b"".join(payload)
    if classes is not None and len(classes) != int(prog.out_elems):
        raise ValueError(f"{len(classes)} class names for a model with {prog.out_elems} outputs")
    header = {
        "name": prog.name, "in_elems": int(prog.in_elems), "out_elems": int(prog.out_elems),
        "flops_per_image": float(prog.flops_per_image), "slot_elems": [int(v) for v in prog.slot_elems],
        "n_ops": len(prog.ops), "op_bytes": C.sizeof(B.Op), "blobs": blobs_meta,
        "classes": list(classes) if classes is not None else None,
        "payload_sha256": hashlib.sha256(payload_bytes).hexdigest(),
    }
    hjson = json.dumps(header).encode()
    ops_raw = b"".join(bytes(op) for op in prog.ops)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<IIQ", VERSION, B.ABI_VERSION, len(hjson)))
        f.write(hjson)
        f.write(ops_raw)
        f.write(payload_bytes)
    os.replace(tmp, path)
    return header


def load_program(path: str, verify: bool = True):
    """Reads a `.b2pw` file.  Returns (Program, classes or None).  Raises ValueError on a damaged or foreign file —
    a model file that does not match the library must stop `init()`, not produce wrong answers."""
    with open(path, "rb") as f:
        data = f.read()
    if len(data) < 20 or data[:4] != MAGIC:
        raise ValueError(f"{path}: not a packed b200predict program (bad magic)")
    version, abi, hlen = struct.unpack_from("<IIQ", data, 4)
    if version != VERSION:
        raise ValueError(f"{path}: file format version {version}, this build reads {VERSION}")
    if abi != B.ABI_VERSION:
        raise ValueError(f"{path}: packed for ABI {abi}, the library is ABI {B.ABI_VERSION}: re-pack the checkpoint")
    pos = 20
    if pos + hlen > len(data):
        raise ValueError(f"{path}: truncated header")
    header = json.loads(data[pos:pos + hlen].decode())
    pos += hlen
    op_bytes = C.sizeof(B.Op)
    if header.get("op_bytes") != op_bytes:
        raise ValueError(f"{path}: op records are {header.get('op_bytes')} bytes, the library expects {op_bytes}")
    n_ops = int(header["n_ops"])
    if pos + n_ops * op_bytes > len(data):
        raise ValueError(f"{path}: truncated op list")
    ops = [B.Op.from_buffer_copy(data, pos + i * op_bytes) for i in range(n_ops)]
    pos += n_ops * op_bytes
    payload = memoryview(data)[pos:]
    if verify and hashlib.sha256(payload).hexdigest() != header["payload_sha256"]:
        raise ValueError(f"{path}: weight payload does not match its checksum")
    blobs: List[torch.Tensor] = []
    for m in header["blobs"]:
        dt = _DTYPES[m["dtype"]]
        raw = payload[m["offset"]:m["offset"] + m["nbytes"]]
        if len(raw) != m["nbytes"]:
            raise ValueError(f"{path}: truncated weight blob")
        t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).view(dt).reshape(m["shape"]) if m["nbytes"] else \
            torch.empty(m["shape"], dtype=dt)
        blobs.append(t)
    prog = Program(ops=ops, slot_elems=list(header["slot_elems"]), blobs=blobs, in_elems=header["in_elems"],
                   out_elems=header["out_elems"], flops_per_image=header["flops_per_image"], name=header["name"])
    return prog, header.get("classes")


def pack_checkpoint(arch: str, out: str, state_dict_path: Optional[str] = None, labels_path: Optional[str] = None,
                    seed: int = 0) -> dict:
    """torchvision module (+ optional real `state_dict`) -> BN-folded bf16 layer program -> `.b2pw` file."""
    from .program import build_program
    from .runtime import seeded_model
    model = seeded_model(arch, seed)
    if state_dict_path:
        sd = torch.load(state_dict_path, map_location="cpu", weights_only=True)
        model.load_state_dict(sd.get("state_dict", sd) if isinstance(sd, dict) else sd)
    prog = build_program(model, arch)
    classes = None
    if labels_path:
        with open(labels_path, encoding="utf-8") as f:
            classes = [ln.rstrip("\n") for ln in f if ln.strip()]
    return save_program(prog, out, classes)


def main(argv=None) -> int:
    import argparse
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--arch", required=True, choices=["resnet18", "resnet34", "resnet50", "resnet101", "vit_b_16"])
    ap.add_argument("--state-dict", help="torch.save'd state_dict of the torchvision module (default: seeded random init)")
    ap.add_argument("--labels", help="text file, one class name per line")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out", required=True)
    a = ap.parse_args(argv)
    h = pack_checkpoint(a.arch, a.out, a.state_dict, a.labels, a.seed)
    size = os.path.getsize(a.out)
    print(f"{a.out}: {h['name']}, {h['n_ops']} ops, {len(h['blobs'])} blobs, {size / 1e6:.1f} MB, "
          f"classes: {'yes' if h['classes'] else 'positional'}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
