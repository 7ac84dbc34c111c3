"""Builds libb200predict.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m mlmicroservicetemplate_b200.build [--force] [--verbose]

The library links the CUDA runtime statically and resolves the two driver entry points it needs
(tensor-map encoders) at run time, so it loads on a machine with no driver installed; every compute
entry point then fails loudly instead of falling back to anything.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB_DIR = PKG / "lib"
LIB_PATH = LIB_DIR / "libb200predict.so"
SOURCES = ["api.cu", "bottleneck56.cu", "conv3x3_c128.cu", "conv3x3_c64.cu", "conv_gemm.cu", "conv_pair.cu", "head_static.cu", "jpeg.cu", "misc_kernels.cu", "preprocess.cu", "stem.cu", "vit.cu"]
HEADERS = [CSRC / "kernels.h", CSRC / "ptx.cuh", PKG.parent / "include" / "b200predict.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


PER_FILE_FLAGS = {}


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; libb200predict.so cannot be built")
    return cand


def _stale() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    deps = [CSRC / s for s in SOURCES] + HEADERS + [Path(__file__)]
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not _stale():
        return LIB_PATH
    LIB_DIR.mkdir(exist_ok=True)
    obj_dir = PKG / "build"
    obj_dir.mkdir(exist_ok=True)
    nvcc = _nvcc()
    procs = []
    objs = []
    for src in SOURCES:
        obj = obj_dir / (src.replace(".cu", ".o"))
        objs.append(str(obj))
        cmd = [nvcc, *NVCC_FLAGS, *PER_FILE_FLAGS.get(src, []), "-c", str(CSRC / src), "-o", str(obj)]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"==== {src}\n{out}")
        failed |= p.returncode != 0
    (obj_dir / "ptxas.log").write_text("\n".join(log))
    if failed:
        sys.stderr.write("\n".join(log))
        raise RuntimeError("nvcc failed; see output above")
    if verbose:
        print("\n".join(log))
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH), *objs,
            "-cudart", "static"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("link failed")
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
