"""The C-ABI library: builds for sm_100a, loads without a GPU, exports exactly what include/b200predict.h
declares, has matching struct layouts, and fails loudly (no fallback) when there is no sm_100 device."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200predict.h")


@pytest.fixture(scope="module")
def lib():
    from mlmicroservicetemplate_b200 import build, backend
    build.build()
    return backend.load_library()


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2p_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from mlmicroservicetemplate_b200 import backend
    declared = _declared_functions()
    assert len(declared) >= 12
    bound = sorted(n for n, _, _ in backend.SYMBOLS)
    assert declared == bound, "backend.SYMBOLS must list exactly the header's functions"
    out = subprocess.run(["nm", "-D", "--defined-only", str(backend.LIB_PATH)], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (b2p_[a-z_0-9]+)", out))
    assert set(declared) <= exported


def test_struct_layouts_match_the_header(tmp_path, lib):
    from mlmicroservicetemplate_b200 import backend as B
    c = tmp_path / "sz.c"
    c.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "b200predict.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                 'sizeof(b2p_op),sizeof(b2p_model_desc),sizeof(b2p_image),sizeof(b2p_stats),offsetof(b2p_model_desc,blob_bytes),'
                 'offsetof(b2p_op,weight),offsetof(b2p_image,res_h));return 0;}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    want = [ctypes.sizeof(B.Op), ctypes.sizeof(B.ModelDesc), ctypes.sizeof(B.Image), ctypes.sizeof(B.Stats),
            B.ModelDesc.blob_bytes.offset, B.Op.weight.offset, B.Image.res_h.offset]
    assert got == want


def test_build_targets_sm100a_with_tcgen05_and_tma(lib):
    from mlmicroservicetemplate_b200 import backend
    sass = subprocess.run(["cuobjdump", "-sass", str(backend.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):     # tcgen05.mma, TMA loads (incl. IM2COL), tcgen05.ld
        assert mnemonic in sass, mnemonic
    assert "UTMALDG.4D.IM2COL" in sass
    assert "HMMA.16" not in sass, "no legacy mma.sync tensor path"


def test_no_gpu_means_loud_failure_not_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from mlmicroservicetemplate_b200 import backend as B, program as P
    prog = P.single_linear_program(torch.zeros(8, 64), torch.zeros(8))
    with pytest.raises(B.B2PError) as ei:
        B.Replica(0, prog, max_batch=4)
    assert ei.value.code in (-2, -3)
    from mlmicroservicetemplate_b200 import runtime
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        runtime.build_gpu_engine("resnet18", None, 8, 1000, 5, "direct", 0)


def test_create_rejects_malformed_programs(lib):
    from mlmicroservicetemplate_b200 import backend as B
    h = ctypes.c_void_p()
    assert lib.b2p_create(0, None, ctypes.byref(h)) == -1
    desc = B.ModelDesc(abi_version=999)
    assert lib.b2p_create(0, ctypes.byref(desc), ctypes.byref(h)) == -1
    assert b"invalid" in lib.b2p_last_error(None)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import it."""
    pkg = os.path.join(ROOT, "mlmicroservicetemplate_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), os.path.join(dp, f)
    # the benchmark scripts measure the product: they may not lean on the oracle either; bench.py may use it in the
    # cpu_baseline / reference-arm leg only (one import, inside cpu_reference_run)
    for f in os.listdir(os.path.join(ROOT, "benchmarks")):
        if f.endswith(".py"):
            txt = open(os.path.join(ROOT, "benchmarks", f)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
    bench = open(os.path.join(ROOT, "bench.py")).read()
    imports = re.findall(r"^\s*(?:from|import)\s+oracle\b.*$", bench, flags=re.M)
    assert len(imports) == 1, imports
    body = bench[bench.index("def cpu_reference_run"):bench.index("def run_reference_arm")]
    assert imports[0].strip() in body, "bench.py may import the oracle only inside cpu_reference_run"
