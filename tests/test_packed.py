"""Packed program files: what init() loads must be byte-identical to what it would have
built, a real state_dict must land in the same place, and a damaged or foreign file must be refused."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from mlmicroservicetemplate_b200 import backend as B, packed, program as P
from mlmicroservicetemplate_b200.runtime import seeded_model


def _same_program(a, b):
    assert (a.name, a.in_elems, a.out_elems, a.slot_elems) == (b.name, b.in_elems, b.out_elems, b.slot_elems)
    assert len(a.ops) == len(b.ops) and len(a.blobs) == len(b.blobs)
    for x, y in zip(a.ops, b.ops):
        assert bytes(x) == bytes(y)
    for x, y in zip(a.blobs, b.blobs):
        assert x.dtype == y.dtype and x.shape == y.shape
        assert torch.equal(x.view(torch.uint8) if x.numel() else x, y.view(torch.uint8) if y.numel() else y)


@pytest.mark.parametrize("arch", ["resnet18", "vit_b_16"])
def test_round_trip_is_byte_identical(tmp_path, arch):
    prog = P.build_program(seeded_model(arch, 0), arch)
    path = str(tmp_path / f"{arch}.b2pw")
    names = [f"label {i}" for i in range(prog.out_elems)]
    hdr = packed.save_program(prog, path, names)
    back, classes = packed.load_program(path)
    _same_program(prog, back)
    assert classes == names and hdr["n_ops"] == len(prog.ops)
    assert abs(os.path.getsize(path) - sum(b.numel() * b.element_size() for b in prog.blobs)) < 1 << 20


def test_state_dict_checkpoint_packs_like_the_module(tmp_path):
    model = seeded_model("resnet18", 7)
    ck = str(tmp_path / "ck.pt")
    torch.save(model.state_dict(), ck)
    labels = str(tmp_path / "labels.txt")
    with open(labels, "w") as f:
        f.write("\n".join(f"thing-{i}" for i in range(1000)) + "\n")
    out = str(tmp_path / "r18.b2pw")
    assert packed.main(["--arch", "resnet18", "--state-dict", ck, "--labels", labels, "--out", out]) == 0
    back, classes = packed.load_program(out)
    _same_program(P.build_program(model, "resnet18"), back)
    assert classes[3] == "thing-3" and len(classes) == 1000


def test_damaged_and_foreign_files_are_refused(tmp_path):
    prog = P.build_program(seeded_model("resnet18", 0), "resnet18")
    path = str(tmp_path / "m.b2pw")
    packed.save_program(prog, path)
    data = bytearray(open(path, "rb").read())
    bad = str(tmp_path / "bad.b2pw")
    flipped = bytearray(data)
    flipped[-1000] ^= 0x40                      # one bit in a weight
    open(bad, "wb").write(flipped)
    with pytest.raises(ValueError, match="checksum"):
        packed.load_program(bad)
    open(bad, "wb").write(data[:len(data) // 2])
    with pytest.raises(ValueError):
        packed.load_program(bad)
    open(bad, "wb").write(b"PK\x03\x04" + bytes(100))
    with pytest.raises(ValueError, match="magic"):
        packed.load_program(bad)
    other_abi = bytearray(data)
    other_abi[8:12] = (B.ABI_VERSION + 1).to_bytes(4, "little")
    open(bad, "wb").write(other_abi)
    with pytest.raises(ValueError, match="ABI"):
        packed.load_program(bad)
    with pytest.raises(ValueError, match="class names"):
        packed.save_program(prog, bad, ["just one"])
