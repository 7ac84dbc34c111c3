"""The reference's model plugin contract filled in on the CPU — oracle / test infrastructure.

Shape follows CodyRichter/MLMicroserviceTemplate src/model/model.py: module-level `init()` (:6-13)
and `predict(image_file)` (:16-31, reads only `image_file.file`, returns
{'classes': [...], 'result': {class: value}}).  Arithmetic: Pillow decode -> fp32 preprocess
(SP/torchvision/transforms/_functional_tensor.py:462-467, :98-99, :928) -> torchvision classifier in
fp32 eval mode (SP/torchvision/models/resnet.py:266-282) -> softmax -> top-k.

Weights: torch.manual_seed(SEED); torchvision.models.<arch>(weights=None).eval()  (random init: no
network, no checkpoints).  The product is handed this same state_dict, so parity never depends on
re-deriving an RNG stream on the GPU.
"""
from __future__ import annotations

import os
from typing import List, Optional

import numpy as np
import torch
import torch.nn.functional as F

SEED = 0
MEAN = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
STD = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)

_state = {"model": None, "arch": None, "classes": None, "topk": 5, "mode": "direct"}


def class_names(n: int) -> List[str]:
    return [f"class_{i:04d}" for i in range(n)]


def build_model(arch: str = "resnet18", seed: int = SEED) -> torch.nn.Module:
    import torchvision
    torch.manual_seed(seed)
    m = getattr(torchvision.models, arch)(weights=None)
    if arch.startswith("vit"):
        # the stock ViT head is all zeros (SP/torchvision/models/vision_transformer.py:264-266): every logit
        # would be exactly 0.  Re-draw it N(0, 0.02) from a fixed generator (bias stays 0) — SURVEY §8c c5.
        g = torch.Generator().manual_seed(123)
        with torch.no_grad():
            m.heads.head.weight.normal_(0.0, 0.02, generator=g)
    return m.eval()


def init(arch: Optional[str] = None, topk: Optional[int] = None, mode: Optional[str] = None) -> None:
    """reference: init() -> None (src/model/model.py:6-13)."""
    arch = arch or os.getenv("MODEL", "resnet18")
    _state["model"] = build_model(arch)
    _state["arch"] = arch
    head = _state["model"].fc if hasattr(_state["model"], "fc") else _state["model"].heads.head
    _state["classes"] = class_names(head.out_features)
    _state["topk"] = int(topk or os.getenv("TOPK", 5))
    _state["mode"] = mode or os.getenv("PREPROCESS", "direct")


def geometry(h: int, w: int, mode: str):
    from .preprocess_np import preset_geometry
    if mode == "preset":
        return preset_geometry(h, w)
    return 224, 224, 0, 0


def preprocess(img_hwc_u8: np.ndarray, res_h: int, res_w: int, top: int, left: int) -> torch.Tensor:
    """P0 of SURVEY §8c: uint8 HWC -> fp32 [1,3,224,224]; no uint8 re-quantisation after the resize."""
    x = torch.from_numpy(np.ascontiguousarray(img_hwc_u8)).permute(2, 0, 1)[None].float()
    if (x.shape[2], x.shape[3]) != (res_h, res_w):
        x = F.interpolate(x, size=(res_h, res_w), mode="bilinear", align_corners=False, antialias=True)
    x = x[:, :, top:top + 224, left:left + 224]
    return (x / 255.0 - MEAN) / STD


@torch.no_grad()
def forward_logits(x: torch.Tensor, model: Optional[torch.nn.Module] = None) -> torch.Tensor:
    return (model or _state["model"])(x.float())


@torch.no_grad()
def predict_arrays(images: List[np.ndarray], mode: Optional[str] = None, k: Optional[int] = None,
                   model: Optional[torch.nn.Module] = None):
    """Batched form used by tests/bench: decoded uint8 HWC images -> (logits, probs[k], idx[k])."""
    mode = mode or _state["mode"]
    k = k or _state["topk"]
    xs = [preprocess(im, *geometry(im.shape[0], im.shape[1], mode)) for im in images]
    logits = forward_logits(torch.cat(xs, 0), model)
    probs, idx = torch.softmax(logits, dim=1).topk(k, dim=1)
    return logits, probs, idx


def predict(image_file) -> dict:
    """reference: predict(image_file: UploadFile = File(...)) -> dict (src/model/model.py:16-31)."""
    from PIL import Image
    image = Image.open(image_file.file).convert("RGB")
    arr = np.asarray(image)
    _, probs, idx = predict_arrays([arr])
    names = _state["classes"]
    return {
        "classes": list(names),
        "result": {names[int(i)]: float(p) for p, i in zip(probs[0], idx[0])},
    }


def bf16_weight_model(model: torch.nn.Module) -> torch.nn.Module:
    """the "bf16 weights, fp32 math" oracle.  A deep copy of a torchvision ResNet whose BatchNorms are
    folded into the preceding convolutions in fp32 (eval statistics) and whose conv / linear WEIGHTS are then rounded
    to bf16 — the values the GPU path multiplies with — while every activation and accumulation stays fp32.  The
    distance GPU <-> this model isolates the GPU's own rounding (bf16 activations between layers, accumulation
    order) from the unavoidable weight/input quantisation."""
    import copy
    from torch.nn.utils.fusion import fuse_conv_bn_eval
    m = copy.deepcopy(model).eval()

    def fold(parent, conv_name, bn_name):
        conv, bn = getattr(parent, conv_name), getattr(parent, bn_name)
        setattr(parent, conv_name, fuse_conv_bn_eval(conv, bn))
        setattr(parent, bn_name, torch.nn.Identity())

    fold(m, "conv1", "bn1")
    for layer in (m.layer1, m.layer2, m.layer3, m.layer4):
        for blk in layer:
            for i in (1, 2, 3):
                if hasattr(blk, f"conv{i}"):
                    fold(blk, f"conv{i}", f"bn{i}")
            if blk.downsample is not None:
                blk.downsample[0] = fuse_conv_bn_eval(blk.downsample[0], blk.downsample[1])
                blk.downsample[1] = torch.nn.Identity()
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, (torch.nn.Conv2d, torch.nn.Linear)):
                mod.weight.copy_(mod.weight.to(torch.bfloat16).float())
    return m
