"""numpy restatement of stage P (antialiased bilinear resize -> crop -> /255 -> (x-mean)/std) —
oracle / test infrastructure, and the written spec the CUDA kernel (csrc/preprocess.cu) follows.

The reference has no such code (src/model/model.py:23 stops at Image.open); the algorithm restated
here is the one torch 2.11.0's CPU kernel applies for
    F.interpolate(x.float(), size, mode="bilinear", align_corners=False, antialias=True)
reached from SP/torchvision/transforms/_functional_tensor.py:462-467, followed by
_functional_tensor.py:98-99 (/255) and :928 (sub(mean).div(std)), SP/torchvision/transforms/_presets.py:45-46
(mean/std).  The published algorithm is ATen's separable antialias upsampling
(aten/src/ATen/native/cpu/UpSampleKernel.cpp, `_compute_indices_min_size_weights_aa` +
`basic_loop_aa_horizontal/vertical`); the exact operation order below — including the
"groups of four taps un-fused, trailing taps fused" accumulation the compiled loop exhibits — was
established by probing the installed build and is checked bit-for-bit against it in
tests/test_oracle.py and by tests/golden/aa_resize_*.npz.
"""
from __future__ import annotations

import math

import numpy as np

f32 = np.float32
f64 = np.float64

MEAN = np.array([0.485, 0.456, 0.406], dtype=f32)
STD = np.array([0.229, 0.224, 0.225], dtype=f32)


def aa_weights(in_size: int, out_size: int):
    """Per output index: (xmin, xsize, fp32 weights).  C++ promotion rules reproduced explicitly."""
    scale = f32(in_size) / f32(out_size)
    support = scale if scale >= 1 else f32(1)
    invscale = f32(f64(1.0) / f64(scale)) if scale >= 1 else f32(1)
    max_taps = int(math.ceil(float(support))) * 2 + 1
    out = []
    for i in range(out_size):
        center = f32(f64(scale) * (i + 0.5))
        xmin = max(int(f64(f32(center - support)) + 0.5), 0)
        xsize = min(int(f64(f32(center + support)) + 0.5), in_size) - xmin
        xsize = min(max(xsize, 0), max_taps)
        w = np.zeros(xsize, f32)
        tot = f32(0)
        for j in range(xsize):
            a = f32(f32(j + xmin) - center)
            t = abs(f32((f64(a) + 0.5) * f64(invscale)))
            w[j] = f32(1.0 - f64(t)) if t < 1.0 else f32(0)
            tot = f32(tot + w[j])
        if tot != 0:
            w = (w / tot).astype(f32)
        out.append((xmin, xsize, w))
    return out


def _pass_1d(x: np.ndarray, wl, axis: int) -> np.ndarray:
    """t = x0*w0; taps 1..4*floor((n-1)/4): t = t + (x*w) with both roundings; remaining taps fused."""
    x = np.moveaxis(x, axis, 0)
    out = np.zeros((len(wl),) + x.shape[1:], f32)
    for o, (xmin, xs, w) in enumerate(wl):
        acc = (x[xmin] * w[0]).astype(f32)
        n = xs - 1
        nmain = (n // 4) * 4
        for jj in range(n):
            j = jj + 1
            if jj < nmain:
                acc = (acc + (x[xmin + j] * w[j]).astype(f32)).astype(f32)
            else:   # fused multiply-add: exact product, one rounding
                acc = (acc.astype(f64) + x[xmin + j].astype(f64) * f64(w[j])).astype(f32)
        out[o] = acc
    return np.moveaxis(out, 0, axis)


def resize_aa(img_hwc_u8: np.ndarray, res_h: int, res_w: int) -> np.ndarray:
    """uint8/float HWC -> fp32 HWC resized (horizontal pass first, then vertical)."""
    h, w, _ = img_hwc_u8.shape
    x = img_hwc_u8.astype(f32)
    if w != res_w:
        x = _pass_1d(x, aa_weights(w, res_w), 1)
    if h != res_h:
        x = _pass_1d(x, aa_weights(h, res_h), 0)
    return x


def preprocess(img_hwc_u8: np.ndarray, res_h: int, res_w: int, top: int, left: int, out: int = 224) -> np.ndarray:
    """-> fp32 CHW [3, out, out], normalised."""
    x = resize_aa(img_hwc_u8, res_h, res_w)[top:top + out, left:left + out, :]
    x = (x / f32(255.0)).astype(f32)
    x = ((x - MEAN).astype(f32) / STD).astype(f32)
    return np.ascontiguousarray(x.transpose(2, 0, 1))


def preset_geometry(h: int, w: int, resize: int = 256, crop: int = 224):
    """ImageClassification preset: short side -> `resize` (SP/torchvision/transforms/functional.py:368-370,
    long side int(resize*long/short)), centre crop (functional.py:592-593)."""
    if w <= h:
        res_w, res_h = resize, int(resize * h / w)
    else:
        res_h, res_w = resize, int(resize * w / h)
    top = int(round((res_h - crop) / 2.0))
    left = int(round((res_w - crop) / 2.0))
    return res_h, res_w, top, left
