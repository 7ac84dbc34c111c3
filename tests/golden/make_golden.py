"""Generates the committed fixtures in tests/golden/.  Run in the BUILD container only (it imports the
reference from /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

1. reference_http_contract.json — the reference's own src/server/main.py, UNMODIFIED, driven through
   fastapi's TestClient under the two import shims of SURVEY §8(c) (pydantic.BaseSettings -> pydantic.v1,
   a stand-in `secrets.API_KEY`); every (request -> status, body) pair of the contract matrix.
2. aa_resize_vectors.npz — torch's F.interpolate(bilinear, antialias=True) + normalise (the oracle's P0)
   on seeded synthetic images, sub-sampled, plus float64 checksums of the full tensors.
3. resnet18_logits.npz — oracle fp32 CPU logits / top-5 of ResNet-18 (seed 0) on seeded images.
4. jpeg_vectors.npz — small JPEG files written by Pillow (4:2:0 / 4:2:2 / 4:4:4 / greyscale, odd sizes, restart
   intervals, optimised tables) and the pixels Pillow itself decodes from them (`Image.open(f).convert('RGB')`,
   the reference's decoder, model.py:23), Pillow + libjpeg-turbo versions recorded.
"""
import io
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def jpeg_bytes(seed=0, h=224, w=224, quality=90):
    from PIL import Image
    from oracle import images
    buf = io.BytesIO()
    Image.fromarray(images.synthetic_image(seed, h, w)).save(buf, format="JPEG", quality=quality)
    return buf.getvalue()


def reference_contract():
    import secrets as _secrets
    import pydantic
    import pydantic.v1
    _secrets.API_KEY = "test"                       # satisfies `from secrets import API_KEY` (ref server_connection.py:6)
    pydantic.BaseSettings = pydantic.v1.BaseSettings   # satisfies ref dependency.py:3,6-7 on pydantic 2
    sys.path.append(REF)                               # appended: ref secrets.py must not shadow the stdlib
    from fastapi.testclient import TestClient
    from src.server import main as ref_main
    from src.server import dependency as ref_dep

    img = jpeg_bytes()
    rows = []

    def rec(name, resp):
        try:
            body = resp.json()
        except Exception:
            body = resp.text
        rows.append({"case": name, "status": resp.status_code, "body": body,
                     "content_type": resp.headers.get("content-type", "")})

    client = TestClient(ref_main.app, raise_server_exceptions=False)
    with client:
        # not ready yet: init() sleeps 1 s on a pool thread (ref model.py:13)
        ref_dep.model_settings.ready_to_predict = False
        rec("GET / (any time)", client.get("/"))
        rec("GET /status before ready", client.get("/status"))
        rec("POST /predict before ready", client.post("/predict", files={"image_file": ("a.jpg", img, "image/jpeg")}))
        t0 = time.time()
        while not ref_dep.model_settings.ready_to_predict and time.time() - t0 < 10:
            time.sleep(0.05)
        ref_dep.model_settings.ready_to_predict = True
        rec("GET /status ready", client.get("/status"))
        rec("POST /predict ready", client.post("/predict", files={"image_file": ("a.jpg", img, "image/jpeg")}))
        rec("POST /predict wrong field", client.post("/predict", files={"filename": ("a.jpg", img, "image/jpeg")}))
        rec("POST /predict bad bytes", client.post("/predict", files={"image_file": ("a.jpg", b"not an image", "image/jpeg")}))
        rec("POST /status (reference has no such route)", client.post("/status"))
        ref_dep.shutdown = True
    json.dump({"generated_from": "CodyRichter/MLMicroserviceTemplate src/server/main.py (unmodified, shims per SURVEY 8c)",
               "cases": rows}, open(os.path.join(HERE, "reference_http_contract.json"), "w"), indent=1)
    print("contract cases:", [(r["case"], r["status"]) for r in rows])


def aa_vectors():
    from oracle import images, model as om
    from oracle.preprocess_np import preset_geometry
    cases = [(11, 320, 320, "direct"), (12, 333, 517, "direct"), (13, 480, 640, "preset"),
             (14, 1080, 1920, "direct"), (15, 301, 227, "preset"), (16, 224, 224, "direct")]
    out = {}
    meta = []
    for seed, h, w, mode in cases:
        im = images.synthetic_image(seed, h, w)
        geom = preset_geometry(h, w) if mode == "preset" else (224, 224, 0, 0)
        x = om.preprocess(im, *geom)[0].numpy()
        key = f"s{seed}_{h}x{w}_{mode}"
        out[key] = x[:, ::7, ::5].copy()
        meta.append({"key": key, "seed": seed, "h": h, "w": w, "mode": mode, "geom": list(geom),
                     "sum": float(x.astype(np.float64).sum()), "abs_sum": float(np.abs(x.astype(np.float64)).sum())})
    np.savez_compressed(os.path.join(HERE, "aa_resize_vectors.npz"), meta=json.dumps(meta), **out)
    print("aa vectors:", [m["key"] for m in meta])


def resnet18_logits():
    import torch
    from oracle import images, model as om
    ims = images.synthetic_batch(1, 4, 224, 224) + [images.synthetic_image(3, 480, 640)]
    model = om.build_model("resnet18")
    logits, probs, idx = om.predict_arrays(ims, mode="direct", k=5, model=model)
    np.savez_compressed(os.path.join(HERE, "resnet18_logits.npz"), logits=logits.numpy(), probs=probs.numpy(),
                        idx=idx.numpy(), note="images: synthetic_batch(1,4,224,224) + synthetic_image(3,480,640); "
                        "weights: torch.manual_seed(0) resnet18(weights=None); torch " + torch.__version__)
    print("resnet18 logits", logits.shape, float(logits.abs().max()))


def jpeg_vectors():
    import PIL
    from PIL import Image, ImageFile, features
    from oracle import images
    ImageFile.MAXBLOCK = 1 << 24
    cases = [(21, 16, 16, dict(quality=90)), (22, 33, 47, dict(quality=75)),
             (23, 33, 47, dict(quality=75, subsampling=0)), (24, 41, 30, dict(quality=60, subsampling=1)),
             (25, 48, 64, dict(quality=85, restart_marker_blocks=3)), (26, 24, 100, dict(quality=30, optimize=True)),
             (27, 37, 53, dict(quality=95, grey=True)), (28, 64, 64, dict(quality=100, subsampling=2))]
    out = {}
    meta = []
    for seed, h, w, kw in cases:
        kw = dict(kw)
        grey = kw.pop("grey", False)
        arr = images.synthetic_image(seed, h, w)
        buf = io.BytesIO()
        Image.fromarray(arr[..., 0] if grey else arr).save(buf, "JPEG", **kw)
        data = buf.getvalue()
        key = f"s{seed}_{h}x{w}"
        out[key + "_file"] = np.frombuffer(data, np.uint8)
        out[key + "_rgb"] = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
        meta.append({"key": key, "h": h, "w": w, "grey": grey, "save_args": kw, "bytes": len(data)})
    note = f"Pillow {PIL.__version__}, libjpeg-turbo {features.version('libjpeg_turbo')}"
    np.savez_compressed(os.path.join(HERE, "jpeg_vectors.npz"), meta=json.dumps(meta), note=note, **out)
    print("jpeg vectors:", note, [m["key"] for m in meta])


if __name__ == "__main__":
    jpeg_vectors()
    aa_vectors()
    resnet18_logits()
    reference_contract()
