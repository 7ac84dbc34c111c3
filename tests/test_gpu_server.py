"""The whole stack on a GPU: HTTP POST /predict (multipart JPEG/PNG) -> decode thread -> micro-batch
queue -> b2p_predict_host on the replica -> result dict, against the oracle's predict() on the very
same encoded bytes (reference contract src/model/model.py:16-31 + src/server/main.py:119-140)."""
import io
import os
import threading
import time
from types import SimpleNamespace

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _encoded(seed, h, w, fmt):
    from PIL import Image
    from oracle import images
    buf = io.BytesIO()
    Image.fromarray(images.synthetic_image(seed, h, w)).save(buf, format=fmt, **({"quality": 90} if fmt == "JPEG" else {}))
    return buf.getvalue()


@pytest.fixture()
def live_server(monkeypatch, cuda_device):
    for k, v in dict(MODEL="resnet18", REGISTER="0", MAX_BATCH="16", MAX_WAIT_US="20000", TOPK="5",
                     PREPROCESS="preset", GPUS="0", SEED="0").items():
        monkeypatch.setenv(k, v)
    import concurrent.futures as cf
    from fastapi.testclient import TestClient
    from mlmicroservicetemplate_b200.src.model import model as plugin
    from mlmicroservicetemplate_b200.src.server import dependency, main
    plugin.shutdown()
    dependency.pool = cf.ThreadPoolExecutor(4)
    main.dep.pool = dependency.pool
    main._init_submitted = False
    dependency.model_settings.ready_to_predict = False
    with TestClient(main.app, raise_server_exceptions=False) as c:
        t0 = time.time()
        while not dependency.model_settings.ready_to_predict and time.time() - t0 < 120:
            time.sleep(0.05)
        assert dependency.model_settings.ready_to_predict, "init() did not finish"
        yield c, plugin
    plugin.shutdown()


def test_concurrent_predict_matches_the_oracle(live_server):
    client, plugin = live_server
    from oracle import model as om
    om.init(arch="resnet18", topk=5, mode="preset")
    sizes = [(224, 224), (256, 256), (300, 400), (480, 640), (333, 517), (720, 1280)]
    payloads = [_encoded(500 + i, *sizes[i % len(sizes)], "JPEG" if i % 3 else "PNG") for i in range(48)]
    want = [om.predict(SimpleNamespace(file=io.BytesIO(b))) for b in payloads]

    got, lat = [None] * len(payloads), [0.0] * len(payloads)
    def post(i):
        t0 = time.perf_counter()
        got[i] = client.post("/predict", files={"image_file": (f"{i}.img", payloads[i], "application/octet-stream")})
        lat[i] = time.perf_counter() - t0
    threads = [threading.Thread(target=post, args=(i,)) for i in range(len(payloads))]
    [t.start() for t in threads]
    [t.join() for t in threads]

    for i, (r, w) in enumerate(zip(got, want)):
        assert r.status_code == 200, r.text
        body = r.json()
        assert body["status"] == "success" and body["result"]["classes"] == w["classes"]
        res = body["result"]["result"]
        top_got = max(res, key=res.get)
        top_want = max(w["result"], key=w["result"].get)
        assert top_got == top_want, f"request {i}: top-1 {top_got} != oracle {top_want}"
        assert abs(res[top_got] - w["result"][top_want]) < 5e-3
    sizes_seen = plugin._state["engine"].batch_sizes
    assert sum(sizes_seen) == len(payloads) and max(sizes_seen) > 1, "concurrent requests must share GPU batches"
    lat.sort()
    print(f"/predict x{len(payloads)} concurrent: batches {sizes_seen}, p50 {1e3 * lat[len(lat) // 2]:.1f} ms, "
          f"p99 {1e3 * lat[-1]:.1f} ms")


def test_blocking_predict_and_error_paths(live_server):
    client, plugin = live_server
    out = plugin.predict(SimpleNamespace(file=io.BytesIO(_encoded(1, 300, 300, "PNG"))))
    assert len(out["classes"]) == 1000 and len(out["result"]) == 5 and abs(sum(out["result"].values()) - 1.0) < 1.0
    assert client.post("/predict", files={"image_file": ("x", b"not an image", "image/jpeg")}).status_code == 500
    assert client.post("/predict", files={"other": ("x", b"1", "image/jpeg")}).status_code == 422
    assert client.get("/status").json()["status"] == "success"
    # greyscale and RGBA uploads are converted to RGB by the decode stage
    from PIL import Image
    for mode in ("L", "RGBA"):
        buf = io.BytesIO()
        Image.new(mode, (260, 300), 128).save(buf, format="PNG")
        assert client.post("/predict", files={"image_file": ("x.png", buf.getvalue(), "image/png")}).status_code == 200


def test_gpu_decoded_and_host_decoded_uploads_give_identical_answers(cuda_device):
    """DECODE=gpu must not change a single bit of the answer: the same JPEG files go through the engine once as
    bytes (decoded by b2p_jpeg_decode into the device slab) and once as Pillow-decoded pixels (host slab + H2D)."""
    from PIL import Image
    from mlmicroservicetemplate_b200 import backend as B, runtime as R
    from oracle import images
    engine, _ = R.build_gpu_engine("resnet18", [0], max_batch=16, max_wait_us=50000, topk=5, mode="preset", seed=0)
    try:
        files = []
        for i, (h, w, sub) in enumerate([(224, 224, 2), (300, 400, 2), (256, 256, 0), (480, 640, 1), (333, 517, 2),
                                          (231, 229, 2), (720, 1280, 2), (224, 224, 0)]):
            buf = io.BytesIO()
            Image.fromarray(images.synthetic_image(700 + i, h, w)).save(buf, "JPEG", quality=88, subsampling=sub)
            files.append(buf.getvalue())
        infos = [B.jpeg_inspect(f) for f in files]
        assert all(i is not None for i in infos)
        as_bytes = engine.predict_many([R.JpegFile(f, i.height, i.width) for f, i in zip(files, infos)])
        as_pixels = engine.predict_many([np.asarray(Image.open(io.BytesIO(f)).convert("RGB")) for f in files])
        for (pa, ia), (pb, ib) in zip(as_bytes, as_pixels):
            assert np.array_equal(ia, ib) and np.array_equal(pa, pb)
        # a mixed batch: GPU-decoded files next to host-decoded pixels, plus one damaged file that must fail alone
        hd_cut = files[1][:len(files[1]) // 2] + b"\xff\xd9"
        info_cut = B.jpeg_inspect(hd_cut)
        assert info_cut is not None
        mixed = [R.JpegFile(files[0], infos[0].height, infos[0].width),
                 np.asarray(Image.open(io.BytesIO(files[2])).convert("RGB")),
                 R.JpegFile(hd_cut, info_cut.height, info_cut.width),
                 R.JpegFile(files[4], infos[4].height, infos[4].width)]
        futs = [engine.submit(m) for m in mixed]
        for j, want in ((0, 0), (1, 2), (3, 4)):
            p, i = futs[j].result(timeout=60)
            assert np.array_equal(i, as_pixels[want][1]) and np.array_equal(p, as_pixels[want][0])
        with pytest.raises(R.CorruptImage):
            futs[2].result(timeout=60)
    finally:
        engine.close()


def test_plugin_init_from_packed_weights_and_labels(tmp_path, monkeypatch, cuda_device):
    """WEIGHTS=<file.b2pw>: init() reads the packed program and its class names; predict() answers with them."""
    from mlmicroservicetemplate_b200 import packed, program as P
    from mlmicroservicetemplate_b200.runtime import seeded_model
    from mlmicroservicetemplate_b200.src.model import model as plugin
    names = [f"species-{i}" for i in range(1000)]
    path = str(tmp_path / "r18.b2pw")
    packed.save_program(P.build_program(seeded_model("resnet18", 0), "resnet18"), path, names)
    for k, v in dict(MODEL="ignored-when-WEIGHTS-is-set", WEIGHTS=path, GPUS="0", MAX_BATCH="4", MAX_WAIT_US="1000",
                     TOPK="3", PREPROCESS="preset").items():
        monkeypatch.setenv(k, v)
    plugin.shutdown()
    plugin.init()
    try:
        out = plugin.predict(SimpleNamespace(file=io.BytesIO(_encoded(9, 256, 320, "JPEG"))))
        assert out["classes"] == names and len(out["result"]) == 3
        assert all(k.startswith("species-") for k in out["result"])
        # same upload through a module-built engine: identical numbers
        plugin.shutdown()
        monkeypatch.delenv("WEIGHTS")
        monkeypatch.setenv("MODEL", "resnet18")
        plugin.init()
        ref = plugin.predict(SimpleNamespace(file=io.BytesIO(_encoded(9, 256, 320, "JPEG"))))
        assert sorted(out["result"].values()) == sorted(ref["result"].values())
    finally:
        plugin.shutdown()
