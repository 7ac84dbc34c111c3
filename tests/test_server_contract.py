"""HTTP contract on CPU: the new server answers exactly like the reference's unmodified server did when
tests/golden/make_golden.py drove it (reference_http_contract.json), for BASELINE.json configs[0]
(template example_model predict on one synthetic 224x224 JPEG via POST /predict)."""
import io
import json
import os
import threading
import time

import numpy as np
import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_http_contract.json")))
CASES = {c["case"]: c for c in GOLD["cases"]}


def _jpeg(seed=0, h=224, w=224):
    from PIL import Image
    from oracle import images
    buf = io.BytesIO()
    Image.fromarray(images.synthetic_image(seed, h, w)).save(buf, format="JPEG", quality=90)
    return buf.getvalue()


@pytest.fixture()
def client(monkeypatch):
    monkeypatch.setenv("MODEL", "template")
    monkeypatch.setenv("REGISTER", "0")
    from fastapi.testclient import TestClient
    from mlmicroservicetemplate_b200.src.server import main, dependency
    from mlmicroservicetemplate_b200.src.model import model as plugin
    import concurrent.futures as cf
    plugin.shutdown()
    dependency.pool = cf.ThreadPoolExecutor(4)
    main.dep.pool = dependency.pool
    main._init_submitted = False
    dependency.model_settings.ready_to_predict = False
    with TestClient(main.app, raise_server_exceptions=False) as c:
        yield c, main, dependency, plugin


def _wait_ready(dep, t=10):
    t0 = time.time()
    while not dep.model_settings.ready_to_predict and time.time() - t0 < t:
        time.sleep(0.01)
    assert dep.model_settings.ready_to_predict


def _same(resp, gold):
    assert resp.status_code == gold["status"], (resp.status_code, gold)
    if isinstance(gold["body"], (dict, list)):
        assert resp.json() == gold["body"]


def test_matrix_matches_the_reference_server(client):
    c, main, dep, plugin = client
    img = _jpeg()
    _wait_ready(dep)
    _same(c.get("/"), CASES["GET / (any time)"])
    dep.model_settings.ready_to_predict = False
    _same(c.get("/status"), CASES["GET /status before ready"])
    _same(c.post("/predict", files={"image_file": ("a.jpg", img, "image/jpeg")}), CASES["POST /predict before ready"])
    dep.model_settings.ready_to_predict = True
    _same(c.get("/status"), CASES["GET /status ready"])
    _same(c.post("/predict", files={"image_file": ("a.jpg", img, "image/jpeg")}), CASES["POST /predict ready"])
    r = c.post("/predict", files={"filename": ("a.jpg", img, "image/jpeg")})
    assert r.status_code == CASES["POST /predict wrong field"]["status"] == 422
    r = c.post("/predict", files={"image_file": ("a.jpg", b"not an image", "image/jpeg")})
    assert r.status_code == CASES["POST /predict bad bytes"]["status"] == 500


def test_post_status_is_additive(client):
    c, main, dep, plugin = client
    assert CASES["POST /status (reference has no such route)"]["status"] == 405
    _wait_ready(dep)
    r = c.post("/status")
    assert r.status_code == 200 and r.json() == CASES["GET /status ready"]["body"]
    dep.model_settings.ready_to_predict = False
    r = c.post("/status")
    assert r.status_code == 503 and r.json() == CASES["GET /status before ready"]["body"]


def test_cors_headers_as_in_the_reference(client):
    c, *_ = client
    r = c.get("/", headers={"Origin": "http://localhost:3000"})
    assert r.headers.get("access-control-allow-origin") == "http://localhost:3000"
    assert r.headers.get("access-control-allow-credentials") == "true"
    r = c.get("/", headers={"Origin": "http://evil.example"})
    assert "access-control-allow-origin" not in r.headers


def test_predict_route_batches_concurrent_clients(client):
    """/predict with a classifier plugged in: requests are decoded off-loop and coalesced by the engine.
    The GPU worker is replaced by a recording stand-in; everything above it is the product code."""
    c, main, dep, plugin = client
    from mlmicroservicetemplate_b200.runtime import Engine
    from concurrent.futures import ThreadPoolExecutor

    class Worker:
        def __init__(self):
            self.batches = []
        def run_batch(self, images):
            self.batches.append(len(images))
            time.sleep(0.01)
            n = len(images)
            # "class" = image height - 224, probability from the mean pixel: lets the test tie results to inputs
            idx = np.array([[im.shape[0] - 224, 999] for im in images], np.int32)
            probs = np.array([[float(im.mean()) / 255.0, 0.0] for im in images], np.float32)
            return probs, idx

    _wait_ready(dep)
    w = Worker()
    plugin._state.update(engine=Engine([w], max_batch=8, max_wait_us=30000), classes=plugin.class_names(1000),
                         arch="resnet18", decode_pool=ThreadPoolExecutor(4))
    try:
        out = [None] * 24
        def post(i):
            out[i] = c.post("/predict", files={"image_file": (f"{i}.jpg", _jpeg(i, 224 + i, 240), "image/jpeg")})
        ts = [threading.Thread(target=post, args=(i,)) for i in range(24)]
        [t.start() for t in ts]; [t.join() for t in ts]
        for i, r in enumerate(out):
            assert r.status_code == 200
            body = r.json()
            # the hand-assembled body (class list rendered once per model) is byte-for-byte what the framework's own
            # JSON response of the same dict would have been
            from starlette.responses import JSONResponse
            assert r.content == JSONResponse(content=body).body
            assert r.headers["content-type"] == "application/json"
            assert body["status"] == "success" and len(body["result"]["classes"]) == 1000
            assert f"class_{i:04d}" in body["result"]["result"], "each client must get ITS image's result"
        assert sum(w.batches) == 24 and max(w.batches) > 1
    finally:
        plugin.shutdown()


def test_blocking_predict_keeps_the_reference_signature(client):
    import inspect
    from types import SimpleNamespace
    c, main, dep, plugin = client
    _wait_ready(dep)
    sig = inspect.signature(plugin.predict)
    assert list(sig.parameters) == ["image_file"] and inspect.signature(plugin.init).parameters == {}
    out = plugin.predict(SimpleNamespace(file=io.BytesIO(_jpeg())))
    assert out == CASES["POST /predict ready"]["body"]["result"]


def _mp(parts, boundary="XbOuNdArY"):
    """parts: [(header lines, payload bytes)] -> (body, content type)"""
    out = b""
    for headers, payload in parts:
        out += b"--" + boundary.encode() + b"\r\n" + "\r\n".join(headers).encode() + b"\r\n\r\n" + payload + b"\r\n"
    return out + b"--" + boundary.encode() + b"--\r\n", f"multipart/form-data; boundary={boundary}"


def test_single_image_part_recognises_only_the_plain_upload():
    from mlmicroservicetemplate_b200.src.server.main import single_image_part
    data = bytes(range(256)) * 40 + b"\r\n--not-the-boundary\r\n"
    cd = 'Content-Disposition: form-data; name="image_file"; filename="a.jpg"'
    body, ct = _mp([([cd, "Content-Type: image/jpeg"], data)])
    assert single_image_part(body, ct) == (data, "a.jpg", "image/jpeg")
    assert single_image_part(body, ct.replace("boundary=", 'boundary="') + '"') == (data, "a.jpg", "image/jpeg")
    assert single_image_part(body[:-2], ct)[0] == data                       # no CRLF after the closing delimiter
    body2, _ = _mp([([cd], data)])
    assert single_image_part(body2, ct) == (data, "a.jpg", "application/octet-stream")
    # everything else goes to the framework's parser
    for parts in ([(['Content-Disposition: form-data; name="other"; filename="a.jpg"'], data)],
                  [(['Content-Disposition: form-data; name="image_file"'], data)],                    # a field, not a file
                  [([cd], data), (['Content-Disposition: form-data; name="x"'], b"1")],               # two parts
                  [(["Content-Disposition: form-data; name=\"image_file\"; filename*=UTF-8''a.jpg"], data)],
                  [(['Content-Disposition: form-data; name=image_file; filename="a.jpg"'], data)],     # unquoted
                  [([cd, "X-Extra: 1"], data)]):
        b, c = _mp(parts)
        assert single_image_part(b, c) is None, parts[0][0]
    assert single_image_part(body, "application/json") is None
    assert single_image_part(body, "multipart/form-data") is None
    assert single_image_part(body[:-8], ct) is None                           # closing delimiter cut off
    assert single_image_part(b"junk" + body, ct) is None


def test_fast_path_answers_exactly_like_the_router(client, monkeypatch):
    """The same uploads through FastPredict and, with the fast path bypassed, through the framework's parser: status,
    body bytes and headers are identical; requests the fast path does not recognise reach the router untouched."""
    c, main, dep, plugin = client
    _wait_ready(dep)
    taken = []
    real = main.single_image_part
    monkeypatch.setattr(main, "single_image_part", lambda b, t: (taken.append(1), real(b, t))[1])
    jpeg = _jpeg(3)

    def drop(h):
        return {k: v for k, v in h.items() if k.lower() not in ("date",)}
    fast = c.post("/predict", files={"image_file": ("a.jpg", jpeg, "image/jpeg")}, headers={"Origin": "http://localhost:3000"})
    assert fast.status_code == 200 and taken
    monkeypatch.setattr(main, "single_image_part", lambda b, t: None)
    slow = c.post("/predict", files={"image_file": ("a.jpg", jpeg, "image/jpeg")}, headers={"Origin": "http://localhost:3000"})
    assert (fast.status_code, fast.content, drop(fast.headers)) == (slow.status_code, slow.content, drop(slow.headers))
    monkeypatch.setattr(main, "single_image_part", real)
    # not recognised -> the router's own answers
    r = c.post("/predict", files={"wrong_name": ("a.jpg", jpeg, "image/jpeg")})
    assert r.status_code == 422 and r.json()["detail"][0]["loc"] == ["body", "image_file"]
    r = c.post("/predict", data={"image_file": "not a file"})
    assert r.status_code == 422
    r = c.post("/predict", files={"image_file": ("a.jpg", jpeg, "image/jpeg")}, data={"extra": "1"})
    assert r.status_code == 200
    bad = c.post("/predict", files={"image_file": ("a.jpg", b"these are not image bytes", "image/jpeg")})
    assert bad.status_code == 500                                              # undecodable upload, as in the reference
    dep.model_settings.ready_to_predict = False
    r = c.post("/predict", files={"image_file": ("a.jpg", jpeg, "image/jpeg")})
    assert r.status_code == 503 and r.json() == main.NOT_READY
