"""CPU checks for the JPEG-decode row: the numpy restatement against Pillow — the
reference's own decoder (src/model/model.py:23) — live and through the committed golden vectors, and the C-ABI's
host-side header parser against the restatement's.  No GPU needed."""
import io
import json
import os

import numpy as np
import pytest
from PIL import Image, ImageFile

from oracle import images, jpeg_np as J

ImageFile.MAXBLOCK = 1 << 24
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "jpeg_vectors.npz")


def _encode(arr, **kw):
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def _pillow(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))


def test_restatement_reproduces_the_golden_vectors():
    z = np.load(GOLDEN)
    meta = json.loads(str(z["meta"]))
    assert len(meta) >= 8
    for m in meta:
        data = z[m["key"] + "_file"].tobytes()
        got = J.decode(data)
        assert got.dtype == np.uint8 and got.shape == (m["h"], m["w"], 3)
        assert np.array_equal(got, z[m["key"] + "_rgb"]), m


@pytest.mark.parametrize("sub", [0, 1, 2])
@pytest.mark.parametrize("size", [(16, 16), (35, 29), (8, 71)])
def test_restatement_matches_live_pillow(size, sub):
    h, w = size
    for q, extra in ((35, {}), (92, {"optimize": True}), (75, {"restart_marker_blocks": 2})):
        data = _encode(images.synthetic_image(5 + q, h, w), quality=q, subsampling=sub, **extra)
        assert np.array_equal(J.decode(data), _pillow(data)), (size, sub, q, extra)


def test_greyscale_and_extreme_content():
    g = images.synthetic_image(3, 40, 24)[..., 1]
    data = _encode(g, quality=80)
    assert np.array_equal(J.decode(data), _pillow(data))
    # saturated checkerboard: exercises the range limiter and the colour clamp
    yy, xx = np.mgrid[0:32, 0:48]
    chk = (((yy // 3 + xx // 5) % 2) * 255).astype(np.uint8)
    rgb = np.stack([chk, 255 - chk, chk], -1)
    for sub in (0, 2):
        data = _encode(rgb, quality=25, subsampling=sub)
        assert np.array_equal(J.decode(data), _pillow(data))


def test_scope_classification():
    arr = images.synthetic_image(1, 32, 32)
    with pytest.raises(J.Unsupported):
        J.parse(_encode(arr, quality=80, progressive=True))
    b = io.BytesIO()
    Image.fromarray(arr).convert("CMYK").save(b, "JPEG")
    with pytest.raises(J.Unsupported):
        J.parse(b.getvalue())
    with pytest.raises(ValueError):
        J.parse(b"\x89PNG\r\n\x1a\n" + bytes(64))
    with pytest.raises(J.Unsupported):
        J.parse(_encode(arr[:, :4], quality=80, subsampling=2))       # chroma only 2 samples wide


def test_c_abi_header_parser_agrees_with_the_restatement():
    """b2p_jpeg_inspect is host-only code inside libb200predict.so: it must load and answer without a GPU."""
    from mlmicroservicetemplate_b200 import backend as B
    arr = images.synthetic_image(2, 50, 70)
    variants = [dict(quality=90), dict(quality=90, subsampling=0), dict(quality=80, subsampling=1),
                dict(quality=80, restart_marker_blocks=2), dict(quality=80, optimize=True)]
    for kw in variants:
        data = _encode(arr, **kw)
        info = B.jpeg_inspect(data)
        hd = J.parse(data)
        assert info is not None
        assert (info.width, info.height, info.components) == (hd.width, hd.height, len(hd.comps))
        assert (info.h_samp, info.v_samp) == (hd.comps[0].h, hd.comps[0].v)
        assert info.restart_interval == hd.restart_interval
        assert info.segments == len(J.restart_segments(data, hd))
        assert info.scan_bytes == hd.scan_end - hd.scan_start
    grey = B.jpeg_inspect(_encode(arr[..., 0], quality=70))
    assert grey is not None and grey.components == 1 and (grey.h_samp, grey.v_samp) == (1, 1)
    assert B.jpeg_inspect(_encode(arr, quality=80, progressive=True)) is None
    assert B.jpeg_inspect(b"not an image") is None
    assert B.jpeg_inspect(_encode(arr, quality=90)[:200]) is None
    z = np.load(GOLDEN)
    for m in json.loads(str(z["meta"])):
        info = B.jpeg_inspect(z[m["key"] + "_file"].tobytes())
        assert info is not None and (info.height, info.width) == (m["h"], m["w"])


def test_header_parser_survives_damaged_files():
    """Uploads are untrusted: whatever bytes arrive, b2p_jpeg_inspect must answer (geometry or 'not for me') and never
    read outside the buffer.  2,400 random mutations of valid files: flipped bytes, cut tails, rewritten segment
    lengths, spliced garbage.  (The buffer handed over is exactly as long as the file, so an over-read would hit the
    allocator's red zone under ASan and, more to the point, the lengths it reports must stay inside the file.)"""
    from mlmicroservicetemplate_b200 import backend as B
    rng = np.random.default_rng(1234)
    arr = images.synthetic_image(4, 40, 56)
    seeds = [_encode(arr, quality=85), _encode(arr, quality=85, subsampling=0, optimize=True),
             _encode(arr, quality=60, restart_marker_blocks=2), _encode(arr[..., 0], quality=70)]
    answered = 0
    for trial in range(2400):
        data = bytearray(seeds[trial % len(seeds)])
        kind = trial % 6
        if kind == 0:                                   # a few random byte flips anywhere
            for _ in range(int(rng.integers(1, 6))):
                data[int(rng.integers(0, len(data)))] = int(rng.integers(0, 256))
        elif kind == 1:                                 # truncation
            data = data[:int(rng.integers(0, len(data)))]
        elif kind == 2:                                 # flips confined to the headers
            for _ in range(int(rng.integers(1, 8))):
                data[int(rng.integers(0, min(len(data), 620)))] = int(rng.integers(0, 256))
        elif kind == 3:                                 # a marker length field rewritten
            pos = data.find(b"\xff", int(rng.integers(2, 300)))
            if 0 <= pos < len(data) - 4:
                data[pos + 2] = int(rng.integers(0, 256))
                data[pos + 3] = int(rng.integers(0, 256))
        elif kind == 4:                                 # garbage spliced in
            at = int(rng.integers(0, len(data)))
            data[at:at] = bytes(rng.integers(0, 256, int(rng.integers(1, 64)), dtype=np.uint8))
        else:                                           # 0xFF bytes sprinkled over the entropy-coded data
            for _ in range(int(rng.integers(1, 10))):
                data[int(rng.integers(len(data) // 2, len(data)))] = 0xFF
        info = B.jpeg_inspect(bytes(data))
        if info is not None:
            answered += 1
            assert 1 <= info.width <= 65535 and 1 <= info.height <= 65535 and info.width * info.height <= 64 << 20
            assert info.components in (1, 3) and info.segments >= 1 and 0 <= info.scan_bytes <= len(data)
    assert answered > 100          # many mutations leave a decodable header: both outcomes are exercised
