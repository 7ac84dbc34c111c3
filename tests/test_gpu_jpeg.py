"""GPU JPEG decode (b2p_jpeg_*) against the reference's own decoder.

The reference decodes with Pillow (src/model/model.py:23); Pillow is present on the GPU box too, so the parity
check is direct: `np.asarray(Image.open(f).convert('RGB'))` versus the bytes the three CUDA kernels write, bit for bit.
"""
import io

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from PIL import Image, ImageFile  # noqa: E402

pytestmark = pytest.mark.gpu
ImageFile.MAXBLOCK = 1 << 24


def _picture(h, w, seed, grey=False):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    a = np.stack([127 + 100 * np.sin(xx / 7.0 + yy / 11.0 + seed), 127 + 90 * np.cos(xx / 5.0 - yy / 3.0),
                  127 + 120 * np.sin(xx * yy / 300.0 + seed)], -1) + rng.normal(0, 12, (h, w, 3))
    a = np.clip(a, 0, 255).astype(np.uint8)
    return a[..., 0] if grey else a


def _encode(arr, **kw):
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def _pillow(data):
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))


def _decode_batch(dec, files):
    from mlmicroservicetemplate_b200 import backend as B
    infos = [B.jpeg_inspect(f) for f in files]
    assert all(i is not None for i in infos)
    offsets, pos = [], 0
    for i in infos:
        offsets.append(pos)
        pos += (i.width * i.height * 3 + 255) // 256 * 256
    slab = torch.full((pos,), 0xAB, dtype=torch.uint8, device="cuda")
    status = torch.full((len(files),), -1, dtype=torch.int32, device="cuda")
    dec.decode(files, slab, pos, offsets, status=status)
    torch.cuda.synchronize()
    host = slab.cpu().numpy()
    outs = [host[o:o + i.width * i.height * 3].reshape(i.height, i.width, 3) for o, i in zip(offsets, infos)]
    return outs, status.cpu().numpy(), host, offsets, infos


@pytest.fixture(scope="module")
def decoder(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    d = B.JpegDecoder(0)
    yield d
    d.close()


CASES = [
    dict(size=(16, 16), kw=dict(quality=90)),
    dict(size=(33, 47), kw=dict(quality=75)),
    dict(size=(33, 47), kw=dict(quality=75, subsampling=0)),
    dict(size=(33, 47), kw=dict(quality=75, subsampling=1)),
    dict(size=(64, 64), kw=dict(quality=100, subsampling=2)),
    dict(size=(17, 100), kw=dict(quality=30, optimize=True)),
    dict(size=(50, 7), kw=dict(quality=85, subsampling=0)),
    dict(size=(224, 224), kw=dict(quality=90)),
    dict(size=(224, 224), kw=dict(quality=90, restart_marker_blocks=5)),
    dict(size=(225, 223), kw=dict(quality=60, subsampling=1, restart_marker_rows=1)),
    dict(size=(480, 640), kw=dict(quality=85)),
    dict(size=(40, 56), kw=dict(quality=85), grey=True),
    dict(size=(37, 53), kw=dict(quality=50), grey=True),
]


def test_mixed_batch_matches_pillow_bit_for_bit(decoder):
    files = [_encode(_picture(*c["size"], seed=i, grey=c.get("grey", False)), **c["kw"]) for i, c in enumerate(CASES)]
    outs, status, host, offsets, infos = _decode_batch(decoder, files)
    assert (status == 0).all(), status
    for i, (f, got) in enumerate(zip(files, outs)):
        ref = _pillow(f)
        assert got.shape == ref.shape
        bad = np.argwhere(got != ref)
        assert bad.size == 0, f"case {i} {CASES[i]}: {len(bad)} differing bytes, first at {bad[:3].tolist()}"
    # nothing outside the images was written
    for o, i, nxt in zip(offsets, infos, offsets[1:] + [len(host)]):
        assert (host[o + i.width * i.height * 3:nxt] == 0xAB).all()


@pytest.mark.parametrize("size,kw", [((16, 16), dict(quality=50)), ((24, 40), dict(quality=95, subsampling=0)),
                                     ((64, 64), dict(quality=90)), ((200, 300), dict(quality=92, subsampling=1))])
def test_single_image_batches(decoder, size, kw):
    """One file per call: the Huffman kernel then runs with the smallest CTAs (32 threads for a tiny file), so the
    table staging and the chunking are exercised at every CTA size, not only the 1024-thread one of a mixed batch."""
    f = _encode(_picture(*size, seed=7), **kw)
    outs, status, *_ = _decode_batch(decoder, [f])
    assert status[0] == 0
    assert np.array_equal(outs[0], _pillow(f))


def test_large_files_spread_over_cta_clusters(decoder):
    """Segments of more than 64 KB no longer fit one CTA's 1024 chunks of 512 bits: the Huffman kernel then runs as a
    cluster of 2, 4 or 8 CTAs that exchange chunk states through distributed shared memory.  Mixed with a tiny file
    (which gets the same cluster size, mostly idle) the pixels must still equal Pillow's."""
    for files in ([_encode(_picture(200, 300, seed=3), quality=93, subsampling=0)],                      # ~80 KB: 2 CTAs
                  [_encode(_picture(300, 400, seed=8), quality=93, subsampling=0)],                      # ~160 KB: 4 CTAs
                  [_encode(_picture(768, 1024, seed=4), quality=95), _encode(_picture(24, 32, seed=5), quality=80)],
                  [_encode(_picture(1080, 1920, seed=6), quality=92), _encode(_picture(1080, 1920, seed=7), quality=75,
                                                                              subsampling=1)]):
        outs, status, *_ = _decode_batch(decoder, files)
        assert (status == 0).all(), status
        for f, got in zip(files, outs):
            assert len(f) > 0 and np.array_equal(got, _pillow(f)), (len(f), got.shape)


def test_committed_golden_vectors(decoder):
    import json
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "jpeg_vectors.npz"))
    meta = json.loads(str(z["meta"]))
    files = [z[m["key"] + "_file"].tobytes() for m in meta]
    outs, status, *_ = _decode_batch(decoder, files)
    assert (status == 0).all()
    for m, got in zip(meta, outs):
        assert np.array_equal(got, z[m["key"] + "_rgb"]), m


def test_two_calls_in_a_row_reuse_the_workspace(decoder):
    a = [_encode(_picture(96, 128, seed=s), quality=80) for s in range(4)]
    b = [_encode(_picture(64, 48, seed=10 + s), quality=95, subsampling=0) for s in range(7)]
    for files in (a, b, a):
        outs, status, *_ = _decode_batch(decoder, files)
        assert (status == 0).all()
        for f, got in zip(files, outs):
            assert np.array_equal(got, _pillow(f))


def test_batch_of_256_at_the_bench_size(decoder):
    files = [_encode(_picture(224, 224, seed=s), quality=90) for s in range(16)] * 16
    outs, status, *_ = _decode_batch(decoder, files)
    assert (status == 0).all()
    refs = [_pillow(f) for f in files[:16]]
    for i, got in enumerate(outs):
        assert np.array_equal(got, refs[i % 16])


def test_truncated_stream_is_flagged_not_fatal(decoder):
    good = _encode(_picture(64, 64, seed=1), quality=90)
    from oracle import jpeg_np as J
    hd = J.parse(good)
    cut = good[:hd.scan_start + (hd.scan_end - hd.scan_start) // 2] + b"\xff\xd9"
    outs, status, *_ = _decode_batch(decoder, [good, cut])
    assert status[0] == 0 and (status[1] & 1) == 1
    assert np.array_equal(outs[0], _pillow(good))


def test_unsupported_files_are_refused_before_any_launch(decoder):
    from mlmicroservicetemplate_b200 import backend as B
    prog = _encode(_picture(32, 32, seed=2), quality=80, progressive=True)
    assert B.jpeg_inspect(prog) is None
    slab = torch.zeros(32 * 32 * 3, dtype=torch.uint8, device="cuda")
    with pytest.raises(B.B2PError) as e:
        decoder.decode([prog], slab, slab.numel(), [0])
    assert e.value.code == B.E_UNSUPPORTED
    with pytest.raises(B.B2PError):
        decoder.decode([_encode(_picture(32, 32, seed=2), quality=80)], slab, 100, [0])     # slab too small


def test_damaged_files_never_take_the_decoder_down(decoder):
    """Corrupt entropy-coded data, corrupt Huffman / quantisation tables: the kernels must stay inside their buffers
    (a stray access would poison the CUDA context and fail everything after it), flag what they can, and the next
    clean file must still decode bit for bit."""
    from mlmicroservicetemplate_b200 import backend as B
    rng = np.random.default_rng(99)
    clean = _encode(_picture(72, 88, seed=3), quality=88)
    seeds = [clean, _encode(_picture(64, 64, seed=4), quality=70, subsampling=0, optimize=True),
             _encode(_picture(48, 80, seed=5), quality=92, restart_marker_blocks=3), _encode(_picture(56, 56, seed=6, grey=True), quality=80)]
    accepted = refused = flagged = 0
    for trial in range(240):
        data = bytearray(seeds[trial % len(seeds)])
        lo = 2 if trial % 3 == 0 else len(data) // 3          # sometimes the tables, mostly the entropy-coded data
        for _ in range(int(rng.integers(1, 12))):
            data[int(rng.integers(lo, len(data) - 2))] = int(rng.integers(0, 256))
        data = bytes(data)
        if B.jpeg_inspect(data) is None:
            refused += 1
            continue
        try:
            outs, status, *_ = _decode_batch(decoder, [data, clean])
        except B.B2PError as e:                               # e.g. a Huffman table that over-subscribes its code space
            assert e.code in (-1, B.E_UNSUPPORTED), e
            refused += 1
            continue
        accepted += 1
        flagged += int(status[0] != 0)
        assert status[1] == 0 and np.array_equal(outs[1], _pillow(clean))
    assert accepted > 60 and flagged > 0, (accepted, refused, flagged)
