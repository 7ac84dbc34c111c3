"""Diagnostic: decode a batch and report, per image, the status word and the number of bytes that differ from Pillow."""
import io
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from benchmarks.jpeg_decode import picture  # noqa: E402


def main():
    import torch
    from PIL import Image
    from mlmicroservicetemplate_b200 import backend as B
    h = w = int(sys.argv[1]) if len(sys.argv) > 1 else 224
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    sub = int(sys.argv[3]) if len(sys.argv) > 3 else 2          # 0 = 4:4:4, 1 = 4:2:2, 2 = 4:2:0, 3 = greyscale
    noise = float(os.environ.get("DIAG_NOISE", "6"))
    quality = int(os.environ.get("DIAG_QUALITY", "90"))
    files = []
    import numpy as _np

    def picture(h, w, seed):          # same pattern as benchmarks/jpeg_decode.picture with adjustable sensor noise
        rng = _np.random.default_rng(seed)
        yy, xx = _np.mgrid[0:h, 0:w]
        a = _np.stack([127 + 100 * _np.sin(xx / 17.0 + yy / 31.0 + seed), 127 + 90 * _np.cos(xx / 9.0 - yy / 13.0),
                       127 + 120 * _np.sin(xx * yy / 3000.0 + seed)], -1) + rng.normal(0, noise, (h, w, 3))
        return _np.clip(a, 0, 255).astype(_np.uint8)
    for s in range(n):
        b = io.BytesIO()
        if sub == 3:
            Image.fromarray(picture(h, w, s)[..., 0]).save(b, "JPEG", quality=quality)
        else:
            Image.fromarray(picture(h, w, s)).save(b, "JPEG", quality=quality, subsampling=sub)
        files.append(b.getvalue())
    dec = B.JpegDecoder(0)
    per = (h * w * 3 + 255) // 256 * 256
    slab = torch.zeros(per * n, dtype=torch.uint8, device="cuda")
    status = torch.zeros(n, dtype=torch.int32, device="cuda")
    for rep in range(3):
        slab.zero_()
        dec.decode(files, slab, slab.numel(), [i * per for i in range(n)], status=status)
        torch.cuda.synchronize()
        host = slab.cpu().numpy()
        st = status.cpu().numpy()
        bad = []
        for i, f in enumerate(files):
            ref = np.asarray(Image.open(io.BytesIO(f)).convert("RGB"))
            got = host[i * per:i * per + h * w * 3].reshape(h, w, 3)
            d = int((ref != got).sum())
            if d or st[i]:
                rows = np.unique(np.argwhere((ref != got).any(axis=2))[:, 0]) if d else []
                bad.append((i, int(st[i]), d, (int(rows.min()), int(rows.max())) if d else None))
        if os.environ.get("B2P_JPEG_DEBUG_ROUNDS"):
            print("chunk", os.environ.get("B2P_JPEG_CHUNK_BITS"), "noise", noise, "q", quality, "bytes/file", sum(map(len, files)) // n,
                  "rounds per image:", sorted((st >> 8).tolist()))
            break
        print("chunk", os.environ.get("B2P_JPEG_CHUNK_BITS"), "rep", rep, "bad images:", bad[:8], "count", len(bad))


if __name__ == "__main__":
    main()
