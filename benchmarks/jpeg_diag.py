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
    files = []
    for s in range(n):
        b = io.BytesIO()
        Image.fromarray(picture(h, w, s)).save(b, "JPEG", quality=90, subsampling=2)
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
            print("chunk", os.environ.get("B2P_JPEG_CHUNK_BITS"), "rounds per image:", sorted((st >> 8).tolist()))
            break
        print("chunk", os.environ.get("B2P_JPEG_CHUNK_BITS"), "rep", rep, "bad images:", bad[:8], "count", len(bad))


if __name__ == "__main__":
    main()
