"""K1 parity on the GPU through the C ABI (b2p_preprocess) against the oracle's fp32 CPU
preprocess (torch F.interpolate antialias path).  Gates from BASELINE.json north_star:
<= 1e-6 in fp32, <= 1 bf16 ulp after the cast."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-6


def _slab(images, align=256):
    """Packs uint8 HWC images into one slab; returns (uint8 tensor, [(offset, h, w)])."""
    offs, total = [], 0
    for im in images:
        offs.append(total)
        total += (im.size + align - 1) // align * align
    slab = np.zeros(total, np.uint8)
    for im, o in zip(images, offs):
        slab[o:o + im.size] = im.reshape(-1)
    return torch.from_numpy(slab), offs


def _dummy_replica(n):
    from mlmicroservicetemplate_b200 import backend as B, program as P
    p = P.Program(name="k1_only")
    p.in_elems = 8
    p.out_elems = 8
    p.ops.append(P._mk_op(B.OP_AVGPOOL, B.SLOT_INPUT, B.SLOT_OUTPUT, in_hwc=(1, 1, 8), out_hwc=(1, 1, 8)))
    return B.Replica(0, p, max_batch=n)


def _bf16_ulp_diff(a_bf16: torch.Tensor, ref_f32: torch.Tensor) -> int:
    """Distance in bf16 ulps between a and RNE(ref), via the ordered-integer trick."""
    def key(t):
        i = t.view(torch.int16).to(torch.int32)
        return torch.where(i < 0, -(i & 0x7FFF), i)
    return (key(a_bf16) - key(ref_f32.to(torch.bfloat16))).abs().max().item()


def _check(images, mode, unaligned=False):
    from mlmicroservicetemplate_b200 import backend as B
    from oracle import model as om
    n = len(images)
    slab, offs = _slab(images, align=1 if unaligned else 256)
    geoms = []
    for im, o in zip(images, offs):
        rh, rw, top, left = om.geometry(im.shape[0], im.shape[1], mode)
        geoms.append((o, im.shape[0], im.shape[1], rh, rw, top, left))
    descs = B.make_images(geoms)
    rep = _dummy_replica(n)
    slab_d = slab.cuda()
    out32 = torch.full((n, 3, 224, 224), float("nan"), device="cuda")
    out16 = torch.empty((n, 3, 224, 224), dtype=torch.bfloat16, device="cuda")
    rep.preprocess(slab_d, descs, n, out32, out_f32=True)
    rep.preprocess(slab_d, descs, n, out16, out_f32=False)
    torch.cuda.synchronize()
    ref = torch.cat([om.preprocess(im, *g[3:]) for im, g in zip(images, geoms)], 0)
    got = out32.cpu()
    err = (got - ref).abs().max().item()
    exact = (got == ref).float().mean().item()
    ulps = _bf16_ulp_diff(out16.cpu(), ref)
    rep.close()
    print(f"K1 {mode} sizes={[im.shape[:2] for im in images][:4]}..: fp32 max_err={err:.3g} exact={exact:.4f} bf16 ulps={ulps}")
    assert err <= FP32_TOL
    assert ulps <= 1


@pytest.mark.parametrize("hw", [(224, 224), (320, 320), (480, 640), (333, 517), (1080, 1920), (1920, 1920), (225, 300)])
@pytest.mark.parametrize("mode", ["direct", "preset"])
def test_k1_single_size(hw, mode, cuda_device):
    from oracle import images
    if mode == "preset" and hw == (224, 224):
        hw = (256, 256)     # preset of a 256^2 image is a pure centre crop (identity resize)
    _check([images.synthetic_image(17 + hw[0], *hw), images.synthetic_image(18 + hw[1], *hw)], mode)


def test_k1_ragged_batch_unaligned(cuda_device):
    """Ragged sizes in one launch, images packed back to back so rows start at odd addresses."""
    from oracle import images
    sizes = [(224, 224), (231, 517), (640, 480), (301, 227), (720, 1280), (257, 999)]
    _check([images.synthetic_image(100 + i, h, w) for i, (h, w) in enumerate(sizes)], "direct", unaligned=True)
    _check([images.synthetic_image(200 + i, h, w) for i, (h, w) in enumerate(sizes)], "preset", unaligned=True)


def test_k1_noise_image(cuda_device):
    """i.i.d. uint8 noise: the hardest case for accumulation-order differences."""
    g = torch.Generator().manual_seed(9)
    ims = [torch.randint(0, 256, (h, w, 3), generator=g, dtype=torch.uint8).numpy() for h, w in [(768, 1024), (400, 400)]]
    _check(ims, "direct")


def test_k1_rejects_bad_geometry(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    rep = _dummy_replica(1)
    slab = torch.zeros(1024, dtype=torch.uint8, device="cuda")
    out = torch.empty((1, 3, 224, 224), dtype=torch.bfloat16, device="cuda")
    with pytest.raises(B.B2PError):
        rep.preprocess(slab, B.make_images([(0, 10, 10, 100, 100, 0, 0)]), 1, out)   # resized < 224
    with pytest.raises(B.B2PError):
        rep.preprocess(slab, B.make_images([(0, 300, 300, 256, 256, 40, 0)]), 1, out)   # crop leaves the image
    rep.close()


def test_k1_very_wide_sources(cuda_device):
    """Panoramas / 45+ Mpixel-camera widths: the staged window of a source row no longer grows with the image width
    (a CTA takes fewer output columns instead), so no width is refused.  Pillow accepts such uploads (reference
    src/model/model.py:23) and the whole micro-batch used to fail with them."""
    from oracle import images
    _check([images.synthetic_image(61, 230, 9000), images.synthetic_image(62, 300, 12000)], "direct")
    _check([images.synthetic_image(63, 260, 10000), images.synthetic_image(64, 2000, 8192)], "preset")


def test_k1_mixed_crop_only_and_resized_batch(cuda_device):
    """Crop-only images and images that need a resize in one call: the table kernel takes the former, the general
    kernel the latter, each skipping the other's images."""
    from oracle import images
    sizes = [(224, 224), (480, 640), (224, 224), (333, 517), (224, 224)]
    _check([images.synthetic_image(400 + i, h, w) for i, (h, w) in enumerate(sizes)], "direct")
    sizes = [(256, 256), (720, 1280), (256, 300), (300, 256)]
    _check([images.synthetic_image(500 + i, h, w) for i, (h, w) in enumerate(sizes)], "preset")
