"""ViT-B/16 on the GPU through the C ABI: the ViT-specific ops one by one against torch fp32 on the
same bf16-rounded operands, then the whole classifier against the oracle's CPU fp32 predict path
(SP/torchvision/models/vision_transformer.py)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf16r(t):
    return t.to(torch.bfloat16).float()


def _one(kind, in_hwc, out_hwc, in_elems, blobs=(), flag0=0, **kw):
    from mlmicroservicetemplate_b200 import backend as B, program as P
    p = P.Program(name="one_op")
    p.in_elems = in_elems
    p.out_elems = out_hwc[0] * out_hwc[1] * out_hwc[2]
    idx = [p.add_blob(b) for b in blobs]
    p.ops.append(P._mk_op(kind, B.SLOT_INPUT, B.SLOT_OUTPUT, in_hwc=in_hwc, out_hwc=out_hwc, flag0=flag0,
                          weight=idx[0] if idx else -1, bias=idx[1] if len(idx) > 1 else -1, **kw))
    return p


def test_patchify(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n = 3
    x = torch.randn(n, 3, 224, 224).to(torch.bfloat16)
    rep = B.Replica(0, _one(B.OP_PATCHIFY, (224, 224, 3), (196, 1, 768), 3 * 224 * 224), n)
    out = torch.empty(n, 196, 768, dtype=torch.bfloat16, device="cuda")
    rep.forward(x.cuda(), n, out)
    torch.cuda.synchronize()
    ref = F.unfold(x.float(), kernel_size=16, stride=16).transpose(1, 2)      # [n, 196, 3*16*16] in (c,ky,kx) order
    rep.close()
    assert torch.equal(out.float().cpu(), ref)


def test_tokens(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n = 3
    proj = _bf16r(torch.randn(n, 196, 768))
    pos, cls = torch.randn(197 * 768) * 0.02, torch.randn(768)
    rep = B.Replica(0, _one(B.OP_TOKENS, (197, 1, 768), (197, 1, 768), 196 * 768, blobs=(pos, cls)), n)
    out = torch.empty(n, 197, 768, dtype=torch.bfloat16, device="cuda")
    rep.forward(proj.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    ref = torch.cat([cls.view(1, 1, 768).expand(n, 1, 768), proj], 1) + pos.view(1, 197, 768)
    rep.close()
    assert torch.equal(out.cpu(), ref.to(torch.bfloat16))


@pytest.mark.parametrize("cls_only", [False, True])
def test_layernorm(cls_only, cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n = 5
    x = _bf16r(torch.randn(n, 197, 768) * 2 + 0.3)
    g, b = torch.rand(768) + 0.5, torch.randn(768) * 0.1
    out_hwc = (1, 1, 768) if cls_only else (197, 1, 768)
    rep = B.Replica(0, _one(B.OP_LAYERNORM, (197, 1, 768), out_hwc, 197 * 768, blobs=(g, b), flag0=int(cls_only)), n)
    out = torch.empty((n, 768) if cls_only else (n, 197, 768), dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    ref = F.layer_norm(x, (768,), g, b, eps=1e-6)
    ref = ref[:, 0] if cls_only else ref
    rep.close()
    assert (out.float().cpu() - ref).abs().max().item() <= 2.0 ** -7 * ref.abs().max().item() + 1e-3


@pytest.mark.parametrize("n", [1, 3])
def test_attention(n, cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    g = torch.Generator().manual_seed(40 + n)
    qkv = _bf16r(torch.randn(n, 197, 2304, generator=g))
    rep = B.Replica(0, _one(B.OP_ATTENTION, (197, 1, 768), (197, 1, 768), 197 * 2304, flag0=12), n)
    out = torch.full((n, 197, 768), float("nan"), dtype=torch.bfloat16, device="cuda")
    rep.forward(qkv.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    q, k, v = [t.view(n, 197, 12, 64).transpose(1, 2) for t in qkv.split(768, dim=2)]
    ref = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(n, 197, 768)
    got = out.float().cpu()
    rep.close()
    err = (got - ref).abs().max().item()
    print(f"attention n={n}: max_err={err:.4g} max_ref={ref.abs().max().item():.4g}")
    assert not torch.isnan(got).any()
    assert err <= 2e-2


def test_linear_gelu_residual(cuda_device):
    """fc1 + GELU and fc2 + residual as 1x1 OP_CONVs over a 197x1 token 'image'."""
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(50)
    n = 3
    x = _bf16r(torch.randn(n, 197, 1, 768, generator=g))
    w = _bf16r(torch.randn(3072, 768, 1, 1, generator=g) / 768 ** 0.5)
    b = torch.randn(3072, generator=g) * 0.1
    prog = P.single_conv_program(w, b, (197, 1, 768), 1, 0, B.ACT_GELU)
    rep = B.Replica(0, prog, max_batch=n)
    out = torch.empty(n, 197, 1, 3072, dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    ref = F.gelu(x.view(n * 197, 768) @ w.view(3072, 768).t() + b).view(n, 197, 1, 3072)
    rep.close()
    err = (out.float().cpu() - ref).abs()
    assert (err > 2.0 ** -8 * ref.abs() + 2e-3).sum().item() == 0
    w2 = _bf16r(torch.randn(768, 768, 1, 1, generator=g) / 768 ** 0.5)
    prog = P.single_conv_program(w2, torch.zeros(768), (197, 1, 768), 1, 0, B.ACT_NONE, with_residual=True)
    rep = B.Replica(0, prog, max_batch=n)
    out = torch.empty(n, 197, 1, 768, dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    ref = (x.view(n * 197, 768) @ w2.view(768, 768).t()).view(n, 197, 1, 768) + x
    rep.close()
    assert ((out.float().cpu() - ref).abs() > 2.0 ** -8 * ref.abs() + 2e-3).sum().item() == 0


def test_vit_b16_end_to_end(cuda_device):
    import numpy as np
    from mlmicroservicetemplate_b200 import backend as B, program as P
    from oracle import images, model as om
    model = om.build_model("vit_b_16")
    prog = P.build_program(model, "vit_b_16")
    assert abs(prog.flops_per_image - 35.128e9) / 35.128e9 < 5e-3
    ims = [images.synthetic_image(700 + i, 224, 224) for i in range(6)] + [images.synthetic_image(710, 480, 640),
                                                                         images.synthetic_image(711, 333, 517)]
    n = len(ims)
    rep = B.Replica(0, prog, max_batch=n)
    offs, total = [], 0
    for im in ims:
        offs.append(total)
        total += (im.size + 255) // 256 * 256
    slab = np.zeros(total, np.uint8)
    geoms = []
    for im, o in zip(ims, offs):
        slab[o:o + im.size] = im.reshape(-1)
        geoms.append((o, im.shape[0], im.shape[1], 224, 224, 0, 0))
    x = torch.empty(n, 3, 224, 224, dtype=torch.bfloat16, device="cuda")
    logits = torch.full((n, 1000), float("nan"), device="cuda")
    rep.preprocess(torch.from_numpy(slab).cuda(), B.make_images(geoms), n, x)
    rep.forward(x, n, logits)
    torch.cuda.synchronize()
    ref, _, _ = om.predict_arrays(ims, mode="direct", k=5, model=model)
    got = logits.cpu()
    rep.close()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item()
    srt = ref.sort(dim=1, descending=True).values
    margins = srt[:, 0] - srt[:, 1]
    agree = got.argmax(1) == ref.argmax(1)
    print(f"vit_b_16 n={n}: logits max_abs_err={err:.4g} max|logit|={scale:.4g} rel={err / scale:.3g} "
          f"top1_agree={agree.float().mean().item():.3f} min_margin={margins.min().item():.4g} "
          f"distinct_top1={ref.argmax(1).unique().numel()}")
    assert not torch.isnan(got).any()
    assert err <= 6e-2 and err / scale <= 3e-2
    # top-1 must agree wherever the fp32 margin exceeds the observed error band
    assert bool(agree[margins > 2 * err].all())


def test_vit_b16_b64_vs_oracle(cuda_device):
    """BASELINE.json configs[4] per-GPU size: 64 distinct uint8 images through the whole chain (b2p_predict) against
    the oracle's CPU fp32 path, every image: logits, top-1 where the fp32 margin exceeds the error band, top-k
    probabilities."""
    import numpy as np
    from mlmicroservicetemplate_b200 import backend as B, program as P
    from oracle import images, model as om
    model = om.build_model("vit_b_16")
    prog = P.build_program(model, "vit_b_16")
    ims = images.synthetic_batch(100, 64, 224, 224)
    n = len(ims)
    rep = B.Replica(0, prog, max_batch=n)
    stride = (224 * 224 * 3 + 255) // 256 * 256
    slab = np.zeros(n * stride, np.uint8)
    for i, im in enumerate(ims):
        slab[i * stride:i * stride + im.size] = im.reshape(-1)
    descs = B.make_images([(i * stride, 224, 224, 224, 224, 0, 0) for i in range(n)])
    probs = torch.empty(n, 5, device="cuda")
    idx = torch.empty(n, 5, dtype=torch.int32, device="cuda")
    rep.predict(torch.from_numpy(slab).cuda(), descs, n, 5, probs, idx)
    torch.cuda.synchronize()
    logits = torch.empty(n, 1000, device="cuda")
    x = torch.empty(n, 3, 224, 224, dtype=torch.bfloat16, device="cuda")
    rep.preprocess(torch.from_numpy(slab).cuda(), descs, n, x)
    rep.forward(x, n, logits)
    torch.cuda.synchronize()
    rep.close()
    ref, ref_probs, ref_idx = om.predict_arrays(ims, mode="direct", k=5, model=model)
    got = logits.cpu()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item()
    srt = ref.sort(dim=1, descending=True).values
    margins = srt[:, 0] - srt[:, 1]
    agree = got.argmax(1) == ref.argmax(1)
    print(f"vit_b_16 b64: logits max_abs_err={err:.4g} max|logit|={scale:.4g} rel={err / scale:.3g} "
          f"top1_agree={agree.float().mean().item():.3f} min_margin={margins.min().item():.4g} "
          f"distinct_top1={ref.argmax(1).unique().numel()}")
    assert not torch.isnan(got).any()
    assert err <= 6e-2 and err / scale <= 3e-2
    assert bool(agree[margins > 2 * err].all())
    assert (probs.cpu() - ref_probs).abs().max().item() < 5e-3
