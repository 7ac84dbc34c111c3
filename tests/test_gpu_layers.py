"""Layer parity on the GPU: every op kind of the layer program, called through the C ABI
(b2p_create / b2p_forward), against a plain torch fp32 CPU reference of the same op fed the
same bf16-rounded inputs and weights.  Shapes are the ResNet-18/50 shapes of SURVEY §8(d)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf16r(t):
    return t.to(torch.bfloat16).float()


def _run_conv(n, hwc, oc, k, stride, pad, relu, residual, seed, pad_hi=None, sub_batch=0):
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(seed)
    h, w, c = hwc
    x = _bf16r(torch.randn(n, h, w, c, generator=g))
    wt = _bf16r(torch.randn(oc, c, k, k, generator=g) / (c * k * k) ** 0.5)
    bias = torch.randn(oc, generator=g) * 0.1
    prog = P.single_conv_program(wt, bias, hwc, stride, pad, relu, with_residual=residual, pad_hi=pad_hi)
    rep = B.Replica(0, prog, max_batch=n, sub_batch=sub_batch)
    xd = x.to(torch.bfloat16).cuda()
    ph = pad if pad_hi is None else pad_hi
    oh = (h + pad + ph - k) // stride + 1
    ow = (w + pad + ph - k) // stride + 1
    out = torch.empty(n, oh, ow, oc, dtype=torch.bfloat16, device="cuda")
    rep.forward(xd, n, out)
    torch.cuda.synchronize()
    xin = F.pad(x.permute(0, 3, 1, 2), (pad, ph, pad, ph))
    ref = F.conv2d(xin, wt, bias, stride=stride)
    if residual:
        ref = ref + x.permute(0, 3, 1, 2)
    if relu:
        ref = ref.relu()
    ref = ref.permute(0, 2, 3, 1)
    got = out.float().cpu()
    rep.close()
    err = (got - ref).abs()
    # the result is fp32-accumulated and rounded ONCE to bf16: half a bf16 ulp (<= 2^-8 |ref|) plus accumulation-order
    # noise near zero
    tol = 2.0 ** -8 * ref.abs() + 2e-3
    bad = (err > tol).sum().item()
    print(f"conv n={n} hwc={hwc} oc={oc} k={k} s={stride} p={pad}: max_err={err.max().item():.4g} "
          f"max_ref={ref.abs().max().item():.4g} bad={bad}/{err.numel()}")
    assert bad == 0, f"{bad} elements off; worst {err.max().item()}"


CONV_CASES = [
    # n, (h, w, c), oc, k, stride, pad, relu, residual
    (2, (56, 56, 64), 64, 1, 1, 0, True, False),          # 1x1 tiled, BLOCK_N 64
    (2, (56, 56, 64), 256, 1, 1, 0, False, False),        # 1x1 expand
    (2, (56, 56, 256), 256, 1, 1, 0, True, True),         # residual + relu
    (3, (28, 28, 512), 128, 1, 1, 0, True, False),        # K = 512 (8 k-blocks, ring wraps)
    (2, (56, 56, 64), 64, 3, 1, 1, True, False),          # 3x3, 56x56x64: patch-resident kernel
    (3, (56, 56, 64), 64, 3, 1, 1, True, True),           # same + residual (ResNet-18 BasicBlock conv2)
    (1, (56, 56, 64), 64, 3, 1, 1, False, False),
    (2, (28, 28, 64), 64, 3, 1, 1, True, False),          # 3x3 generic im2col path
    (2, (28, 28, 128), 128, 3, 1, 1, True, False),        # 3x3, 28x28x128: patch-resident kernel (streamed weights)
    (1, (28, 28, 128), 128, 3, 1, 1, False, False),
    (45, (28, 28, 128), 128, 3, 1, 1, True, False),       # 315 tiles: every CTA wraps its patch / weight / TMEM rings
    (2, (56, 56, 128), 128, 3, 2, 1, True, False),        # 3x3 stride 2
    (2, (56, 56, 256), 512, 1, 2, 0, False, False),       # 1x1 stride 2 downsample
    (5, (7, 7, 512), 512, 3, 1, 1, True, True),           # tiny spatial, tiles span images, M tail
    (3, (14, 14, 256), 256, 3, 1, 1, True, False),
    (2, (14, 14, 1024), 2048, 1, 2, 0, False, False),
]


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: f"n{c[0]}_{c[1][0]}x{c[1][2]}_oc{c[2]}_k{c[3]}s{c[4]}")
def test_conv(case, cuda_device):
    n, hwc, oc, k, stride, pad, relu, residual = case
    _run_conv(n, hwc, oc, k, stride, pad, relu, residual, seed=hash(case) % 1000)


def test_conv_stem_s2d_shape(cuda_device):
    # the 4x4 / pad (2,1) / 16-channel conv the stem is rewritten into (32-byte swizzle path)
    _run_conv(2, (112, 112, 16), 64, 4, 1, 2, True, False, seed=7, pad_hi=1)


# Batch sizes for the pooled variant pick every segment length of the launcher's cost model on a 148-SM part
# (1, 2, 4, 7, 14 and 28 pooled rows per segment for n = 1, 3, 8, 18, 37 and 74): the TMEM slot ring wraps differently
# in each.
@pytest.mark.parametrize("pool,n", [(False, 1), (False, 3), (True, 1), (True, 3), (True, 8), (True, 18), (True, 37), (True, 74)],
                         ids=lambda v: f"pool{int(v)}" if isinstance(v, bool) else f"n{v}")
def test_stem_dedicated_kernel(pool, n, cuda_device):
    """OP_STEM_CONV: 4x4/pad(2,1) conv over the s2d image + bias + ReLU (+ fused 3x3/s2/p1 max pool)."""
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(31 + n)
    x = _bf16r(torch.randn(n, 112, 112, 16, generator=g))
    w = _bf16r(torch.randn(64, 16, 4, 4, generator=g) / 16.0)
    b = torch.randn(64, generator=g) * 0.1
    rep = B.Replica(0, P.single_stem_program(w, b, pool), max_batch=n)
    side = 56 if pool else 112
    out = torch.full((n, side, side, 64), float("nan"), dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    ref = F.conv2d(F.pad(x.permute(0, 3, 1, 2), (2, 1, 2, 1)), w, b).relu()
    if pool:
        ref = F.max_pool2d(ref, 3, 2, 1)
    ref = ref.permute(0, 2, 3, 1)
    got = out.float().cpu()
    rep.close()
    err = (got - ref).abs()
    bad = (err > 2.0 ** -7 * ref.abs() + 2e-2).sum().item()
    print(f"stem pool={pool} n={n}: max_err={err.max().item():.4g} max_ref={ref.abs().max().item():.4g} bad={bad}/{err.numel()}")
    assert not torch.isnan(got).any()
    assert bad == 0


@pytest.mark.parametrize("k1,n2,hw,n,sub", [(64, 64, 56, 2, 1), (128, 128, 28, 3, 1), (256, 256, 14, 5, 1), (64, 128, 56, 1, 1),
                                              (64, 128, 56, 3, 2), (128, 128, 28, 2, 2)])
def test_conv_pair_fused(k1, n2, hw, n, sub, cuda_device):
    """OP_CONV_PAIR (conv3 + residual + ReLU -> next conv1 + ReLU in one kernel) against the two convs in torch.
    Both outputs are exported through an identity 1x1 conv (exact in bf16) because a program has one output."""
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(k1 + n2 + hw)
    n1 = 4 * k1
    t2 = _bf16r(torch.randn(n, hw, hw, k1, generator=g))
    w_id = _bf16r(torch.randn(n1, k1, 1, 1, generator=g) / k1 ** 0.5)          # makes the "identity" tensor from t2
    w3 = _bf16r(torch.randn(n1, k1, 1, 1, generator=g) / k1 ** 0.5)
    b3 = torch.randn(n1, generator=g) * 0.1
    w1 = _bf16r(torch.randn(n2, n1, 1, 1, generator=g) / n1 ** 0.5)
    b1 = torch.randn(n2, generator=g) * 0.1

    def build(export_second):
        bld = P._Builder("pair")
        p = bld.p
        p.in_elems = hw * hw * k1
        idn, _ = bld.conv(B.SLOT_INPUT, (hw, hw, k1), w_id, torch.zeros(n1), 1, 0, False)
        t2s, _ = bld.conv(B.SLOT_INPUT, (hw, hw, k1), torch.eye(k1).view(k1, k1, 1, 1), torch.zeros(k1), 1, 0, False)
        wi, bi = p.add_blob(P.pack_conv_weight(w3)), p.add_blob(P.pad_bias(b3, n1))
        wi2, bi2 = p.add_blob(P.pack_conv_weight(w1)), p.add_blob(P.pad_bias(b1, n2))
        # sub = 2: the first output keeps only its even rows / columns (a compact [hw/2, hw/2, n1] tensor)
        out1, out2 = p.alloc((hw // sub) * (hw // sub) * n1), p.alloc(hw * hw * n2)
        p.ops.append(P._mk_op(B.OP_CONV_PAIR, t2s, out1, idn, (hw, hw, k1), (hw, hw, n1), (1, 1), 1, (0, 0), 1, wi, bi,
                              reserved=(out2, wi2, bi2, n2), reserved2=sub))
        c = n2 if export_second else n1
        side = hw if export_second else hw // sub
        bld.conv(out2 if export_second else out1, (side, side, c), torch.eye(c).view(c, c, 1, 1), torch.zeros(c), 1, 0, False,
                 dst=B.SLOT_OUTPUT)
        p.out_elems = side * side * c
        return p

    x = t2.to(torch.bfloat16).cuda()
    outs = []
    for second in (False, True):
        rep = B.Replica(0, build(second), max_batch=n)
        c = n2 if second else n1
        side = hw if second else hw // sub
        out = torch.full((n, side, side, c), float("nan"), dtype=torch.bfloat16, device="cuda")
        rep.forward(x, n, out)
        torch.cuda.synchronize()
        outs.append(out.float().cpu())
        rep.close()
    xin = t2.permute(0, 3, 1, 2)
    idn = _bf16r(F.conv2d(xin, w_id))
    ref1 = (F.conv2d(xin, w3, b3) + idn).relu()
    ref2 = F.conv2d(_bf16r(ref1), w1, b1).relu().permute(0, 2, 3, 1)
    ref1 = ref1.permute(0, 2, 3, 1)[:, ::sub, ::sub, :]
    e1, e2 = (outs[0] - ref1).abs(), (outs[1] - ref2).abs()
    print(f"pair k1={k1} n2={n2}: out1 max_err={e1.max().item():.4g} out2 max_err={e2.max().item():.4g}")
    assert not torch.isnan(outs[0]).any() and not torch.isnan(outs[1]).any()
    # the residual operand is itself a bf16-rounded conv output (|idn| up to ~5): where the GPU and torch round it to
    # different neighbours (fp32 accumulation order at a tie) the sum moves by one ulp OF THE OPERAND, 1.6e-2 at 2-4,
    # whatever the size of the sum - hence the absolute term
    assert (e1 > 2.0 ** -7 * ref1.abs() + 2e-2).sum().item() == 0
    assert (e2 > 2.0 ** -6 * ref2.abs() + 3e-2).sum().item() == 0      # second conv sees a bf16-rounded input


@pytest.mark.parametrize("c_main,c_sc,oc,hw_out,stride2,n", [(64, 64, 256, 56, 1, 2), (128, 256, 512, 28, 2, 3), (512, 1024, 2048, 7, 2, 5)])
def test_conv_with_folded_shortcut(c_main, c_sc, oc, hw_out, stride2, n, cuda_device):
    """OP_CONV with a second GEMM source: out = relu(W3 t2 + Wd x(strided) + b): the projection shortcut as extra K."""
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(c_main + c_sc)
    hw_in = hw_out * stride2
    x = _bf16r(torch.randn(n, hw_in, hw_in, c_sc, generator=g))                  # external input = the block input
    w_t2 = _bf16r(torch.randn(c_main, c_sc, 1, 1, generator=g) / c_sc ** 0.5)    # makes t2 (at output resolution) from x
    w3 = _bf16r(torch.randn(oc, c_main, 1, 1, generator=g) / c_main ** 0.5)
    wd = _bf16r(torch.randn(oc, c_sc, 1, 1, generator=g) / c_sc ** 0.5)
    b3, bd = torch.randn(oc, generator=g) * 0.1, torch.randn(oc, generator=g) * 0.1
    bld = P._Builder("dual")
    p = bld.p
    p.in_elems = hw_in * hw_in * c_sc
    # the second source must live in a slot: copy x through an identity conv, and derive t2 from it
    xs, _ = bld.conv(B.SLOT_INPUT, (hw_in, hw_in, c_sc), torch.eye(c_sc).view(c_sc, c_sc, 1, 1), torch.zeros(c_sc), 1, 0, False)
    t2, hwc2 = bld.conv(xs, (hw_in, hw_in, c_sc), w_t2, torch.zeros(c_main), stride2, 0, False)
    bld.conv(t2, hwc2, w3, b3, 1, 0, True, second=(xs, wd, bd, stride2), dst=B.SLOT_OUTPUT)
    p.out_elems = hw_out * hw_out * oc
    rep = B.Replica(0, p, max_batch=n)
    out = torch.full((n, hw_out, hw_out, oc), float("nan"), dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    rep.close()
    xin = x.permute(0, 3, 1, 2)
    t2_ref = _bf16r(F.conv2d(xin, w_t2, stride=stride2))
    ref = (F.conv2d(t2_ref, w3, b3) + F.conv2d(xin, wd, bd, stride=stride2)).relu().permute(0, 2, 3, 1)
    err = (out.float().cpu() - ref).abs()
    print(f"dual-source conv {c_main}+{c_sc}->{oc} s{stride2}: max_err={err.max().item():.4g}")
    assert (err > 2.0 ** -8 * ref.abs() + 2e-3).sum().item() == 0


def test_conv_sub_batched(cuda_device):
    # same op run in chunks of 2 images: exercises the external-source re-encode and chunk offsets
    _run_conv(5, (28, 28, 128), 128, 3, 1, 1, True, False, seed=11, sub_batch=2)


def test_linear(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(3)
    n, fin, classes = 70, 2048, 1000
    x = _bf16r(torch.randn(n, fin, generator=g))
    w = _bf16r(torch.randn(classes, fin, generator=g) / fin ** 0.5)
    b = torch.randn(classes, generator=g)
    rep = B.Replica(0, P.single_linear_program(w, b), max_batch=n)
    out = torch.full((n, classes), float("nan"), device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, out)
    torch.cuda.synchronize()
    ref = x @ w.t() + b
    err = (out.cpu() - ref).abs().max().item()
    print("linear max_err", err)
    rep.close()
    assert err < 2e-3


def _one_op_program(kind, in_hwc, out_hwc, in_elems, k=(1, 1), stride=1, pad=(0, 0)):
    from mlmicroservicetemplate_b200 import backend as B, program as P
    p = P.Program(name="one_op")
    p.in_elems = in_elems
    p.out_elems = out_hwc[0] * out_hwc[1] * out_hwc[2]
    p.ops.append(P._mk_op(kind, B.SLOT_INPUT, B.SLOT_OUTPUT, in_hwc=in_hwc, out_hwc=out_hwc, k=k, stride=stride, pad=pad))
    return p


def test_maxpool(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n = 3
    x = torch.randn(n, 112, 112, 64).to(torch.bfloat16)
    rep = B.Replica(0, _one_op_program(B.OP_MAXPOOL, (112, 112, 64), (56, 56, 64), 112 * 112 * 64, (3, 3), 2, (1, 1)), n)
    out = torch.empty(n, 56, 56, 64, dtype=torch.bfloat16, device="cuda")
    rep.forward(x.cuda(), n, out)
    torch.cuda.synchronize()
    ref = F.max_pool2d(x.float().permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1)
    rep.close()
    assert torch.equal(out.float().cpu(), ref)


def test_avgpool(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n = 5
    x = torch.randn(n, 7, 7, 2048).to(torch.bfloat16)
    rep = B.Replica(0, _one_op_program(B.OP_AVGPOOL, (7, 7, 2048), (1, 1, 2048), 49 * 2048), n)
    out = torch.empty(n, 2048, dtype=torch.bfloat16, device="cuda")
    rep.forward(x.cuda(), n, out)
    torch.cuda.synchronize()
    ref = x.float().mean(dim=(1, 2))
    rep.close()
    assert (out.float().cpu() - ref).abs().max().item() < 4e-3


def test_stem_pack(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n = 2
    x = torch.randn(n, 3, 224, 224).to(torch.bfloat16)
    rep = B.Replica(0, _one_op_program(B.OP_STEM_PACK, (224, 224, 3), (112, 112, 16), 3 * 224 * 224), n)
    out = torch.empty(n, 112, 112, 16, dtype=torch.bfloat16, device="cuda")
    rep.forward(x.cuda(), n, out)
    torch.cuda.synchronize()
    ref = torch.zeros(n, 112, 112, 16)
    xf = x.float()
    for di in range(2):
        for dj in range(2):
            for c in range(3):
                ref[..., (di * 2 + dj) * 3 + c] = xf[:, c, di::2, dj::2]
    rep.close()
    assert torch.equal(out.float().cpu(), ref)


def test_softmax_topk(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    n, classes, k = 37, 1000, 5
    g = torch.Generator().manual_seed(5)
    logits = torch.randn(n, classes, generator=g) * 3
    logits[3, 10] = logits[3, 500] = 50.0     # an exact tie above every other logit: lowest index must win
    rep = B.Replica(0, _one_op_program(B.OP_AVGPOOL, (1, 1, 8), (1, 1, 8), 8), n)   # any handle will do
    probs = torch.empty(n, k, device="cuda")
    idx = torch.empty(n, k, dtype=torch.int32, device="cuda")
    rep.softmax_topk(logits.cuda(), n, classes, k, probs, idx)
    torch.cuda.synchronize()
    ref_p, ref_i = torch.softmax(logits, 1).topk(k, dim=1)
    rep.close()
    assert idx[3, 0].item() == 10 and idx[3, 1].item() == 500
    mask = torch.ones(n, dtype=torch.bool)
    mask[3] = False
    assert torch.equal(idx.cpu().long()[mask], ref_i[mask])
    assert torch.allclose(probs.cpu()[mask], ref_p[mask], rtol=1e-5, atol=1e-8)
    assert abs(probs[3, 0].item() - 0.5) < 1e-5 and abs(probs[3, 1].item() - 0.5) < 1e-5


@pytest.mark.parametrize("case", [
    (8, (56, 56, 256), 256, 1, 1, 0, True, True),         # 196 tiles of 128 x 256: 148 + 48, residual stream
    (112, (14, 14, 256), 256, 3, 1, 1, True, False),      # 172 tiles, im2col operand
    (5, (56, 56, 64), 256, 1, 1, 0, False, False),        # 123 tiles (< one wave): nothing to split, flag must be inert
], ids=["1x1_res", "3x3_im2col", "single_wave"])
def test_conv_with_tail_split(case, cuda_device, monkeypatch):
    """B2P_TAIL_SPLIT=1: the tiles of the partly filled last wave are computed as 64-column pieces by other CTAs."""
    monkeypatch.setenv("B2P_TAIL_SPLIT", "1")
    n, hwc, oc, k, stride, pad, relu, residual = case
    _run_conv(n, hwc, oc, k, stride, pad, relu, residual, seed=77)


@pytest.mark.parametrize("proj,n", [(False, 2), (False, 3), (False, 5)])
def test_bottleneck56_fused_block(proj, n, cuda_device):
    """OP_BOTTLENECK - conv1x1 -> conv3x3 -> conv1x1 + identity shortcut of a 56x56 bottleneck in one kernel - against
    torch fp32 with the same bf16 roundings of the two intermediates (SP/torchvision/models/resnet.py:143-163)."""
    from mlmicroservicetemplate_b200 import backend as B, program as P
    g = torch.Generator().manual_seed(56 + int(proj) + n)
    cin = 64 if proj else 256
    x = _bf16r(torch.randn(n, 56, 56, cin, generator=g))
    w1 = _bf16r(torch.randn(64, cin, 1, 1, generator=g) / cin ** 0.5)
    w2 = _bf16r(torch.randn(64, 64, 3, 3, generator=g) / 576 ** 0.5)
    w3 = _bf16r(torch.randn(256, 64, 1, 1, generator=g) / 64 ** 0.5)
    wd = _bf16r(torch.randn(256, cin, 1, 1, generator=g) / cin ** 0.5)
    b1, b2, b3, bd = (torch.randn(c, generator=g) * 0.1 for c in (64, 64, 256, 256))
    p = P.Program(name="bneck")
    p.in_elems = 56 * 56 * cin
    p.out_elems = 56 * 56 * 256
    xs, out = p.alloc(56 * 56 * cin), p.alloc(56 * 56 * 256)
    # the block works slot to slot: a 1x1 max pool is an exact copy in and out of the workspace
    p.ops.append(P._mk_op(B.OP_MAXPOOL, B.SLOT_INPUT, xs, in_hwc=(56, 56, cin), out_hwc=(56, 56, cin), k=(1, 1), stride=1))
    P.bottleneck_op(p, xs, out, 56, cin, w1, b1, w2, b2, w3, b3, shortcut=(wd, bd) if proj else None)
    p.ops.append(P._mk_op(B.OP_MAXPOOL, out, B.SLOT_OUTPUT, in_hwc=(56, 56, 256), out_hwc=(56, 56, 256), k=(1, 1), stride=1))
    rep = B.Replica(0, p, max_batch=n)
    got = torch.full((n, 56, 56, 256), float("nan"), dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, got)
    torch.cuda.synchronize()
    got2 = torch.full((n, 56, 56, 256), float("nan"), dtype=torch.bfloat16, device="cuda")
    rep.forward(x.to(torch.bfloat16).cuda(), n, got2)          # second call: every barrier phase / buffer reused
    torch.cuda.synchronize()
    rep.close()
    xin = x.permute(0, 3, 1, 2)
    t1 = _bf16r(F.conv2d(xin, w1, b1).relu())
    t2 = _bf16r(F.conv2d(t1, w2, b2, padding=1).relu())
    ref = F.conv2d(t2, w3, b3) + (F.conv2d(xin, wd, bd) if proj else xin)
    ref = ref.relu().permute(0, 2, 3, 1)
    g_ = got.float().cpu()
    err = (g_ - ref).abs()
    # two bf16-rounded intermediates upstream: an input ulp flips now and then, so allow two result ulps
    bad = (err > 2.0 ** -6 * ref.abs() + 3e-2).sum().item()
    print(f"bottleneck56 proj={proj} n={n}: max_err={err.max().item():.4g} max_ref={ref.abs().max().item():.4g} bad={bad}")
    assert not torch.isnan(g_).any()
    assert bad == 0
    assert torch.equal(got, got2)
