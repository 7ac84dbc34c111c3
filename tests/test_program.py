"""Host-side weight packing and layer-program construction, checked on CPU by interpreting the program
with plain torch ops (test-only interpreter) against the torchvision module it was built from."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from mlmicroservicetemplate_b200 import backend as B, program as P


def interpret(prog: P.Program, x_nchw: torch.Tensor) -> torch.Tensor:
    """fp32 torch interpretation of the op list (NHWC slots) using the PACKED blobs."""
    slots = {B.SLOT_INPUT: x_nchw}
    out = None
    for op in prog.ops:
        src = slots[op.src]
        if op.kind == B.OP_STEM_PACK:
            n = src.shape[0]
            y = torch.zeros(n, op.out_h, op.out_w, 16)
            for di in range(2):
                for dj in range(2):
                    for c in range(3):
                        y[..., (di * 2 + dj) * 3 + c] = src[:, c, di::2, dj::2]
        elif op.kind == B.OP_CONV:
            w = prog.blobs[op.weight].float()[: op.out_c]                 # [oc, kh, kw, ic]
            b = prog.blobs[op.bias][: op.out_c]
            xin = F.pad(src.permute(0, 3, 1, 2), (op.pad_lo, op.pad_hi, op.pad_lo, op.pad_hi))
            if op.src2_c > 0:      # weight rows are [kh*kw*in_c | src2_c]
                k1 = op.kh * op.kw * op.in_c
                wflat = w.reshape(op.out_c, -1)
                w, w_sc = wflat[:, :k1].reshape(op.out_c, op.kh, op.kw, op.in_c), wflat[:, k1:]
            y = F.conv2d(xin, w.permute(0, 3, 1, 2), b, stride=op.stride).permute(0, 2, 3, 1)
            if op.src2_c > 0:
                x2 = slots[op.src2].permute(0, 3, 1, 2)
                y = y + F.conv2d(x2, w_sc.reshape(op.out_c, op.src2_c, 1, 1), stride=op.src2_stride).permute(0, 2, 3, 1)
            if op.res != B.SLOT_NONE:
                y = y + slots[op.res]
            if op.relu == 1:
                y = y.relu()
            elif op.relu == 2:
                y = F.gelu(y)
        elif op.kind == B.OP_CONV_PAIR:
            wflat = prog.blobs[op.weight].float()[: op.out_c].reshape(op.out_c, -1)
            b = prog.blobs[op.bias][: op.out_c]
            y = F.conv2d(src.permute(0, 3, 1, 2), wflat[:, : op.in_c].reshape(op.out_c, op.in_c, 1, 1), b).permute(0, 2, 3, 1)
            if op.src2_c > 0:
                y = y + F.conv2d(slots[op.src2].permute(0, 3, 1, 2), wflat[:, op.in_c:].reshape(op.out_c, op.src2_c, 1, 1),
                                 stride=op.src2_stride).permute(0, 2, 3, 1)
            if op.res != B.SLOT_NONE:
                y = y + slots[op.res]
            y = y.relu()
            dst2, wi2, bi2, n2 = op.reserved[0], op.reserved[1], op.reserved[2], op.reserved[3]
            w2 = prog.blobs[wi2].float()[:n2]
            # the fused kernel feeds the second conv the bf16-rounded first result (as the un-fused path would)
            y2 = F.conv2d(y.permute(0, 3, 1, 2), w2.permute(0, 3, 1, 2), prog.blobs[bi2][:n2]).relu().permute(0, 2, 3, 1)
            assert y2[0].numel() <= prog.slot_elems[dst2]
            slots[dst2] = y2
            if op.reserved2 == 2:          # only the even rows / columns of the first result are kept
                y = y[:, ::2, ::2, :].contiguous()
        elif op.kind == B.OP_BOTTLENECK:
            mid = op.reserved2
            xin = src.permute(0, 3, 1, 2)
            w1 = prog.blobs[op.weight].float()[:mid].permute(0, 3, 1, 2)
            w2 = prog.blobs[op.reserved[0]].float()[:mid].permute(0, 3, 1, 2)
            w3 = prog.blobs[op.reserved[2]].float()[: op.out_c].reshape(op.out_c, -1)
            t1 = F.conv2d(xin, w1, prog.blobs[op.bias][:mid]).relu()
            t2 = F.conv2d(t1, w2, prog.blobs[op.reserved[1]][:mid], padding=1).relu()
            y = F.conv2d(t2, w3[:, :mid].reshape(op.out_c, mid, 1, 1), prog.blobs[op.reserved[3]][: op.out_c])
            if op.res == B.SLOT_NONE:     # projection shortcut folded into conv3 as extra K columns
                y = y + F.conv2d(xin, w3[:, mid:].reshape(op.out_c, op.in_c, 1, 1))
            else:
                y = y + xin
            y = y.relu().permute(0, 2, 3, 1)
        elif op.kind == B.OP_STEM_CONV:
            w = prog.blobs[op.weight].float()[: op.out_c]
            b = prog.blobs[op.bias][: op.out_c]
            y = F.conv2d(F.pad(src.permute(0, 3, 1, 2), (2, 1, 2, 1)), w.permute(0, 3, 1, 2), b).relu()
            if op.reserved[0]:
                y = F.max_pool2d(y, 3, 2, 1)
            y = y.permute(0, 2, 3, 1)
        elif op.kind == B.OP_PATCHIFY:
            y = F.unfold(src, kernel_size=16, stride=16).transpose(1, 2).unsqueeze(2)        # [n,196,1,768]
        elif op.kind == B.OP_TOKENS:
            n = src.shape[0]
            pos = prog.blobs[op.weight].view(1, op.in_h, op.in_c)
            cls = prog.blobs[op.bias].view(1, 1, op.in_c).expand(n, 1, op.in_c)
            y = (torch.cat([cls, src.view(n, op.in_h - 1, op.in_c)], 1) + pos).unsqueeze(2)
        elif op.kind == B.OP_LAYERNORM:
            y = F.layer_norm(src.squeeze(2), (op.in_c,), prog.blobs[op.weight], prog.blobs[op.bias], eps=1e-6)
            y = y[:, 0] if op.reserved[0] else y.unsqueeze(2)
        elif op.kind == B.OP_ATTENTION:
            n, heads = src.shape[0], op.reserved[0]
            q, k, v = [t.view(n, op.in_h, heads, op.in_c // heads).transpose(1, 2) for t in src.squeeze(2).split(op.in_c, dim=2)]
            y = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(n, op.in_h, 1, op.in_c)
        elif op.kind == B.OP_MAXPOOL:
            y = F.max_pool2d(src.permute(0, 3, 1, 2), op.kh, op.stride, op.pad_lo).permute(0, 2, 3, 1)
        elif op.kind == B.OP_AVGPOOL:
            y = src.mean(dim=(1, 2))
        elif op.kind == B.OP_LINEAR:
            y = src @ prog.blobs[op.weight].float()[: op.out_c].t() + prog.blobs[op.bias][: op.out_c]
        else:
            raise AssertionError(op.kind)
        sub = 2 if (op.kind == B.OP_CONV_PAIR and op.reserved2 == 2) else 1
        assert tuple(y.shape[1:]) in ((op.out_h // sub, op.out_w // sub, op.out_c), (op.out_c,)), (op.kind, tuple(y.shape))
        if op.dst == B.SLOT_OUTPUT:
            out = y
        else:
            assert y[0].numel() <= prog.slot_elems[op.dst], "slot too small for what the op writes"
            slots[op.dst] = y
    return out


@pytest.mark.parametrize("arch,flops", [("resnet18", 3.6281e9), ("resnet50", 8.1784e9)])
def test_program_reproduces_the_torchvision_forward(arch, flops):
    import torchvision
    torch.manual_seed(0)
    model = getattr(torchvision.models, arch)(weights=None).eval()
    # non-trivial BN statistics so that the fold is really exercised
    g = torch.Generator().manual_seed(1)
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.weight.data.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.bias.data.copy_(torch.randn(m.num_features, generator=g) * 0.1)
    prog = P.build_resnet_program(model, arch)
    assert abs(prog.flops_per_image - flops) / flops < 1e-4
    x = torch.randn(2, 3, 224, 224, generator=g).to(torch.bfloat16).float()
    with torch.no_grad():
        ref = model(x)
        got = interpret(prog, x)
    # only difference: weights rounded to bf16
    assert (got - ref).abs().max().item() < 0.02 * ref.abs().max().item()
    assert torch.equal(got.argmax(1), ref.argmax(1))


def test_vit_program_reproduces_the_torchvision_forward():
    from mlmicroservicetemplate_b200.runtime import seeded_model
    model = seeded_model("vit_b_16", 0)
    prog = P.build_program(model, "vit_b_16")
    assert abs(prog.flops_per_image - 35.128e9) / 35.128e9 < 5e-3
    g = torch.Generator().manual_seed(4)
    x = torch.randn(2, 3, 224, 224, generator=g).to(torch.bfloat16).float()
    with torch.no_grad():
        ref = model(x)
        got = interpret(prog, x)
    assert ref.abs().max().item() > 0.5, "the re-initialised head must give non-trivial logits"
    assert (got - ref).abs().max().item() < 0.03 * ref.abs().max().item()


@pytest.mark.parametrize("arch", ["resnet50", "vit_b_16"])
def test_slots_are_never_clobbered_while_live(arch):
    from mlmicroservicetemplate_b200.runtime import seeded_model
    prog = P.build_program(seeded_model(arch, 0), arch)
    last_read = {}
    for i, op in enumerate(prog.ops):
        for s in (op.src, op.res, op.src2 if op.src2_c > 0 else B.SLOT_NONE):
            if s not in (B.SLOT_NONE,):
                last_read[s] = i
    def dsts(o):
        return [o.dst] + ([o.reserved[0]] if o.kind == B.OP_CONV_PAIR else [])
    writer = {}
    for i, op in enumerate(prog.ops):
        for d_ in dsts(op):
            assert d_ != op.src and (op.res == B.SLOT_NONE or d_ != op.res), "no in-place ops"
            writer[d_] = i
    # replay: a slot written at i must not be overwritten before its readers ran
    live_until = {}
    for i, op in enumerate(prog.ops):
        for d_ in dsts(op):
            if d_ in live_until:
                assert live_until[d_] < i, f"op {i} overwrites slot {d_} still needed by op {live_until[d_]}"
            nxt = [j for j, o in enumerate(prog.ops) if j > i and (o.src == d_ or o.res == d_ or (o.src2_c > 0 and o.src2 == d_))]
            rewrite = [j for j, o in enumerate(prog.ops) if j > i and d_ in dsts(o)]
            horizon = rewrite[0] if rewrite else len(prog.ops)
            readers = [j for j in nxt if j <= horizon]
            if d_ != B.SLOT_OUTPUT:
                assert readers, f"op {i} writes slot {d_} that nobody reads"
                live_until[d_] = max(readers)
    assert len(prog.slot_elems) <= 6


def test_stem_rewrite_equals_the_7x7_conv():
    g = torch.Generator().manual_seed(2)
    w = torch.randn(64, 3, 7, 7, generator=g)
    x = torch.randn(2, 3, 224, 224, generator=g)
    ref = F.conv2d(x, w, stride=2, padding=3)
    ws = P.stem_s2d_weight(w)                                       # [64, 4, 4, 16]
    s2d = torch.zeros(2, 16, 112, 112)
    for di in range(2):
        for dj in range(2):
            for c in range(3):
                s2d[:, (di * 2 + dj) * 3 + c] = x[:, c, di::2, dj::2]
    got = F.conv2d(F.pad(s2d, (2, 1, 2, 1)), ws.permute(0, 3, 1, 2))
    assert got.shape == ref.shape and (got - ref).abs().max().item() < 1e-4


def test_fold_bn_and_packing_layout():
    g = torch.Generator().manual_seed(3)
    conv = torch.nn.Conv2d(8, 16, 3, bias=False)
    bn = torch.nn.BatchNorm2d(16).eval()
    bn.running_mean.copy_(torch.randn(16, generator=g)); bn.running_var.copy_(torch.rand(16, generator=g) + 0.5)
    bn.weight.data.copy_(torch.rand(16, generator=g) + 0.5); bn.bias.data.copy_(torch.randn(16, generator=g))
    x = torch.randn(1, 8, 10, 10, generator=g)
    w, b = P.fold_bn(conv.weight, bn)
    with torch.no_grad():
        assert (F.conv2d(x, w, b) - bn(conv(x))).abs().max().item() < 1e-5
    packed = P.pack_conv_weight(w)
    assert packed.shape == (64, 3, 3, 8) and packed.dtype == torch.bfloat16
    assert torch.equal(packed[:16].float(), w.permute(0, 2, 3, 1).to(torch.bfloat16).float())
    assert packed[16:].abs().sum().item() == 0


def test_resnet50_transition_output_is_subsampled_once(monkeypatch):
    """The stage-1 -> stage-2 CONV_PAIR keeps only the even rows / columns of its first output (its other reader is the
    next block's stride-2 shortcut, which then reads the compact tensor with stride 1); B2P_PAIR_SUBSAMPLE=0 restores the
    full tensor and the stride-2 read.  Both programs compute the same network (the interpreter test above runs the
    default one)."""
    from mlmicroservicetemplate_b200.runtime import seeded_model
    model = seeded_model("resnet50", 0)
    prog = P.build_program(model, "resnet50")
    pairs = [op for op in prog.ops if op.kind == B.OP_CONV_PAIR]
    sub = [op for op in pairs if op.reserved2 == 2]
    assert len(pairs) == 5 and len(sub) == 1
    op = sub[0]
    assert (op.out_h, op.out_w, op.out_c, op.reserved[3]) == (56, 56, 256, 128)
    assert prog.slot_elems[op.dst] >= 28 * 28 * 256
    # the ops that see this tensor: from the pair op to the next writer of its slot (slot ids are recycled)
    i0 = list(prog.ops).index(op)
    live = []
    for o in prog.ops[i0 + 1:]:
        if o.dst == op.dst or (o.kind == B.OP_CONV_PAIR and o.reserved[0] == op.dst):
            break
        live.append(o)
    readers = [o for o in live if o.src2_c > 0 and o.src2 == op.dst]
    assert len(readers) == 1 and readers[0].src2_stride == 1 and readers[0].src2_c == 256 and readers[0].out_h == 28
    assert not any(o.src == op.dst or o.res == op.dst for o in live), "nobody reads the compact tensor at full resolution"
    monkeypatch.setenv("B2P_PAIR_SUBSAMPLE", "0")
    full = P.build_program(model, "resnet50")
    assert not any(o.kind == B.OP_CONV_PAIR and o.reserved2 == 2 for o in full.ops)
    strided = [o for o in full.ops if o.src2_c == 256 and o.out_h == 28]
    assert len(strided) == 1 and strided[0].src2_stride == 2
