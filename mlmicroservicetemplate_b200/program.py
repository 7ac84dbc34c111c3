"""Turns a torchvision classifier into the flat layer program + packed weight blobs that
libb200predict.so executes (include/b200predict.h, `b2p_op` / `b2p_model_desc`).

What init() (reference src/model/model.py:6-13) has to prepare per replica happens here on the host:
  * eval-mode BatchNorm is folded into the preceding conv in fp32
    (w' = w * gamma / sqrt(var + eps), b' = beta - mean * gamma / sqrt(var + eps);
    SP/torchvision/models/resnet.py:89-105, 143-163 for where the BNs sit),
  * conv weights are re-laid as [out_c][kh][kw][in_c] bf16 (K-major rows for tcgen05),
  * the 7x7/s2 stem is rewritten as a 4x4/s1 conv over a 2x2 space-to-depth input with 16-channel
    pixels so that TMA im2col can feed it,
  * activation slots are assigned by liveness so the workspace stays small.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import torch

from . import backend as B


@dataclass
class Program:
    ops: List[B.Op] = field(default_factory=list)
    slot_elems: List[int] = field(default_factory=lambda: [0])   # slot 0 = external input
    blobs: List[torch.Tensor] = field(default_factory=list)     # CPU tensors, kept alive for b2p_create
    in_elems: int = 0
    out_elems: int = 0
    flops_per_image: float = 0.0          # conv/linear 2*MACs only (SURVEY §8d accounting)
    name: str = ""

    # ---- slot allocator (greedy, by liveness)
    _free: List[int] = field(default_factory=list)

    def alloc(self, elems: int) -> int:
        best = None
        for s in self._free:
            if self.slot_elems[s] >= elems and (best is None or self.slot_elems[s] < self.slot_elems[best]):
                best = s
        if best is None and self._free:
            best = max(self._free, key=lambda s: self.slot_elems[s])
            self.slot_elems[best] = elems
        if best is not None:
            self._free.remove(best)
            return best
        self.slot_elems.append(elems)
        return len(self.slot_elems) - 1

    def release(self, slot: int) -> None:
        if slot > 0 and slot not in self._free:
            self._free.append(slot)

    def add_blob(self, t: torch.Tensor) -> int:
        self.blobs.append(t.contiguous())
        return len(self.blobs) - 1


def _bf16(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).contiguous()


def fold_bn(conv_w: torch.Tensor, bn: Optional[torch.nn.BatchNorm2d], conv_b: Optional[torch.Tensor] = None):
    """Returns (w', b') in fp32 with eval-mode BN folded into the conv."""
    w = conv_w.detach().float()
    b = conv_b.detach().float() if conv_b is not None else torch.zeros(w.shape[0])
    if bn is None:
        return w, b
    inv = bn.weight.detach().float() / torch.sqrt(bn.running_var.detach().float() + bn.eps)
    return w * inv.view(-1, 1, 1, 1), (b - bn.running_mean.detach().float()) * inv + bn.bias.detach().float()


def pack_conv_weight(w: torch.Tensor, pad_rows_to: int = 64) -> torch.Tensor:
    """[out_c, in_c, kh, kw] fp32 -> [out_c_pad, kh, kw, in_c] bf16."""
    oc = w.shape[0]
    ocp = (oc + pad_rows_to - 1) // pad_rows_to * pad_rows_to
    p = w.permute(0, 2, 3, 1).contiguous()
    if ocp != oc:
        p = torch.cat([p, torch.zeros(ocp - oc, *p.shape[1:])], 0)
    return _bf16(p)


def pad_bias(b: torch.Tensor, rows: int) -> torch.Tensor:
    out = torch.zeros(rows, dtype=torch.float32)
    out[: b.numel()] = b.float()
    return out


def stem_s2d_weight(w: torch.Tensor) -> torch.Tensor:
    """7x7/s2/p3 weight [oc,3,7,7] -> 4x4/s1 (pad 2 before, 1 after) weight [oc,4,4,16] over the 2x2
    space-to-depth input whose channel order is (di*2+dj)*3 + c.  Tap (a,b), sub-pixel (di,dj) maps
    to original tap (2a+di-1, 2b+dj-1); taps that fall outside the 7x7 window are zero."""
    oc = w.shape[0]
    out = torch.zeros(oc, 4, 4, 16, dtype=torch.float32)
    for a in range(4):
        for di in range(2):
            r = 2 * a + di - 1
            if not 0 <= r < 7:
                continue
            for b in range(4):
                for dj in range(2):
                    s = 2 * b + dj - 1
                    if not 0 <= s < 7:
                        continue
                    for c in range(3):
                        out[:, a, b, (di * 2 + dj) * 3 + c] = w[:, c, r, s]
    return out


def _mk_op(kind, src, dst, res=B.SLOT_NONE, in_hwc=(0, 0, 0), out_hwc=(0, 0, 0), k=(1, 1), stride=1, pad=(0, 0),
           relu=0, weight=-1, bias=-1, flag0=0, reserved=None, src2=B.SLOT_NONE, src2_c=0, src2_stride=1, reserved2=0) -> B.Op:
    r = reserved if reserved is not None else (flag0, 0, 0, 0)
    return B.Op(kind, src, dst, res, in_hwc[0], in_hwc[1], in_hwc[2], out_hwc[0], out_hwc[1], out_hwc[2],
                k[0], k[1], stride, pad[0], pad[1], relu, weight, bias, (B.C.c_int32 * 4)(*[int(v) for v in r]),
                src2, src2_c, src2_stride, int(reserved2))


class _Builder:
    def __init__(self, name: str):
        self.p = Program(name=name)

    def conv(self, src, hwc, w, b, stride, pad, relu, res=B.SLOT_NONE, dst=None, pad_hi=None, second=None):
        """w: fp32 [oc, ic, kh, kw] (BN already folded); returns (dst_slot, out_hwc).
        second = (slot, w2 [oc, c2, 1, 1], b2, stride2): a 1x1/stride2 conv of another slot accumulated into the
        same output (the projection shortcut), packed as extra K columns after the main weight."""
        h, wd, c = hwc
        oc, ic, kh, kw = w.shape
        assert ic == c, (ic, c)
        pad_hi = pad if pad_hi is None else pad_hi
        oh = (h + pad + pad_hi - kh) // stride + 1
        ow = (wd + pad + pad_hi - kw) // stride + 1
        packed = pack_conv_weight(w)
        src2, c2, stride2 = B.SLOT_NONE, 0, 1
        if second is not None:
            src2, w2, b2, stride2 = second
            c2 = w2.shape[1]
            assert w2.shape[0] == oc and tuple(w2.shape[2:]) == (1, 1) and c2 % 64 == 0
            ocp = packed.shape[0]
            w2p = torch.zeros(ocp, c2)
            w2p[:oc] = w2.float().view(oc, c2)
            packed = torch.cat([packed.view(ocp, -1), _bf16(w2p)], dim=1).contiguous()
            b = b + b2
            self.p.flops_per_image += 2.0 * oh * ow * oc * c2
        wi = self.p.add_blob(packed)
        bi = self.p.add_blob(pad_bias(b, packed.shape[0]))
        if dst is None:
            dst = self.p.alloc(oh * ow * oc)
        self.p.ops.append(_mk_op(B.OP_CONV, src, dst, res, (h, wd, c), (oh, ow, oc), (kh, kw), stride,
                                 (pad, pad_hi), int(relu), wi, bi, src2=src2, src2_c=c2, src2_stride=stride2))
        self.p.flops_per_image += 2.0 * oh * ow * oc * ic * kh * kw
        return dst, (oh, ow, oc)


def bottleneck_op(p: Program, src: int, dst: int, hw: int, cin: int, w1, b1, w2, b2, w3, b3, shortcut=None) -> None:
    """Appends one OP_BOTTLENECK (a whole 56x56 bottleneck block in one kernel).  w1 [64,cin,1,1], w2 [64,64,3,3],
    w3 [256,64,1,1] fp32 with BN folded; shortcut = None for an identity block (the residual is the block input).  The op
    format also describes a projection block (shortcut = (wd [256,cin,1,1], bd), folded into conv3 as extra K columns);
    the library's kernel covers identity blocks and b2p_create refuses the other form."""
    mid, cout = w1.shape[0], w3.shape[0]
    w3p, b3p = pack_conv_weight(w3).view(cout, -1), b3
    if shortcut is not None:
        wd, bd = shortcut
        w3p = torch.cat([w3p, _bf16(wd.float().view(cout, -1))], dim=1).contiguous()
        b3p = b3 + bd
        p.flops_per_image += 2.0 * hw * hw * cout * wd.shape[1]
    blobs = [p.add_blob(pack_conv_weight(w1)), p.add_blob(pad_bias(b1, mid)), p.add_blob(pack_conv_weight(w2)),
             p.add_blob(pad_bias(b2, mid)), p.add_blob(w3p), p.add_blob(pad_bias(b3p, cout))]
    op = _mk_op(B.OP_BOTTLENECK, src, dst, B.SLOT_NONE if shortcut is not None else src, (hw, hw, cin), (hw, hw, cout),
                (3, 3), 1, (1, 1), 1, blobs[0], blobs[1], reserved=(blobs[2], blobs[3], blobs[4], blobs[5]))
    op.reserved2 = mid
    p.ops.append(op)
    p.flops_per_image += 2.0 * hw * hw * (mid * cin + mid * mid * 9 + cout * mid)


def build_resnet_program(model: torch.nn.Module, name: str = "resnet", fused_stem: bool = True,
                         fuse_pairs: bool = True, fold_shortcut: bool = True, fuse_blocks: Optional[bool] = None,
                         subsample_pair_out: Optional[bool] = None) -> Program:
    """Walks a torchvision ResNet (SP/torchvision/models/resnet.py:266-282: conv1-bn1-relu-maxpool,
    layer1..4, avgpool, fc) and emits the program.  Works for BasicBlock and Bottleneck nets."""
    import os
    if subsample_pair_out is None:
        subsample_pair_out = os.getenv("B2P_PAIR_SUBSAMPLE", "1") != "0"
    from torchvision.models.resnet import BasicBlock, Bottleneck

    model = model.eval()
    bld = _Builder(name)
    p = bld.p
    p.in_elems = 3 * 224 * 224

    # stem: NCHW -> s2d NHWC16 -> 4x4 conv (== 7x7/s2/p3) + folded bn1 + relu
    s2d = p.alloc(112 * 112 * 16)
    p.ops.append(_mk_op(B.OP_STEM_PACK, B.SLOT_INPUT, s2d, in_hwc=(224, 224, 3), out_hwc=(112, 112, 16)))
    w, b = fold_bn(model.conv1.weight, model.bn1)
    ws = stem_s2d_weight(w).permute(0, 3, 1, 2).contiguous()          # back to [oc, ic=16, 4, 4]
    mpool = model.maxpool
    can_fuse = (fused_stem and tuple(model.conv1.kernel_size) == (7, 7) and model.conv1.out_channels == 64
                and mpool.kernel_size == 3 and mpool.stride == 2 and mpool.padding == 1)
    if can_fuse:
        # dedicated stem kernel: conv + bias + ReLU + 3x3/s2 max pool in one pass, [n,56,56,64] out
        wi = p.add_blob(pack_conv_weight(ws))
        bi = p.add_blob(pad_bias(b, 64))
        mp = p.alloc(56 * 56 * 64)
        p.ops.append(_mk_op(B.OP_STEM_CONV, s2d, mp, in_hwc=(112, 112, 16), out_hwc=(56, 56, 64), k=(4, 4), stride=1,
                            pad=(2, 1), relu=1, weight=wi, bias=bi, flag0=1))
        p.flops_per_image += 2.0 * 112 * 112 * 64 * 147
        p.release(s2d)
    else:
        c1, hwc = bld.conv(s2d, (112, 112, 16), ws, b, stride=1, pad=2, pad_hi=1, relu=True)
        # the stem's FLOPs are accounted on the original 7x7x3 window, not the zero-padded s2d one
        p.flops_per_image += 2.0 * 112 * 112 * 64 * (147 - 256)
        p.release(s2d)
        # maxpool 3x3/s2/p1
        mp = p.alloc(56 * 56 * 64)
        p.ops.append(_mk_op(B.OP_MAXPOOL, c1, mp, in_hwc=hwc, out_hwc=(56, 56, 64), k=(3, 3), stride=2, pad=(1, 1)))
        p.release(c1)
    x, hwc = mp, (56, 56, 64)

    import os
    if fuse_blocks is None:
        fuse_blocks = os.getenv("B2P_FUSE_BLOCKS", "0") != "0"
    def fusable(blk, in_hwc):
        """An identity bottleneck of the 56x56 stage (256 -> 64 -> 64 -> 256): the shape bottleneck56.cu covers."""
        return (fuse_blocks and isinstance(blk, Bottleneck) and tuple(in_hwc) == (56, 56, 256) and blk.downsample is None
                and blk.conv1.out_channels == 64 and blk.conv3.out_channels == 256 and blk.conv2.stride == (1, 1)
                and blk.conv2.groups == 1 and blk.conv2.dilation == (1, 1))

    blocks = [blk for layer in (model.layer1, model.layer2, model.layer3, model.layer4) for blk in layer]
    carry = None          # (t1 slot, hwc1) when the previous block's CONV_PAIR already computed this block's conv1
    x_sub = 1             # 2: x holds only the even rows / columns of the block input (see the CONV_PAIR branch)
    for bi_, blk in enumerate(blocks):
        if True:
            # A 56x56 bottleneck (ResNet-50/101/152 layer1) runs as ONE kernel: only the block input is read and the
            # block output written (measured against the layer-by-layer form with its conv3 -> next-conv1 pairing).
            if carry is None and fusable(blk, hwc):
                w1, b1 = fold_bn(blk.conv1.weight, blk.bn1)
                w2, b2 = fold_bn(blk.conv2.weight, blk.bn2)
                w3, b3 = fold_bn(blk.conv3.weight, blk.bn3)
                out = p.alloc(56 * 56 * 256)
                bottleneck_op(p, x, out, 56, hwc[2], w1, b1, w2, b2, w3, b3)
                p.release(x)
                x, hwc = out, (56, 56, 256)
                continue
            if isinstance(blk, Bottleneck):
                w1, b1 = fold_bn(blk.conv1.weight, blk.bn1)
                w2, b2 = fold_bn(blk.conv2.weight, blk.bn2)
                w3, b3 = fold_bn(blk.conv3.weight, blk.bn3)
                # The projection shortcut (1x1/stride-s conv of x) is not a layer of its own: it is accumulated into
                # conv3's accumulator as extra K columns read from x (out = relu(W3 t2 + Wd x + b3 + bd)), so the
                # "identity" tensor of a downsampling block is never written to or read from HBM.
                shortcut = None
                if blk.downsample is not None:
                    wd, bd = fold_bn(blk.downsample[0].weight, blk.downsample[1])
                    ds_stride = blk.downsample[0].stride[0] // x_sub      # a pre-subsampled x is read with stride 1
                    ds_presub = x_sub != 1
                    if fold_shortcut and wd.shape[1] % 64 == 0 and tuple(blk.downsample[0].kernel_size) == (1, 1):
                        shortcut = (x, wd, bd, ds_stride)
                        idn = B.SLOT_NONE
                    else:
                        idn, _ = bld.conv(x, hwc, wd, bd, ds_stride, 0, False)
                else:
                    assert x_sub == 1
                    idn = x
                x_sub = 1
                if carry is not None:
                    t1, hwc1 = carry
                    carry = None
                else:
                    t1, hwc1 = bld.conv(x, hwc, w1, b1, 1, 0, True)
                if blk.downsample is not None and shortcut is None:
                    p.release(x)
                t2, hwc2 = bld.conv(t1, hwc1, w2, b2, blk.conv2.stride[0], 1, True)
                p.release(t1)
                # conv3 (+ residual + ReLU): fused with the NEXT block's conv1 when that is a plain 1x1/s1 conv
                nxt = blocks[bi_ + 1] if bi_ + 1 < len(blocks) else None
                n1 = w3.shape[0]
                can_pair = (fuse_pairs and isinstance(nxt, Bottleneck) and nxt.conv1.kernel_size == (1, 1)
                            and not fusable(nxt, (hwc2[0], hwc2[1], w3.shape[0]))     # a fused next block has its own conv1
                            # a STRIDED folded shortcut stays out of the pair kernel: its gathered (stride-2 im2col) operand
                            # loads starve the pair's two-stage ring - measured 170 us paired vs 73 + 40 us as conv3(+shortcut)
                            # and a separate conv1 for ResNet-50's first 28x28 block (profiles/r02v_pair_shortcut_ab.txt)
                            and not (shortcut is not None and (shortcut[3] != 1 or
                                                               (ds_presub and os.getenv("B2P_PAIR_AFTER_SUB", "0") == "0")))
                            and nxt.conv1.stride == (1, 1) and nxt.conv1.out_channels in (64, 128)
                            and n1 % 128 == 0 and w3.shape[1] in (64, 128) and tuple(blk.conv3.kernel_size) == (1, 1))
                # (wider pairs - K1 >= 256 or N2 = 256, i.e. ResNet-50's stage 3 - were measured slower than the two
                #  separate kernels: 90 vs 77 us, the fused kernel's two-stage operand ring is latency bound there)
                if can_pair:
                    wn, bn_ = fold_bn(nxt.conv1.weight, nxt.bn1)
                    n2 = wn.shape[0]
                    oh, ow = hwc2[0], hwc2[1]
                    w3p, b3p, sc_kw = pack_conv_weight(w3).view(n1, -1), b3, {}
                    if shortcut is not None:      # [W3 | Wd] along K, biases summed
                        sx, swd, sbd, sstride = shortcut
                        w3p = torch.cat([w3p, _bf16(swd.float().view(n1, -1))], dim=1).contiguous()
                        b3p = b3 + sbd
                        sc_kw = dict(src2=sx, src2_c=swd.shape[1], src2_stride=sstride)
                        p.flops_per_image += 2.0 * oh * ow * n1 * swd.shape[1]
                    wi = p.add_blob(w3p)
                    bi = p.add_blob(pad_bias(b3p, n1))
                    wi2 = p.add_blob(pack_conv_weight(wn))
                    bi2 = p.add_blob(pad_bias(bn_, n2))
                    # When the next block downsamples, its conv1 (this kernel's second GEMM) is the only full-resolution
                    # reader of this block's output; the other one is the 1x1/stride-2 projection shortcut.  So only the
                    # even rows / columns are written (a quarter of the tensor: 103 instead of 411 MB per 256 images at
                    # 56x56x256) and the shortcut reads the compact tensor with stride 1.
                    sub = 1
                    if (subsample_pair_out and nxt.downsample is not None and tuple(nxt.downsample[0].kernel_size) == (1, 1)
                            and tuple(nxt.downsample[0].stride) == (2, 2) and nxt.conv2.stride == (2, 2)
                            and oh % 2 == 0 and ow % 2 == 0):
                        sub = 2
                    out = p.alloc((oh // sub) * (ow // sub) * n1)
                    t1n = p.alloc(oh * ow * n2)
                    p.ops.append(_mk_op(B.OP_CONV_PAIR, t2, out, idn, hwc2, (oh, ow, n1), (1, 1), 1, (0, 0), 1, wi, bi,
                                        reserved=(t1n, wi2, bi2, n2), reserved2=sub, **sc_kw))
                    x_sub = sub
                    if shortcut is not None:
                        p.release(x)
                    p.flops_per_image += 2.0 * oh * ow * n1 * w3.shape[1] + 2.0 * oh * ow * n2 * n1
                    hwc3 = (oh // sub, ow // sub, n1)
                    carry = (t1n, (oh, ow, n2))
                elif shortcut is not None:
                    out, hwc3 = bld.conv(t2, hwc2, w3, b3, 1, 0, True, second=shortcut)
                    p.release(x)
                else:
                    out, hwc3 = bld.conv(t2, hwc2, w3, b3, 1, 0, True, res=idn)
                p.release(t2)
                if idn != B.SLOT_NONE:
                    p.release(idn)
                x, hwc = out, hwc3
            elif isinstance(blk, BasicBlock):
                w1, b1 = fold_bn(blk.conv1.weight, blk.bn1)
                w2, b2 = fold_bn(blk.conv2.weight, blk.bn2)
                t1, hwc1 = bld.conv(x, hwc, w1, b1, blk.conv1.stride[0], 1, True)
                if blk.downsample is not None:
                    wd, bd = fold_bn(blk.downsample[0].weight, blk.downsample[1])
                    idn, _ = bld.conv(x, hwc, wd, bd, blk.downsample[0].stride[0], 0, False)
                    p.release(x)
                else:
                    idn = x
                out, hwc2 = bld.conv(t1, hwc1, w2, b2, 1, 1, True, res=idn)
                p.release(t1)
                p.release(idn)
                x, hwc = out, hwc2
            else:
                raise TypeError(f"unsupported block {type(blk)}")

    feat = p.alloc(hwc[2])
    p.ops.append(_mk_op(B.OP_AVGPOOL, x, feat, in_hwc=hwc, out_hwc=(1, 1, hwc[2])))
    p.release(x)
    fc_w = model.fc.weight.detach().float()                            # [classes, in]
    classes, fin = fc_w.shape
    rows = (classes + 127) // 128 * 128
    wp = torch.zeros(rows, fin)
    wp[:classes] = fc_w
    wi = p.add_blob(_bf16(wp))
    bi = p.add_blob(pad_bias(model.fc.bias.detach(), rows))
    p.ops.append(_mk_op(B.OP_LINEAR, feat, B.SLOT_OUTPUT, in_hwc=(1, 1, fin), out_hwc=(1, 1, classes),
                        weight=wi, bias=bi))
    p.flops_per_image += 2.0 * classes * fin
    p.out_elems = classes
    return p


def build_vit_program(model: torch.nn.Module, name: str = "vit_b_16") -> Program:
    """Walks a torchvision VisionTransformer (SP/torchvision/models/vision_transformer.py:160-305:
    conv_proj patchify, class token, positional embedding, encoder blocks :110-119, final LN, head).
    Token matrices are "images" of 197 x 1 so every linear layer is a 1x1 OP_CONV on the GEMM core."""
    model = model.eval()
    if model.patch_size != 16 or model.image_size != 224 or model.hidden_dim != 768:
        raise ValueError("only ViT-B/16 at 224 is supported")
    bld = _Builder(name)
    p = bld.p
    p.in_elems = 3 * 224 * 224
    seq, dim, heads = model.seq_length, model.hidden_dim, model.encoder.layers[0].num_heads
    tok = (seq, 1, dim)

    def linear(src, hwc, lin_w, lin_b, act=B.ACT_NONE, res=B.SLOT_NONE):
        w = lin_w.detach().float()
        return bld.conv(src, hwc, w.view(w.shape[0], w.shape[1], 1, 1), lin_b.detach().float(), 1, 0, act, res=res)

    def layernorm(src, ln, cls_only=False):
        wi = p.add_blob(ln.weight.detach().float())
        bi = p.add_blob(ln.bias.detach().float())
        dst = p.alloc(dim if cls_only else seq * dim)
        p.ops.append(_mk_op(B.OP_LAYERNORM, src, dst, in_hwc=tok, out_hwc=(1, 1, dim) if cls_only else tok, weight=wi,
                            bias=bi, flag0=int(cls_only)))
        return dst

    # patch embedding: patchify -> GEMM [n*196, 768] x W^T + b -> tokens (+cls, +pos)
    patches = p.alloc((seq - 1) * dim)
    p.ops.append(_mk_op(B.OP_PATCHIFY, B.SLOT_INPUT, patches, in_hwc=(224, 224, 3), out_hwc=(seq - 1, 1, dim)))
    cw = model.conv_proj.weight.detach().float().reshape(dim, 3 * 16 * 16)
    proj, _ = linear(patches, (seq - 1, 1, dim), cw, model.conv_proj.bias)
    p.release(patches)
    pos = p.add_blob(model.encoder.pos_embedding.detach().float().reshape(seq * dim))
    cls = p.add_blob(model.class_token.detach().float().reshape(dim))
    x = p.alloc(seq * dim)
    p.ops.append(_mk_op(B.OP_TOKENS, proj, x, in_hwc=tok, out_hwc=tok, weight=pos, bias=cls))
    p.release(proj)

    for blk in model.encoder.layers:
        att = blk.self_attention
        y = layernorm(x, blk.ln_1)
        qkv, _ = linear(y, tok, att.in_proj_weight, att.in_proj_bias)
        p.release(y)
        a = p.alloc(seq * dim)
        p.ops.append(_mk_op(B.OP_ATTENTION, qkv, a, in_hwc=tok, out_hwc=tok, flag0=heads))
        p.flops_per_image += 2.0 * 2.0 * seq * seq * (dim // heads) * heads
        p.release(qkv)
        x2, _ = linear(a, tok, att.out_proj.weight, att.out_proj.bias, res=x)
        p.release(a)
        p.release(x)
        y = layernorm(x2, blk.ln_2)
        h, hwc_h = linear(y, tok, blk.mlp[0].weight, blk.mlp[0].bias, act=B.ACT_GELU)
        p.release(y)
        x, _ = linear(h, hwc_h, blk.mlp[3].weight, blk.mlp[3].bias, res=x2)
        p.release(h)
        p.release(x2)

    feat = layernorm(x, model.encoder.ln, cls_only=True)
    p.release(x)
    head = model.heads.head
    classes, fin = head.weight.shape
    rows = (classes + 127) // 128 * 128
    wp = torch.zeros(rows, fin)
    wp[:classes] = head.weight.detach().float()
    wi = p.add_blob(_bf16(wp))
    bi = p.add_blob(pad_bias(head.bias.detach(), rows))
    p.ops.append(_mk_op(B.OP_LINEAR, feat, B.SLOT_OUTPUT, in_hwc=(1, 1, fin), out_hwc=(1, 1, classes), weight=wi, bias=bi))
    p.flops_per_image += 2.0 * classes * fin
    p.out_elems = classes
    return p


def build_program(model: torch.nn.Module, name: str = "") -> Program:
    """Dispatches on the torchvision module type."""
    cls = type(model).__name__
    if cls == "ResNet":
        return build_resnet_program(model, name or "resnet")
    if cls == "VisionTransformer":
        return build_vit_program(model, name or "vit")
    raise TypeError(f"no layer-program builder for {cls}")


def single_conv_program(w: torch.Tensor, b: torch.Tensor, hwc, stride: int, pad: int, relu: bool,
                        with_residual: bool = False, pad_hi: Optional[int] = None) -> Program:
    """One CONV op from the external NHWC input to the external bf16 NHWC output; with_residual adds
    the input itself as the residual (needs matching shapes).  Used by the layer-parity tests."""
    bld = _Builder("single_conv")
    p = bld.p
    h, wd, c = hwc
    p.in_elems = h * wd * c
    _, ohwc = bld.conv(B.SLOT_INPUT, hwc, w.float(), b.float(), stride, pad, relu,
                       res=B.SLOT_INPUT if with_residual else B.SLOT_NONE, dst=B.SLOT_OUTPUT, pad_hi=pad_hi)
    p.out_elems = ohwc[0] * ohwc[1] * ohwc[2]
    return p


def single_stem_program(w4: torch.Tensor, b: torch.Tensor, pool: bool) -> Program:
    """One STEM_CONV op (external s2d NHWC16 input -> external output); w4 is [64, 16, 4, 4] fp32."""
    p = Program(name="single_stem")
    p.in_elems = 112 * 112 * 16
    wi = p.add_blob(pack_conv_weight(w4.float()))
    bi = p.add_blob(pad_bias(b, 64))
    side = 56 if pool else 112
    p.out_elems = side * side * 64
    p.ops.append(_mk_op(B.OP_STEM_CONV, B.SLOT_INPUT, B.SLOT_OUTPUT, in_hwc=(112, 112, 16), out_hwc=(side, side, 64),
                        k=(4, 4), stride=1, pad=(2, 1), relu=1, weight=wi, bias=bi, flag0=int(pool)))
    return p


def single_linear_program(w: torch.Tensor, b: torch.Tensor) -> Program:
    p = Program(name="single_linear")
    classes, fin = w.shape
    rows = (classes + 127) // 128 * 128
    wp = torch.zeros(rows, fin)
    wp[:classes] = w.float()
    wi = p.add_blob(_bf16(wp))
    bi = p.add_blob(pad_bias(b, rows))
    p.in_elems = fin
    p.out_elems = classes
    p.ops.append(_mk_op(B.OP_LINEAR, B.SLOT_INPUT, B.SLOT_OUTPUT, in_hwc=(1, 1, fin), out_hwc=(1, 1, classes),
                        weight=wi, bias=bi))
    return p
