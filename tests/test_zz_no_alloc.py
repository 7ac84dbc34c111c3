"""b2p_head_static — the classifier head for callers that cannot allocate device memory (a process
whose own pool already owns the whole GPU, or one that keeps cudaMalloc off its request path because
it stalls the device).  The function keeps its intermediate (the pooled features) in a fixed 1 MiB
array inside the library instead of a cudaMalloc'd workspace slot.

Checked here through the C ABI against the regular entry point (a two-op AVGPOOL -> LINEAR program
run by b2p_forward, whose intermediate lives in workspace allocated by b2p_create), on two calls in a
row that share the array — the second smaller than the first, so stale rows from the first call are
still sitting in it.  This file sorts last on purpose: a run that stops at the first failure has
already been through every other test."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-3          # same bound tests/test_gpu_layers.py::test_linear uses for fp32 logits


def _head_program(w, b, hw, c):
    from mlmicroservicetemplate_b200 import backend as B, program as P
    p = P.Program(name="head")
    classes = w.shape[0]
    rows = (classes + 127) // 128 * 128
    wp = torch.zeros(rows, c)
    wp[:classes] = w
    wi = p.add_blob(wp.to(torch.bfloat16))
    bi = p.add_blob(P.pad_bias(b, rows))
    pooled = p.alloc(c)
    side = int(hw ** 0.5)
    p.in_elems = hw * c
    p.out_elems = classes
    p.ops.append(P._mk_op(B.OP_AVGPOOL, B.SLOT_INPUT, pooled, in_hwc=(side, side, c), out_hwc=(1, 1, c)))
    p.ops.append(P._mk_op(B.OP_LINEAR, pooled, B.SLOT_OUTPUT, in_hwc=(1, 1, c), out_hwc=(1, 1, classes), weight=wi, bias=bi))
    return p, p.blobs[wi], p.blobs[bi], rows


def test_head_static_matches_the_allocating_path_on_two_calls_in_a_row(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    g = torch.Generator().manual_seed(21)
    hw, c, classes = 49, 2048, 1000
    w = torch.randn(classes, c, generator=g) / c ** 0.5
    b = torch.randn(classes, generator=g)
    prog, w_packed, b_padded, rows = _head_program(w, b, hw, c)
    rep = B.Replica(0, prog, max_batch=B.HEAD_STATIC_MAX_IMAGES)
    w_dev, b_dev = w_packed.cuda(), b_padded.cuda()          # caller-owned device copies for the static entry point

    for n in (200, 37):                                       # second call is smaller: rows 37..199 hold stale data
        feat = torch.randn(n, 7, 7, c, generator=g).to(torch.bfloat16).cuda()
        want = torch.full((n, classes), float("nan"), device="cuda")
        got = torch.full((n, classes), float("nan"), device="cuda")
        rep.forward(feat, n, want)
        B.head_static(feat, n, hw, c, w_dev, rows, b_dev, classes, got)
        torch.cuda.synchronize()
        cpu = feat.float().cpu().mean(dim=(1, 2)).to(torch.bfloat16).float() @ w_packed[:classes].float().t() + b
        assert not torch.isnan(got).any()
        assert (got - want).abs().max().item() <= TOL, "static-scratch head differs from the workspace-backed one"
        assert (got.cpu() - cpu).abs().max().item() <= 2e-2
    rep.close()


def test_head_static_rejects_sizes_beyond_its_fixed_scratch(cuda_device):
    from mlmicroservicetemplate_b200 import backend as B
    x = torch.zeros(8, device="cuda")
    with pytest.raises(B.B2PError) as e:
        B.head_static(x, B.HEAD_STATIC_MAX_IMAGES + 1, 49, 2048, x, 1024, x, 1000, x)
    assert e.value.code == -5 and "256 images" in e.value.detail
    with pytest.raises(B.B2PError) as e:
        B.head_static(x, 4, 49, B.HEAD_STATIC_MAX_CHANNELS + 64, x, 1024, x, 1000, x)
    assert e.value.code == -5
