"""Times one 1x1 convolution layer shape through the C ABI, with and without a residual stream.

    python benchmarks/layer_time.py --n 256 --hw 14 --cin 256 --cout 1024 [--iters 200]

Program A: x -> conv -> slot (no residual).  Program B: A followed by the same conv with A's output as residual.
time(B) - time(A) is the residual variant's time.  Environment switches of the library (B2P_CTA2, B2P_TAIL_SPLIT,
B2P_LIB ...) apply, so this is the A/B tool for single-kernel changes.  Prints one JSON line.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--hw", type=int, default=14)
    ap.add_argument("--cin", type=int, default=256)
    ap.add_argument("--cout", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    import torch
    from mlmicroservicetemplate_b200 import backend as B, program as P
    torch.cuda.set_device(0)
    g = torch.Generator().manual_seed(0)
    w = torch.randn(args.cout, args.cin, 1, 1, generator=g) / args.cin ** 0.5
    b = torch.randn(args.cout, generator=g) * 0.1
    hwc = (args.hw, args.hw, args.cin)

    def build(with_res):
        bld = P._Builder("layer_time")
        p = bld.p
        p.in_elems = args.hw * args.hw * args.cin
        s1, ohwc = bld.conv(B.SLOT_INPUT, hwc, w, b, 1, 0, True)
        if with_res:
            bld.conv(B.SLOT_INPUT, hwc, w, b, 1, 0, True, res=s1, dst=B.SLOT_OUTPUT)
        else:
            bld.conv(B.SLOT_INPUT, hwc, w, b, 1, 0, True, dst=B.SLOT_OUTPUT)
        p.out_elems = ohwc[0] * ohwc[1] * ohwc[2]
        return p

    x = (torch.rand(args.n, args.hw, args.hw, args.cin, generator=g) - 0.5).to(torch.bfloat16).cuda()
    out = torch.empty(args.n, args.hw, args.hw, args.cout, dtype=torch.bfloat16, device="cuda")
    res = {}
    for name, with_res in (("two_plain", False), ("plain_then_residual", True)):
        rep = B.Replica(0, build(with_res), max_batch=args.n)
        for _ in range(10):
            rep.forward(x, args.n, out)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.iters):
            rep.forward(x, args.n, out)
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / args.iters * 1e3
        rep.close()
    plain = res["two_plain"] / 2
    with_res_us = res["plain_then_residual"] - plain
    flop = 2.0 * args.n * args.hw * args.hw * args.cin * args.cout
    print(json.dumps({"shape": vars(args), "plain_us": plain, "residual_us": with_res_us,
                      "plain_tflops": flop / plain / 1e6, "residual_tflops": flop / with_res_us / 1e6}))


if __name__ == "__main__":
    main()
