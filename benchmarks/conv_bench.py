#!/usr/bin/env python
"""Per-layer timing of the tcgen05 convolution kernels (csrc/conv_tc.cu) on the ResNet-50 layer
shapes at the bench batch size, next to cuDNN on the same tensors. CUDA events on the launching
stream, a 1 GiB write between timed iterations so nothing is served from L2, median of --iters.

    python benchmarks/conv_bench.py --batch 256 --out gpurun_out/conv_bench.json
"""
import argparse
import json
import os
import statistics
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops import conv as C  # noqa: E402

CL = torch.channels_last
# (H, Cin, Cout, k, stride, count in ResNet-50)
LAYERS = [
    (56, 64, 64, 1, 1, 1), (56, 64, 64, 3, 1, 3), (56, 64, 256, 1, 1, 4), (56, 256, 64, 1, 1, 2),
    (56, 256, 128, 1, 1, 1), (56, 128, 128, 3, 2, 1), (28, 128, 512, 1, 1, 4), (56, 256, 512, 1, 2, 1),
    (28, 512, 128, 1, 1, 3), (28, 128, 128, 3, 1, 3),
    (28, 512, 256, 1, 1, 1), (28, 256, 256, 3, 2, 1), (14, 256, 1024, 1, 1, 6), (28, 512, 1024, 1, 2, 1),
    (14, 1024, 256, 1, 1, 5), (14, 256, 256, 3, 1, 5),
    (14, 1024, 512, 1, 1, 1), (14, 512, 512, 3, 2, 1), (7, 512, 2048, 1, 1, 3), (14, 1024, 2048, 1, 2, 1),
    (7, 2048, 512, 1, 1, 2), (7, 512, 512, 3, 1, 2),
]


def timeit(fn, flush, iters):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(iters):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-cudnn", action="store_true")
    ap.add_argument("--only", default="", help="fwd,dgrad,wgrad subset")
    args = ap.parse_args()
    only = set(args.only.split(",")) if args.only else {"fwd", "dgrad", "wgrad"}
    torch.backends.cudnn.benchmark = True
    n = args.batch
    flush = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    rows = []
    tot = {"fwd": [0.0, 0.0], "dgrad": [0.0, 0.0], "wgrad": [0.0, 0.0]}
    for h, cin, cout, k, s, cnt in LAYERS:
        x = torch.randn((n, cin, h, h), device="cuda").bfloat16().contiguous(memory_format=CL)
        w = (torch.randn((cout, cin, k, k), device="cuda") / (cin * k * k) ** 0.5).bfloat16()
        w = w.contiguous(memory_format=CL)
        oh, ow = C._out_hw(h, h, k, s)
        dy = torch.randn((n, cout, oh, ow), device="cuda").bfloat16().contiguous(memory_format=CL)
        stats = torch.zeros(2, cout, dtype=torch.float32, device="cuda")
        flops = 2.0 * n * oh * ow * cout * cin * k * k
        row = {"layer": f"{h}x{h} {cin}->{cout} k{k} s{s}", "count": cnt, "gflop": flops / 1e9}
        ours = {"fwd": lambda: C.conv_fwd(x, w, s, stats),
                "dgrad": lambda: C.conv_dgrad(dy, w, x.shape, s),
                "wgrad": lambda: C.conv_wgrad(x, dy, w, s)}
        lib = {"fwd": lambda: F.conv2d(x, w, stride=s, padding=k // 2),
               "dgrad": lambda: torch.ops.aten.convolution_backward(
                   dy, x, w, None, [s, s], [k // 2] * 2, [1, 1], False, [0, 0], 1, [True, False, False]),
               "wgrad": lambda: torch.ops.aten.convolution_backward(
                   dy, x, w, None, [s, s], [k // 2] * 2, [1, 1], False, [0, 0], 1, [False, True, False])}
        for what in ("fwd", "dgrad", "wgrad"):
            if what not in only:
                continue
            t = timeit(ours[what], flush, args.iters)
            row[what + "_us"] = round(t, 1)
            row[what + "_tflops"] = round(flops / t / 1e6, 1)
            tot[what][0] += t * cnt
            if not args.no_cudnn:
                tl = timeit(lib[what], flush, args.iters)
                row[what + "_cudnn_us"] = round(tl, 1)
                tot[what][1] += tl * cnt
        rows.append(row)
        print(json.dumps(row), flush=True)
    summary = {k: {"ours_ms": round(v[0] / 1e3, 3), "cudnn_ms": round(v[1] / 1e3, 3)} for k, v in tot.items()}
    print(json.dumps({"per_step_weighted": summary}))
    if args.out:
        json.dump({"batch": n, "layers": rows, "per_step_weighted": summary}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
