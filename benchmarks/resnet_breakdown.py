#!/usr/bin/env python
"""Per-kernel time breakdown of one eager ResNet-50 training step of the native engine
(torch.profiler CUPTI records). Diagnostic only - never a bench value."""
import argparse
import json
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.models.resnet import ResNet  # noqa: E402
from tf_operator_b200.parallel.engine import DataParallelEngine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--out", default="")
    ap.add_argument("--stem-bwd", default="", help="gather | scatter (ops.bn.STEM_BWD_MODE)")
    args = ap.parse_args()
    if args.stem_bwd:
        from tf_operator_b200.ops import bn as _bn
        _bn.STEM_BWD_MODE = args.stem_bwd
    torch.manual_seed(0)
    B = args.batch_size
    eng = DataParallelEngine(ResNet(50), batch_size=B, image_size=224, lr=0.1, graph=False,
                             stem_s2d=True)
    x = torch.randint(0, 256, (B, 224, 224, 3), dtype=torch.uint8).pin_memory()
    y = torch.randint(0, 1000, (B,), dtype=torch.int64).pin_memory()
    for _ in range(4):
        eng.step(x, y).result()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        eng.step(x, y).result()
        torch.cuda.synchronize()
    agg = defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            a = agg[ev.name[:120]]
            a[0] += ev.device_time
            a[1] += 1
    rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
    total = sum(v[0] for _, v in rows)
    groups = defaultdict(float)
    for k, v in rows:
        lk = k.lower()
        g = ("gemm / conv-stem (ours)" if "conv_stem" in lk
             else "conv fwd+dgrad tcgen05 (ours)" if "conv_igemm" in lk
             else "conv wgrad tcgen05 (ours)" if "conv_wgrad" in lk
             else "conv (cuDNN/cutlass)" if any(t in lk for t in ("implicit_gemm", "cudnn", "xmma", "conv", "nchwtonhwc", "nhwctonchw"))
             else "bn (ours)" if "bn_" in lk else "pool (ours)" if "pool" in lk
             else "optimizer/pack (ours)" if any(t in lk for t in ("sgd", "pack", "adamw")) else
             "gemm / conv-stem (ours)" if ("gemm_tn" in lk or "conv_stem" in lk) else "preproc (ours)" if ("stem" in lk or "u8" in lk)
             else "memcpy" if "memcpy" in lk else "other torch")
        groups[g] += v[0]
    print(f"total kernel time {total / 1e3:.3f} ms")
    for g, t in sorted(groups.items(), key=lambda kv: -kv[1]):
        print(f"  {t / total * 100:5.1f}%  {t / 1e3:8.3f} ms  {g}")
    for k, v in rows[:args.top]:
        print(f"  {v[0] / 1e3:9.3f} ms  x{v[1]:<4d} {k}")
    for k, v in rows:
        if "stem_bn" in k or "maxpool" in k:
            print(f"  [stem tail] {v[0] / 1e3:9.3f} ms  x{v[1]:<4d} {k[:90]}")
    if args.out:
        json.dump({"total_kernel_ms": total / 1e3, "groups_ms": {g: t / 1e3 for g, t in groups.items()},
                   "kernels": [{"name": k, "ms": v[0] / 1e3, "calls": v[1]} for k, v in rows[:args.top]]},
                  open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
