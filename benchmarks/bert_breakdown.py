#!/usr/bin/env python
"""Per-kernel time breakdown of one eager BERT training step (native tcgen05 path vs the
all-library twin), from torch.profiler's CUPTI kernel records. Diagnostic only: numbers taken
under a profiler are never reported as bench values."""
import argparse
import json
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tf_operator_b200.models.bert import BertConfig, BertForPreTraining  # noqa: E402
from tf_operator_b200.parallel.engine import FlatDataParallel  # noqa: E402


def run(impl, model_name, B, S, top):
    dev = torch.device("cuda", 0)
    cfg = getattr(BertConfig, model_name)()
    cfg.tc_gemm = impl == "native"
    torch.manual_seed(1234)
    model = BertForPreTraining(cfg).to(dev).to(torch.bfloat16)
    ids = torch.randint(0, cfg.vocab_size, (B, S), device=dev)
    labels = torch.where(torch.rand(B, S, device=dev) < 0.15, ids, torch.full_like(ids, -100))
    nsp = torch.randint(0, 2, (B,), device=dev)
    if impl == "native":
        dp = FlatDataParallel(model, comm=None, optimizer="adamw", lr=1e-4, weight_decay=0.01,
                              bucket_bytes=64 << 20)
        step = lambda: dp.train_step(lambda: model(ids, labels, nsp))  # noqa: E731
    else:
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=0.01, fused=True)

        def step():
            opt.zero_grad(set_to_none=True)
            loss = model(ids, labels, nsp)
            loss.backward()
            opt.step()
            return loss.detach()
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    # record the (m, n, k) of every gemm_tn call in issue order to pair with the kernel records
    shapes = []
    if impl == "native":
        from tf_operator_b200.ops import gemm as G
        real, real_g = G.gemm_tn, G.gemm

        def spy(a, b, *args, **kw):
            shapes.append((a.shape[0], b.shape[0], a.shape[1]))
            return real(a, b, *args, **kw)

        def spy_g(a, b, *args, a_t=False, b_t=False, **kw):
            m, k = (a.shape[1], a.shape[0]) if a_t else a.shape
            shapes.append((m, b.shape[1] if b_t else b.shape[0], k))
            return real_g(a, b, *args, a_t=a_t, b_t=b_t, **kw)
        G.gemm_tn, G.gemm = spy, spy_g
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    agg = defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            a = agg[ev.name[:110]]
            a[0] += ev.device_time
            a[1] += 1
    by_shape = {}
    if impl == "native":
        G.gemm_tn, G.gemm = real, real_g
        evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA
                      and "gemm_tn_kernel" in e.name), key=lambda e: e.time_range.start)
        if len(evs) == len(shapes):
            acc = defaultdict(lambda: [0.0, 0])
            for e, sh in zip(evs, shapes):
                acc[sh][0] += e.device_time
                acc[sh][1] += 1
            by_shape = {f"{m}x{n}x{k}": {"calls": c, "avg_us": round(t / c, 1),
                                          "tflops": round(2 * m * n * k * c / t / 1e6, 1)}
                        for (m, n, k), (t, c) in sorted(acc.items(), key=lambda kv: -kv[1][0])}
        else:
            by_shape = {"mismatch": [len(evs), len(shapes)]}
    rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
    total = sum(v[0] for _, v in rows)
    out = {"impl": impl, "total_kernel_ms": round(total / 1e3, 3), "gemm_by_shape": by_shape,
           "kernels": [{"name": k, "ms": round(v[0] / 1e3, 3), "calls": v[1]} for k, v in rows[:top]]}
    del model
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="large")
    ap.add_argument("--batch-size", type=int, default=16)
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--impls", default="native,stock")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    res = [run(i, args.model, args.batch_size, args.seq_len, args.top) for i in args.impls.split(",")]
    for r in res:
        print(f"== {r['impl']}: {r['total_kernel_ms']} ms of kernels")
        for sh, v in r["gemm_by_shape"].items():
            print(f"  gemm {sh}: {v}")
        for k in r["kernels"]:
            print(f"  {k['ms']:9.3f} ms  x{k['calls']:<5d} {k['name']}")
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
