#!/usr/bin/env python
"""Back-to-back GEMM throughput (no flush, no sync between calls) with SM clock and board power
sampled through NVML: separates burst speed (kernel_bench.py) from what a training step sees
when the tensor cores stay busy for tens of milliseconds."""
import argparse
import json
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_operator_b200.ops.gemm import gemm_tn  # noqa: E402


class Sampler(threading.Thread):
    def __init__(self):
        super().__init__(daemon=True)
        import pynvml
        pynvml.nvmlInit()
        self.nv = pynvml
        self.h = pynvml.nvmlDeviceGetHandleByIndex(0)
        self.rows = []
        self.stop = False

    def run(self):
        while not self.stop:
            try:
                self.rows.append((self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM),
                                  self.nv.nvmlDeviceGetPowerUsage(self.h) / 1e3))
            except Exception:
                pass
            time.sleep(0.005)


def sustained(fn, iters):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    s = Sampler()
    s.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    s.stop = True
    s.join()
    clk = sorted(r[0] for r in s.rows) or [0]
    pw = sorted(r[1] for r in s.rows) or [0]
    return e0.elapsed_time(e1) / iters * 1e3, clk[len(clk) // 2], pw[len(pw) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=400)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    res = []
    for (m, n, k) in [(8192, 3072, 1024), (8192, 1024, 4096), (4096, 1024, 8192), (8192, 8192, 8192)]:
        it = args.iters if m * n * k < 1 << 38 else args.iters // 8
        for data in ("randn", "small"):
            a = torch.randn(m, k, device="cuda").bfloat16()
            b = torch.randn(n, k, device="cuda").bfloat16()
            if data == "small":      # activations / weights of a real net: narrow value range
                a, b = (a * 0.02).bfloat16(), (b * 0.02).bfloat16()
            row = {"m": m, "n": n, "k": k, "data": data}
            for name, fn in (("tcgen05", lambda: gemm_tn(a, b)), ("cublas", lambda: torch.matmul(a, b.t()))):
                us, clk, pw = sustained(fn, it)
                row[name] = {"us": round(us, 1), "tflops": round(2 * m * n * k / us / 1e6, 1),
                             "sm_mhz": clk, "watts": round(pw)}
            res.append(row)
            print(json.dumps(row), flush=True)
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
