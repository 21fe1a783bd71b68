#!/usr/bin/env python
"""BERT-large data-parallel pre-training as TFJob replicas (BASELINE.json config 3: TFJob
Chief + Workers, TF_CONFIG contract, bf16, bucketed fused all-reduce).

Ranks come from TF_CONFIG (chief first, then workers; pkg/controller.v1/tensorflow/
tensorflow.go:97-173). All projection / MLP GEMMs run on the hand-written tcgen05 kernel; the
gradient buckets are all-reduced in place in symmetric memory as backward produces them.
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.models.bert import BertConfig, BertForPreTraining  # noqa: E402
from tf_operator_b200.parallel.engine import FlatDataParallel  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="large", choices=["tiny", "base", "large"])
    ap.add_argument("--batch-size", type=int, default=16, help="sequences per GPU")
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "nccl"])
    ap.add_argument("--no-graph", action="store_true", help="run eager instead of one CUDA graph per step")
    ap.add_argument("--fused-wgrad", action="store_true",
                    help="weight gradients of the encoder's linear layers: GEMM and cross-rank sum in "
                         "ONE kernel (in-switch reduction) instead of GEMM + bucket all-reduce")
    ap.add_argument("--no-nvls", action="store_true",
                    help="reduce gradient buckets with the P2P two-shot kernel instead of NVLS multimem")
    ap.add_argument("--ckpt", default="", help="checkpoint file (a mounted path in the pod template): "
                    "loaded at start when it exists, written by rank 0 at the end - a replica the "
                    "operator restarts under the same name resumes from it")
    args = ap.parse_args()

    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    cfg = getattr(BertConfig, args.model)()
    cfg.tc_gemm = args.impl == "native"       # --impl nccl is the all-library baseline
    torch.manual_seed(1234)
    model = BertForPreTraining(cfg).to(dev).to(torch.bfloat16)
    nparam = sum((p.numel() + 7) // 8 * 8 for p in model.parameters())
    B, S = args.batch_size, min(args.seq_len, cfg.max_position)
    g = torch.Generator(device=dev).manual_seed(10 + ctx.rank)
    ids = torch.randint(0, cfg.vocab_size, (B, S), device=dev, generator=g)
    labels = torch.where(torch.rand(B, S, device=dev, generator=g) < 0.15, ids, torch.full_like(ids, -100))
    nsp = torch.randint(0, 2, (B,), device=dev, generator=g)

    use_nvls = False
    if args.impl == "nccl":
        if ctx.world_size > 1:
            ctx.init_process_group("nccl")
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[0])
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=0.01, fused=True)

        def step():
            opt.zero_grad(set_to_none=True)
            loss = model(ids, labels, nsp)
            loss.backward()
            opt.step()
            return loss.detach()
        launches = lambda: 0  # noqa: E731
    else:
        comm = None
        if ctx.world_size > 1:
            # gradient bucket in symmetric memory; with NVLS it is also bound to a multicast
            # object so the NVSwitch reduces it in place (comm.nvls_bytes == 0 when unsupported)
            comm = ctx.symmetric_comm(nparam * 2 + (8 << 20),
                                      nvls_bytes=0 if args.no_nvls else nparam * 2 + (1 << 20))
        use_nvls = comm is not None and getattr(comm, "nvls_bytes", 0) > 0
        dp = FlatDataParallel(model, comm=comm, optimizer="adamw", lr=1e-4, weight_decay=0.01,
                              bucket_bytes=64 << 20, use_nvls=use_nvls,
                              fused_wgrad=args.fused_wgrad)

        if args.ckpt and os.path.exists(args.ckpt):
            print(json.dumps({"resumed_from_step": dp.load_checkpoint(args.ckpt)}), flush=True)

        graph = {"g": None, "loss": None}

        def eager():
            return dp.train_step(lambda: model(ids, labels, nsp))

        def step():
            if args.no_graph:
                return eager()
            if graph["g"] is None:
                side = torch.cuda.Stream()
                with torch.cuda.stream(side):
                    for _ in range(2):
                        eager()
                side.synchronize()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    graph["loss"] = eager()
                graph["g"] = g
            # the AdamW step counter lives in device memory: bump it, then replay
            dp._hp_host[7] = float(dp._steps + 1)
            dp.hp.copy_(dp._hp_host, non_blocking=True)
            dp._steps += 1
            graph["g"].replay()
            return graph["loss"]
        launches = lambda: dp.kernel_launches_per_step  # noqa: E731

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    out = {"role": ctx.task_type, "rank": ctx.rank, "world": ctx.world_size, "impl": args.impl,
           "model": f"bert-{args.model}", "params_M": round(nparam / 1e6, 1), "seq_len": S,
           "per_gpu_batch": B, "ms_per_step": round(ms, 3),
           "seq_per_s_job": round(B * ctx.world_size / ms * 1e3, 1), "loss": round(float(loss), 4),
           "kernel_launches_per_step": launches(),
           "fused_wgrad": bool(getattr(dp, "fused_wgrad", False)) if args.impl == "native" else False,
           "allreduce": ("nvls" if (args.impl == "native" and ctx.world_size > 1 and use_nvls) else
                         "p2p two-shot" if (args.impl == "native" and ctx.world_size > 1) else
                         "nccl" if ctx.world_size > 1 else None)}
    print(json.dumps(out), flush=True)
    if args.ckpt and args.impl == "native" and ctx.rank == 0:
        dp.save_checkpoint(args.ckpt)
    time.sleep(0.5)
    return 0


if __name__ == "__main__":
    sys.exit(main())
