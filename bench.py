#!/usr/bin/env python
"""Headline benchmark: ResNet-50 img/s via the PyTorchJob replica runtime (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # our runtime
    python bench.py --impl reference ...                      # the unmodified reference (unavailable: see DESIGN.md)
    python bench.py --impl stock ...                          # stock torch DDP + NCCL on the same box

For N > 1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 --master-port P bench.py --gpus N ...`` (one rank per GPU).
Rank 0 prints ONE JSON line. Synthetic uint8 224x224 images, random-init weights, bf16.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

METRIC = "ResNet-50 img/s via PyTorchJob (bf16, synthetic 224x224), whole job"


def _reference_arm():
    # The reference is a Go Kubernetes operator plus a thin Python SDK. Neither
    # contains a trainer, and neither can run here (DESIGN.md "Reference arm").
    print(json.dumps({
        "impl": "reference",
        "unavailable": "reference is a Go k8s operator (no Go toolchain, no kube-apiserver); "
                       "`pip install /root/reference` fails (no setup.py/pyproject at the root); its "
                       "sdk/python installs with --no-deps but imports the absent `kubernetes` "
                       "package and needs a live cluster; it ships no ResNet-50/PyTorch trainer",
    }))
    return 0


def _through_the_product(args, world):
    """The same job through the product's public API: a PyTorchJob submitted with
    PyTorchJobClient to a LocalCluster (operator + launcher + rendezvous broker), one replica
    process per GPU. Returns the launch latencies on real GPU replicas (p50 of --launch-trials)
    and the throughput the launched ResNet-50 job reports (device-timed, max over ranks)."""
    import tempfile
    import yaml
    from tf_operator_b200.cluster import LocalCluster
    from tf_operator_b200.launcher.kubelet import KubeletOptions
    from tf_operator_b200.sdk import PyTorchJobClient
    from tf_operator_b200.testing import launch_probe
    repo = os.path.dirname(os.path.abspath(__file__))
    out = {}
    cl = LocalCluster(kubelet_options=KubeletOptions(cwd=repo))
    try:
        if args.launch_trials > 0:
            out["launch"] = launch_probe.measure(cl, world, args.launch_trials)
        if not args.no_via_launcher:
            job = yaml.safe_load(open(os.path.join(repo, "examples", "resnet50_ddp", "pytorchjob.yaml")))
            res = os.path.join(tempfile.mkdtemp(prefix="tfop-bench-"), "result.json")
            cmd = ["python", "examples/resnet50_ddp/train.py", f"--steps={args.steps}",
                   f"--warmup={max(args.warmup, 3)}", f"--device-steps={args.steps}",
                   f"--batch-size={args.batch_size}", f"--image-size={args.image_size}",
                   f"--depth={args.depth}", f"--result-file={res}"]
            specs = job["spec"]["pytorchReplicaSpecs"]
            for rt in list(specs):
                specs[rt]["template"]["spec"]["containers"][0]["command"] = cmd
                specs[rt]["restartPolicy"] = "Never"
            if world > 1:
                specs["Worker"]["replicas"] = world - 1
            else:
                specs.pop("Worker", None)
            cli = PyTorchJobClient(cluster=cl)
            t0 = time.time()
            cli.create(job)
            name = job["metadata"]["name"]
            cli.wait_for_job(name, polling_interval=0.2, timeout_seconds=600)
            if not cli.is_job_succeeded(name):
                raise RuntimeError(f"launched job failed: {cli.get_logs(name, master=True)}")
            r = json.load(open(res))
            out["via_launcher"] = {"value": r["device_img_per_s"], "unit": "img/s",
                                   "ms_per_step": r["device_ms_per_step"],
                                   "e2e_value": r["img_per_s"], "e2e_ms_per_step": r["ms_per_step"],
                                   "allreduce": r["allreduce"], "job_wall_s": round(time.time() - t0, 1),
                                   "api": "PyTorchJobClient.create -> LocalCluster -> "
                                          "examples/resnet50_ddp/train.py replicas"}
    finally:
        cl.stop()
    return out


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._t = threading.Thread(target=self._read, daemon=True)
            self._t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference", "stock"])
    ap.add_argument("--batch-size", type=int, default=256, help="per-GPU batch")
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--depth", type=int, default=50)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--no-fused-bn", action="store_true")
    ap.add_argument("--no-s2d-stem", action="store_true",
                    help="feed the stem conv plain NHWC x3 input instead of the space-to-depth layout")
    ap.add_argument("--no-nvls", action="store_true",
                    help="use the P2P two-shot all-reduce even when NVLS multicast is available")
    ap.add_argument("--comm-blocks", type=int, default=32)
    ap.add_argument("--bucket-mb", type=int, default=16)
    ap.add_argument("--diag-step-times", action="store_true",
                    help="DIAGNOSTIC: also record every step's device time on every rank and report "
                         "the spread (how much a per-step synchronisation of jittering GPUs costs)")
    ap.add_argument("--diag-no-allreduce", action="store_true",
                    help="DIAGNOSTIC, not a benchmark: skip the gradient all-reduce so every rank "
                         "free-runs; shows how far apart the GPUs are on their own")
    ap.add_argument("--launch-trials", type=int, default=10,
                    help="PyTorchJobs of N real GPU replicas launched for the launch-to-Running / "
                         "launch-to-first-step p50 (0 = skip)")
    ap.add_argument("--no-via-launcher", action="store_true",
                    help="skip re-running the config as a PyTorchJob through the operator + launcher")
    ap.add_argument("--stock-mode", default="autocast", choices=["autocast", "purebf16"],
                    help="--impl stock: fp32 weights under torch.autocast(bf16) (the usual recipe) "
                         "or a model cast to bf16")
    args = ap.parse_args()
    if args.impl == "reference":
        return _reference_arm()
    args.warmup = max(args.warmup, 3)

    # Libraries (NCCL_DEBUG=VERSION, torchrun) may print to stdout; keep the real stdout
    # for the single JSON line and send everything else to stderr.
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "error": "no CUDA device"}))
        return 1
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        host_pg = dist.new_group(backend="gloo")     # waits that must not occupy the GPUs
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(1234 + rank)

    B, S = args.batch_size, args.image_size
    # synthetic host batches in pinned memory (uint8 NHWC, like a decoded JPEG batch)
    n_host = 4
    host_x = [torch.randint(0, 256, (B, S, S, 3), dtype=torch.uint8).pin_memory()
              for _ in range(n_host)]
    host_y = [torch.randint(0, 1000, (B,), dtype=torch.int64).pin_memory() for _ in range(n_host)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def host_barrier():
        # the launched replicas below run on these same GPUs: a rank waiting in an NCCL barrier
        # would time-slice its GPU with them
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=host_pg)

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    launches_per_step = 0
    use_nvls = False
    if args.impl == "native":
        from tf_operator_b200.models.resnet import ResNet
        from tf_operator_b200.parallel.engine import DataParallelEngine
        from tf_operator_b200.runtime.symm import SymmetricComm

        model = ResNet(args.depth, fused=not args.no_fused_bn)
        comm = None
        use_nvls = False
        rdv_server = None
        if world > 1:
            # The peer handles (CUDA IPC blobs, the NVLS multicast descriptor) travel through the
            # launcher's own rendezvous store and descriptor broker - the code a PyTorchJob
            # replica runs - hosted here by rank 0; torch.distributed only carries its address.
            from tf_operator_b200.launcher.rendezvous import RendezvousServer
            from tf_operator_b200.runtime.env import rendezvous_comm
            where = [None]
            if rank == 0:
                rdv_server = RendezvousServer().start()
                where = [(rdv_server.addr, rdv_server.uds_path)]
            dist.broadcast_object_list(where, src=0)
            nparam = sum((p.numel() + 7) // 8 * 8 for p in model.parameters())
            comm = rendezvous_comm(where[0][0], where[0][1], f"bench/resnet/{os.getppid()}/comm0",
                                   rank, world, nparam * 2 + (8 << 20),
                                   nvls_bytes=0 if args.no_nvls else nparam * 2 + (1 << 20))
            use_nvls = comm.nvls_bytes > 0
            if args.diag_no_allreduce:
                comm.all_reduce_nvls = lambda *a, **k: None
                comm.all_reduce = lambda *a, **k: None
        eng = DataParallelEngine(model, comm=comm, batch_size=B, image_size=S, lr=0.1,
                                 momentum=0.9, weight_decay=1e-4, graph=not args.no_graph,
                                 overlap=not args.no_overlap, comm_blocks=args.comm_blocks,
                                 bucket_bytes=args.bucket_mb << 20,
                                 stem_s2d=not args.no_s2d_stem, use_nvls=use_nvls)

        def device_step():
            eng.run_device_step()

        def e2e_loop(k):
            pending = None
            eng.prefetch(host_x[0], host_y[0])
            last = 0.0
            for i in range(k):
                h = eng.step()
                if i + 1 < k:
                    eng.prefetch(host_x[(i + 1) % n_host], host_y[(i + 1) % n_host])
                if pending is not None:
                    last = pending.result()
                pending = h
            last = pending.result()
            return last

        # first batch resident for the device-only region
        eng.x_u8.copy_(host_x[0])
        eng.labels.copy_(host_y[0])
        torch.cuda.synchronize()
        main_stream = eng.main_stream
        h2d, d2h = eng.h2d_bytes_per_step(), eng.d2h_bytes_per_step()
    else:
        # "reference mode": stock torchvision-style graph, torch DDP over NCCL, torch SGD
        from tf_operator_b200.models.resnet import ResNet
        model = ResNet(args.depth, fused=False).to(dev).to(memory_format=torch.channels_last)
        amp = args.stock_mode == "autocast"
        if not amp:
            model = model.to(torch.bfloat16)
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank],
                                                              gradient_as_bucket_view=True)
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
        mean = torch.tensor([0.485, 0.456, 0.406], device=dev).view(1, 3, 1, 1) * 255
        std = torch.tensor([0.229, 0.224, 0.225], device=dev).view(1, 3, 1, 1) * 255
        x_dev = host_x[0].to(dev)
        y_dev = host_y[0].to(dev)
        main_stream = torch.cuda.current_stream()

        def train(xu8, y):
            x = ((xu8.permute(0, 3, 1, 2).float() - mean) / std).to(torch.bfloat16)
            x = x.contiguous(memory_format=torch.channels_last)
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=amp):
                loss = torch.nn.functional.cross_entropy(model(x).float(), y)
            loss.backward()
            opt.step()
            return loss

        def device_step():
            train(x_dev, y_dev)

        def e2e_loop(k):
            last = 0.0
            for i in range(k):
                xu8 = host_x[i % n_host].to(dev, non_blocking=True)
                y = host_y[i % n_host].to(dev, non_blocking=True)
                last = float(train(xu8, y).item())
            return last

        h2d, d2h = B * S * S * 3 + B * 8, 4

    # ---------------- device-timed region ----------------
    for _ in range(args.warmup):
        device_step()
    barrier()
    if args.impl == "native":
        launches_per_step = eng.kernel_launches_per_step
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    marks = []
    e0.record(main_stream)
    for _ in range(args.steps):
        device_step()
        if args.diag_step_times:
            marks.append(torch.cuda.Event(enable_timing=True))
            marks[-1].record(main_stream)
    e1.record(main_stream)
    barrier()
    ms_mine = e0.elapsed_time(e1)
    ms_total = max_over_ranks(ms_mine)
    per_rank = [ms_mine / args.steps]
    if world > 1:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, ms_mine / args.steps)
    step_stats = None
    if args.diag_step_times:
        mine = [a.elapsed_time(b) for a, b in zip([e0] + marks[:-1], marks)]
        allr = [mine]
        if world > 1:
            allr = [None] * world
            dist.all_gather_object(allr, mine)
        if rank == 0:
            import statistics as st
            step_stats = {
                "per_rank_mean_ms": [round(st.mean(r), 4) for r in allr],
                "per_rank_stdev_ms": [round(st.pstdev(r), 4) for r in allr],
                "per_rank_max_ms": [round(max(r), 4) for r in allr],
                # what a per-step synchronisation of these free-running ranks would cost
                "mean_of_stepwise_max_ms": round(st.mean(max(c) for c in zip(*allr)), 4),
                "max_of_means_ms": round(max(st.mean(r) for r in allr), 4)}
    clocks = sampler.stop() if sampler else None

    # ---------------- end-to-end region (pinned host -> device every step, loss read back) ---
    e2e_loop(max(2, args.warmup // 2))
    barrier()
    t0 = time.perf_counter()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(main_stream)
    last_loss = e2e_loop(args.steps)
    f1.record(main_stream)
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max_over_ranks(max(f0.elapsed_time(f1), wall_ms))
    barrier()

    # ---------------- the same job through the product (operator + launcher), rank 0 drives ----
    product = {}
    if args.impl == "native" and rank == 0 and (args.launch_trials > 0 or not args.no_via_launcher):
        try:
            product = _through_the_product(args, world)
        except Exception as e:  # noqa: BLE001 - the headline numbers above stand on their own
            product = {"error": f"{type(e).__name__}: {e}"[:500]}
    host_barrier()

    if rank == 0:
        gb = B * world
        ms_step = ms_total / args.steps
        out = {
            "metric": METRIC, "value": round(gb / (ms_step / 1e3), 2), "unit": "img/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": args.impl,
            "config": {"model": f"ResNet-{args.depth}", "global_batch": gb, "per_gpu_batch": B,
                       "image": f"{S}x{S}x3 uint8", "seq_len": None,
                       "parallelism": f"dp{world}", "optimizer": "SGD momentum 0.9 wd 1e-4",
                       "l2": "per-step working set (>10 GB of activations) far exceeds the "
                             "126 MB L2; host inputs rotate over 4 pinned batches",
                       "cuda_graph": args.impl == "native" and not args.no_graph,
                       "fused_bn": args.impl == "native" and not args.no_fused_bn,
                       "stem": "space-to-depth 4x4 conv (same 7x7 parameters)"
                               if args.impl == "native" and not args.no_s2d_stem else "7x7/2",
                       "allreduce": ("nvls multimem in place" if (args.impl == "native" and world > 1 and use_nvls)
                                     else "p2p two-shot in place" if (args.impl == "native" and world > 1)
                                     else None),
                       "stock_mode": args.stock_mode if args.impl == "stock" else None},
            "clocks": clocks,
            "e2e": {"value": round(gb / (e2e_ms / args.steps / 1e3), 2), "unit": "img/s",
                    "ms_per_step": round(e2e_ms / args.steps, 4),
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "last_loss": round(last_loss, 4)},
            "gpu_launches": launches_per_step * args.steps,
        }
        if world > 1:
            out["ms_per_step_per_rank"] = [round(x, 4) for x in per_rank]
        if step_stats:
            out["step_time_spread"] = step_stats
        if args.diag_no_allreduce:
            out["diagnostic"] = "gradient all-reduce skipped: NOT a benchmark number"
        if product:
            out.update({k: v for k, v in product.items()})
            if "launch" in product:
                out["launch_to_running_ms"] = product["launch"]["launch_to_running_ms"]
                out["launch_to_first_step_ms"] = product["launch"]["launch_to_first_step_ms"]
        real_stdout.write(json.dumps(out) + "\n")
        real_stdout.flush()
    if args.impl == "native":
        if comm is not None:
            comm.close()                 # exit barrier at the rendezvous store, then free
        if rdv_server is not None:
            rdv_server.stop()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
