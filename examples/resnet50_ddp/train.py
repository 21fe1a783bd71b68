#!/usr/bin/env python
"""ResNet-50 data-parallel training as a PyTorchJob replica (BASELINE.json config 2:
1 Master + 7 Workers, one B200 per replica, bf16, synthetic 224x224 images).

The replica reads the operator's env contract (MASTER_ADDR / MASTER_PORT / WORLD_SIZE / RANK,
pkg/controller.v1/pytorch/pytorch.go:41-65), joins its gang's rendezvous generation at the
launcher, maps its peers' symmetric gradient buffers (CUDA IPC) and the NVSwitch multicast buffer
(descriptor passed through the launcher's broker) and trains with the hand-written all-reduce
kernels. No torch.distributed process group, no NCCL.

Restart contract (pkg/controller.v1/tensorflow/pod.go:140-159, restartPolicy ExitCode): when a
peer dies, this replica's next gradient exchange times out, the step handle raises CommError and
the process exits with code 138 (retryable) - the operator re-creates every replica of the gang,
they meet in the next rendezvous generation and resume from ``--checkpoint``.

``TFOP_FAULT_INJECT=<rank>:<step>:<marker file>`` kills that rank with SIGKILL at that step, once
(the marker file records that it happened): the e2e restart test's fault injector, the analogue
of the reference test-server's ``/exit?exitCode=`` (test/test-server/test_app.py:47-53).
"""
import argparse
import json
import os
import signal
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.models.resnet import ResNet  # noqa: E402
from tf_operator_b200.parallel.engine import DataParallelEngine  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402
from tf_operator_b200.runtime.symm import CommError  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--image-size", type=int, default=224)
    ap.add_argument("--depth", type=int, default=50)
    ap.add_argument("--classes", type=int, default=1000)
    ap.add_argument("--lr", type=float, default=0.1)
    ap.add_argument("--no-nvls", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--checkpoint", default="", help="resume from / save to this file (rank 0 saves)")
    ap.add_argument("--checkpoint-every", type=int, default=0)
    ap.add_argument("--comm-timeout-ms", type=int, default=0)
    ap.add_argument("--result-file", default="", help="rank 0 writes its final JSON line here too")
    ap.add_argument("--device-steps", type=int, default=0,
                    help="also time this many steps on the resident batch (no host traffic)")
    args = ap.parse_args()

    t_start = time.time()
    ctx = ReplicaContext.from_env()
    torch.cuda.set_device(0)            # the launcher exposes exactly one GPU per replica
    torch.manual_seed(1234)             # identical initial weights on every replica
    model = ResNet(args.depth, num_classes=args.classes)
    nparam = sum((p.numel() + 7) // 8 * 8 for p in model.parameters())
    comm = None
    if ctx.world_size > 1:
        comm = ctx.symmetric_comm(nparam * 2 + (8 << 20),
                                  nvls_bytes=0 if args.no_nvls else nparam * 2 + (1 << 20))
        if args.comm_timeout_ms:
            comm.set_timeout_ms(args.comm_timeout_ms)
    use_nvls = comm is not None and comm.nvls_bytes > 0
    eng = DataParallelEngine(model, comm=comm, batch_size=args.batch_size, image_size=args.image_size,
                             lr=args.lr * ctx.world_size, stem_s2d=True, use_nvls=use_nvls,
                             graph=not args.no_graph)
    start_step = 0
    if args.checkpoint and os.path.exists(args.checkpoint):
        start_step = eng.load_checkpoint(args.checkpoint)
    if comm is not None and getattr(comm, "_rdv", None) is not None:
        # meet once more before the first collective: building the engine and loading the
        # checkpoint take different times on different replicas, and the first all-reduce should
        # not spend its spin limit on that. A peer that is gone by now means another restart round.
        from tf_operator_b200.launcher.rendezvous import Superseded
        cli, prefix = comm._rdv
        scope, gen = getattr(comm, "_rdv_gen", ("", -1))
        try:
            cli.barrier(prefix + "/start", ctx.world_size, timeout=120, scope=scope, gen=gen)
        except (Superseded, TimeoutError) as e:
            print(f"rank {ctx.rank}: gang incomplete before the first step ({type(e).__name__}); "
                  f"exiting 138 (retryable)", flush=True)
            os._exit(138)
    torch.manual_seed(100 + ctx.rank)
    B, S = args.batch_size, args.image_size
    xs = [torch.randint(0, 256, (B, S, S, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
    ys = [torch.randint(0, args.classes, (B,), dtype=torch.int64).pin_memory() for _ in range(2)]
    print(f"rank {ctx.rank}/{ctx.world_size} on {torch.cuda.get_device_name(0)} "
          f"generation {getattr(comm, 'generation', 0)} resume_step {start_step} "
          f"allreduce {'nvls' if use_nvls else 'p2p' if comm else 'none'}", flush=True)

    fault = os.environ.get("TFOP_FAULT_INJECT", "")
    f_rank, f_step, f_marker = (fault.split(":", 2) + ["", ""])[:3] if fault else ("", "", "")

    def maybe_die(step):
        if fault and int(f_rank) == ctx.rank and step == int(f_step) and not os.path.exists(f_marker):
            open(f_marker, "w").write(str(os.getpid()))
            print(f"fault injection: rank {ctx.rank} kills itself at step {step}", flush=True)
            os.kill(os.getpid(), signal.SIGKILL)

    def max_over_ranks(tag, value):
        """device-timed numbers are the slowest rank's (values meet at the rendezvous store)"""
        if comm is None or getattr(comm, "_rdv", None) is None:
            return value
        cli, prefix = comm._rdv
        vals = cli.all_gather(f"{prefix}/max/{tag}", ctx.rank, ctx.world_size, repr(value).encode())
        return max(float(v) for v in vals)

    dev_ms = None
    try:
        first = True
        for _ in range(args.warmup):
            eng.step(xs[0], ys[0]).result()
            if first and ctx.rank == 0:
                # create -> first completed step (rendezvous, handle mapping, first all-reduce)
                print(json.dumps({"event": "first_step", "unix_time": time.time(),
                                  "since_process_start_s": round(time.time() - t_start, 3)}),
                      flush=True)
            first = False
        torch.cuda.synchronize()
        if args.device_steps:
            d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if comm is not None:
                comm._exit_barrier()
            d0.record(eng.main_stream)
            for _ in range(args.device_steps):
                eng.run_device_step()
            d1.record(eng.main_stream)
            torch.cuda.synchronize()
            dev_ms = max_over_ranks("dev", d0.elapsed_time(d1) / args.device_steps)
            if comm is not None:
                comm.raise_if_failed()
                comm._exit_barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(eng.main_stream)
        pending, loss = None, float("nan")
        for i in range(start_step, args.steps):
            maybe_die(i)
            h = eng.step(xs[i % 2], ys[i % 2])
            if pending is not None:
                loss = pending.result()
                if ctx.rank == 0 and i % 10 == 0:
                    print(f"step {i} loss {loss:.4f}", flush=True)
            pending = h
            if args.checkpoint and args.checkpoint_every and (i + 1) % args.checkpoint_every == 0:
                loss = pending.result()
                pending = None
                if ctx.rank == 0:
                    eng._steps = i + 1
                    eng.save_checkpoint(args.checkpoint)
        if pending is not None:
            loss = pending.result()
        e1.record(eng.main_stream)
        torch.cuda.synchronize()
    except CommError as e:
        print(f"rank {ctx.rank}: {e}; exiting {e.exit_code} (retryable)", flush=True)
        os._exit(e.exit_code)           # peers are gone: no exit barrier, no destructor
    ms = max_over_ranks("e2e", e0.elapsed_time(e1) / max(1, args.steps - start_step))
    if ctx.rank == 0:
        line = json.dumps({"img_per_s": round(B * ctx.world_size / (ms / 1e3), 1),
                           "ms_per_step": round(ms, 3), "world": ctx.world_size,
                           "device_ms_per_step": round(dev_ms, 3) if dev_ms else None,
                           "device_img_per_s": round(B * ctx.world_size / (dev_ms / 1e3), 1) if dev_ms else None,
                           "kernel_launches_per_step": eng.kernel_launches_per_step,
                           "h2d_bytes_per_step": eng.h2d_bytes_per_step(),
                           "d2h_bytes_per_step": eng.d2h_bytes_per_step(),
                           "allreduce": "nvls" if use_nvls else ("p2p" if comm else None),
                           "generation": getattr(comm, "generation", 0),
                           "resumed_from": start_step, "final_loss": round(loss, 4)})
        print(line, flush=True)
        if args.result_file:
            with open(args.result_file, "w") as f:
                f.write(line + "\n")
    if comm is not None:
        comm.close()                    # device sync + rendezvous exit barrier, then free
    return 0


if __name__ == "__main__":
    sys.exit(main())
