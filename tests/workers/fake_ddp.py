#!/usr/bin/env python
"""CPU stand-in for examples/resnet50_ddp/train.py in the control-plane version of the
kill-and-restart test: the same life cycle (form the communicator through the launcher's
rendezvous store, step in lock-step with the gang, checkpoint, SIGKILL by fault injection, exit 138
when a peer is gone, resume from the checkpoint in the next generation) with the GPU parts faked:
the "communicator" only exchanges strings and a step is a store barrier with a time-out."""
import json
import os
import signal
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import tf_operator_b200.runtime.symm as symm  # noqa: E402
from tf_operator_b200.launcher.rendezvous import Replaced, Superseded  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402


class FakeComm:
    def __init__(self, rank, world, handles):
        self.rank, self.world, self.handles, self.nvls_bytes = rank, world, handles, 0

    @classmethod
    def create(cls, rank, world, nbytes, exchange, device=None):
        return cls(rank, world, exchange(f"handle-of-{rank}-{os.getpid()}".encode()))

    def enable_nvls(self, nbytes):
        cli, prefix = self._rdv
        scope, gen = self._rdv_gen
        for phase in ("created", "added", "bound"):
            cli.barrier(f"{prefix}/nvls/{phase}", self.world, scope=scope, gen=gen)
        self.nvls_bytes = nbytes
        return True


def main():
    steps, ckpt, result = int(sys.argv[1]), sys.argv[2], sys.argv[3]
    step_timeout = float(sys.argv[4]) if len(sys.argv) > 4 else 2.0
    symm.SymmetricComm = FakeComm
    ctx = ReplicaContext.from_env()
    jitter = float(os.environ.get("FAKE_START_JITTER_S", "0"))
    if jitter:
        # what eight processes creating CUDA contexts at once look like: staggered, seconds apart
        import random
        time.sleep(random.Random(os.getpid()).uniform(0, jitter))
    comm = ctx.symmetric_comm(1 << 20, nvls_bytes=4096)
    cli, prefix = comm._rdv
    scope, gen = comm._rdv_gen
    start = json.load(open(ckpt))["step"] if os.path.exists(ckpt) else 0
    print(f"rank {ctx.rank}/{ctx.world_size} generation {comm.generation} resume_step {start}", flush=True)
    fault = os.environ.get("TFOP_FAULT_INJECT", "")
    f_rank, f_step, f_marker = (fault.split(":", 2) + ["", ""])[:3] if fault else ("", "", "")
    try:
        cli.barrier(prefix + "/start", ctx.world_size, timeout=60, scope=scope, gen=gen)   # as train.py
        for step in range(start, steps):
            if fault and int(f_rank) == ctx.rank and step == int(f_step) and not os.path.exists(f_marker):
                open(f_marker, "w").write(str(os.getpid()))
                print(f"fault injection: rank {ctx.rank} kills itself at step {step}", flush=True)
                os.kill(os.getpid(), signal.SIGKILL)
            time.sleep(0.005)                                   # the step's compute
            cli.barrier(f"{prefix}/step", ctx.world_size, timeout=step_timeout, scope=scope, gen=gen)
            if ctx.rank == 0 and (step + 1) % 10 == 0:
                tmp = ckpt + ".tmp"
                json.dump({"step": step + 1}, open(tmp, "w"))
                os.replace(tmp, ckpt)
    except (TimeoutError, Superseded, Replaced) as e:
        print(f"rank {ctx.rank}: peer lost ({type(e).__name__}); exiting 138 (retryable)", flush=True)
        os._exit(138)
    if ctx.rank == 0:
        json.dump({"world": ctx.world_size, "generation": comm.generation, "resumed_from": start},
                  open(result, "w"))
    comm._exit_barrier()
    return 0


if __name__ == "__main__":
    sys.exit(main())
