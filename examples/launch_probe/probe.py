#!/usr/bin/env python
"""Smallest real GPU replica: what stands between `create` and a job's first collective.

Reads the operator's env contract, creates its CUDA context, allocates a symmetric buffer, joins
its gang's rendezvous generation at the launcher, maps every peer's buffer (CUDA IPC) and runs one
in-place two-shot all-reduce over NVLink - then prints the wall-clock time at which that
all-reduce had completed. No torch import (it alone costs seconds): the C ABI of libtfop_b200 is
driven through ctypes, exactly as include/tfop_b200.h documents it.

Used by bench.py / benchmarks/launch_latency.py for `launch_to_first_step_ms` (BASELINE.json:
"job launch-to-Running p50"; the reference's own observation cadence is 30 s polling,
py/kubeflow/tf_operator/tf_job_client.py:121-122).
"""
import ctypes as C
import json
import os
import sys
import time

T0 = time.time()
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from tf_operator_b200.launcher.rendezvous import RendezvousClient  # noqa: E402
from tf_operator_b200.ops import _lib  # noqa: E402
from tf_operator_b200.runtime.env import ReplicaContext  # noqa: E402

TWO_SHOT, BF16 = 2, 1


def main():
    ctx = ReplicaContext.from_env()
    lib = _lib.load()
    nbytes = 4 << 20
    h = C.c_void_p()
    _lib.check(lib.tfop_comm_create(ctx.rank, ctx.world_size, nbytes, C.byref(h)))   # CUDA context
    t_ctx = time.time()
    world = ctx.world_size
    if world > 1:
        hb = lib.tfop_comm_handle_bytes()
        mine = C.create_string_buffer(hb)
        _lib.check(lib.tfop_comm_get_handle(h, mine))
        cli = RendezvousClient(ctx.raw["TFOP_RENDEZVOUS"])
        scope = ctx.rendezvous_scope("probe")
        gen = cli.join(scope, ctx.rank, world)
        prefix = f"{scope}/g{gen}"
        handles = cli.all_gather(prefix + "/ipc", ctx.rank, world, mine.raw)
        _lib.check(lib.tfop_comm_open_peers(h, C.create_string_buffer(b"".join(handles), hb * world)))
        cli.barrier(prefix + "/ready", world)
    t_mapped = time.time()
    n = (1 << 20) // 2
    _lib.check(lib.tfop_allreduce(h, None, 0, None, 0, BF16, n, 0, 1.0 / world, TWO_SHOT, 0, None))
    flag = C.c_int(0)
    _lib.check(lib.tfop_comm_check_error(h, C.byref(flag)))          # device -> host read: syncs
    t_done = time.time()
    if flag.value:
        print("all-reduce timed out", flush=True)
        return 138
    if ctx.rank == 0:
        print(json.dumps({"event": "first_step", "unix_time": t_done,
                          "process_start_unix_time": T0,
                          "cuda_context_s": round(t_ctx - T0, 4),
                          "rendezvous_and_map_s": round(t_mapped - t_ctx, 4),
                          "first_allreduce_s": round(t_done - t_mapped, 4)}), flush=True)
    if world > 1:
        cli.barrier(prefix + "/exit", world, timeout=60)
        _lib.check(lib.tfop_comm_quiesce(h))
        cli.close()
    _lib.check(lib.tfop_comm_destroy(h))
    return 0


if __name__ == "__main__":
    sys.exit(main())
