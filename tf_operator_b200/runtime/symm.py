"""Symmetric peer memory + fused P2P all-reduce front-end (csrc/symm.cu, csrc/allreduce.cu).

One ``SymmetricComm`` per replica process. Each rank allocates a data buffer and
a signal pad, publishes their CUDA IPC handles through any byte-exchange
callable (the launcher's rendezvous store, or torch.distributed), and maps all
peers. After that, collectives are single kernels that read/write peer HBM over
NVLink — NCCL is only the measured baseline.

Replaces the DNS + TCP addressing the reference hands to the workload
(pkg/controller.v1/pytorch/pytorch.go:13-68).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Optional, Sequence

import torch

from ..ops import _lib

AUTO, ONE_SHOT, TWO_SHOT = 0, 1, 2

# Exit code of a replica whose collective gave up on a peer. 138 (= 128 + SIGUSR1) is the
# "user-defined retryable" code of the operator's ExitCode restart policy
# (train_util.IsRetryableExitCode): the operator re-creates the replica, and
# because every surviving rank hits the same timeout, the whole gang re-forms.
RETRYABLE_EXIT_CODE = 138


_ORPHANS: list = []       # handles of communicators whose formation was interrupted


class CommError(RuntimeError):
    """A cross-rank barrier of a collective timed out: a peer is dead or wedged. The collective
    was abandoned (no data was reduced or stored); the replica should exit with
    ``RETRYABLE_EXIT_CODE`` so the operator restarts it."""
    exit_code = RETRYABLE_EXIT_CODE


class _RawCudaBuffer:
    """Expose a raw device pointer to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2,
            "strides": None,
        }


def tensor_from_ptr(ptr: int, nbytes: int, device, owner=None) -> torch.Tensor:
    return torch.as_tensor(_RawCudaBuffer(ptr, nbytes, owner), device=device)


class SymmetricComm:
    def __init__(self, handle: int, rank: int, world: int, device: torch.device,
                 emulated: bool = False):
        self._h = C.c_void_p(handle)
        self.rank, self.world, self.device, self.emulated = rank, world, device, emulated
        self._lib = _lib.load()
        self.nbytes = self._lib.tfop_comm_buffer_bytes(self._h)
        self._views = {}

    # ---- construction ------------------------------------------------------
    @classmethod
    def create(cls, rank: int, world: int, nbytes: int,
               exchange: Callable[[bytes], Sequence[bytes]],
               device: Optional[torch.device] = None) -> "SymmetricComm":
        """``exchange(my_bytes) -> [bytes of rank 0, ..., bytes of rank world-1]``."""
        lib = _lib.load()
        device = device or torch.device("cuda", torch.cuda.current_device())
        with torch.cuda.device(device):
            h = C.c_void_p()
            _lib.check(lib.tfop_comm_create(rank, world, nbytes, C.byref(h)), "comm_create")
            hb = lib.tfop_comm_handle_bytes()
            mine = C.create_string_buffer(hb)
            _lib.check(lib.tfop_comm_get_handle(h, mine), "comm_get_handle")
            try:
                allh = exchange(mine.raw)
            except BaseException:
                # a peer may already hold this handle: the allocation stays until process exit
                _ORPHANS.append(h)
                raise
            assert len(allh) == world and all(len(b) == hb for b in allh)
            blob = C.create_string_buffer(b"".join(allh), hb * world)
            if world > 1:
                _lib.check(lib.tfop_comm_open_peers(h, blob), "comm_open_peers")
        return cls(h.value, rank, world, device)

    @classmethod
    def from_torch_distributed(cls, nbytes: int, group=None) -> "SymmetricComm":
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)

        def exchange(b: bytes):
            out: List[Optional[bytes]] = [None] * world
            dist.all_gather_object(out, b, group=group)
            return out

        comm = cls.create(rank, world, nbytes, exchange)
        dist.barrier(group=group)
        comm._exit_barrier = lambda: dist.barrier(group=group)
        return comm

    @classmethod
    def emulated(cls, world: int, nbytes: int,
                 device: Optional[torch.device] = None) -> List["SymmetricComm"]:
        """All ranks on ONE GPU in ONE process (tests; single cooperative launches)."""
        lib = _lib.load()
        device = device or torch.device("cuda", torch.cuda.current_device())
        arr = (C.c_void_p * world)()
        with torch.cuda.device(device):
            _lib.check(lib.tfop_comm_create_emulated(world, nbytes, arr), "comm_create_emulated")
        return [cls(arr[r], r, world, device, emulated=True) for r in range(world)]

    # ---- buffers -----------------------------------------------------------
    def buffer(self, peer: Optional[int] = None) -> torch.Tensor:
        """uint8 view of rank ``peer``'s symmetric buffer as mapped in this process."""
        peer = self.rank if peer is None else peer
        if peer not in self._views:
            p = self._lib.tfop_comm_buffer(self._h, peer)
            self._views[peer] = tensor_from_ptr(p, self.nbytes, self.device, owner=self)
        return self._views[peer]

    def view(self, byte_off: int, numel: int, dtype: torch.dtype,
             peer: Optional[int] = None) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        return self.buffer(peer)[byte_off:byte_off + numel * es].view(dtype)

    # ---- collectives -------------------------------------------------------
    def all_reduce(self, n: int, *, inp: Optional[torch.Tensor] = None,
                   out: Optional[torch.Tensor] = None, sym_off: int = 0,
                   wire_dtype: torch.dtype = torch.bfloat16, scale: float = 1.0,
                   algo: int = AUTO, max_blocks: int = 0, stream=None) -> None:
        """Fused sum-all-reduce * scale. ``inp=None``: data already in the symmetric buffer at
        ``sym_off`` (wire dtype). ``out=None``: leave the result there (two-shot, in place)."""
        assert not self.emulated, "use all_reduce_emulated for emulated groups"
        _lib.check(self._lib.tfop_allreduce(
            self._h, _lib.ptr(inp), _lib.dtype_tag(inp.dtype) if inp is not None else 0,
            _lib.ptr(out), _lib.dtype_tag(out.dtype) if out is not None else 0,
            _lib.dtype_tag(wire_dtype), n, sym_off, float(scale), algo, max_blocks,
            _lib.stream_handle(stream)), "allreduce")

    # ---- NVLS (multicast) ------------------------------------------------------
    @staticmethod
    def nvls_supported() -> bool:
        flag = C.c_int(0)
        rc = _lib.load().tfop_nvls_supported(C.byref(flag))
        return rc == 0 and bool(flag.value)

    def enable_nvls(self, nbytes: int, group=None) -> bool:
        """Create/import the multicast object, bind this rank's memory and map both views.
        Returns False when the box has no NVLS.

        A communicator that was formed through the launcher's rendezvous store (``_rdv`` set by
        ReplicaContext.symmetric_comm) passes the multicast file descriptor through the
        launcher's descriptor broker (SCM_RIGHTS over its Unix socket) and barriers on the store:
        no torch.distributed process group is needed. Otherwise (torchrun-style launches) the
        descriptor travels from rank 0 over an ad-hoc Unix socket and torch.distributed provides
        the barriers."""
        if getattr(self, "_rdv", None) is not None and group is None:
            return self._enable_nvls_brokered(nbytes)
        import os
        import socket
        import time

        import torch.distributed as dist
        self.nvls_bytes = 0
        if self.world == 1 or not self.nvls_supported():
            return False
        lib = self._lib
        cpu_group = dist.get_backend(group) == "gloo"
        ok = torch.tensor([1], device="cpu" if cpu_group else self.device)
        name = [f"\0tfop-nvls-{os.getpid()}-{id(self) & 0xffff}"]
        dist.broadcast_object_list(name, src=0, group=group)
        with torch.cuda.device(self.device):
            if self.rank == 0:
                fd = C.c_int(-1)
                rc = lib.tfop_nvls_create(self._h, nbytes, C.byref(fd))
                if rc != 0:
                    ok[0] = 0
                srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                srv.bind(name[0])
                srv.listen(self.world)
                dist.barrier(group=group)            # listening
                for _ in range(self.world - 1):
                    conn, _ = srv.accept()
                    if rc == 0:
                        socket.send_fds(conn, [b"fd"], [fd.value])
                    else:
                        conn.sendall(b"no")
                    conn.close()
                srv.close()
                if rc == 0:
                    os.close(fd.value)
            else:
                dist.barrier(group=group)
                cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                for _ in range(200):
                    try:
                        cli.connect(name[0])
                        break
                    except OSError:
                        time.sleep(0.05)
                msg, fds, _, _ = socket.recv_fds(cli, 16, 1)
                cli.close()
                if not fds or lib.tfop_nvls_import(self._h, nbytes, fds[0]) != 0:
                    ok[0] = 0
                for f in fds:
                    os.close(f)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                return False
            _lib.check(lib.tfop_nvls_add_device(self._h), "nvls_add_device")
            dist.barrier(group=group)
            _lib.check(lib.tfop_nvls_bind(self._h), "nvls_bind")
            dist.barrier(group=group)
        self.nvls_bytes = lib.tfop_nvls_bytes(self._h)
        return True

    def _enable_nvls_brokered(self, nbytes: int) -> bool:
        import os
        cli, prefix = self._rdv
        # every wait names the rendezvous generation this communicator belongs to: a peer that
        # restarts in the middle ends the wait with Superseded instead of a timeout
        sg = dict(zip(("scope", "gen"), getattr(self, "_rdv_gen", ("", -1))))
        self.nvls_bytes = 0
        if self.world == 1:
            return False
        lib, base = self._lib, f"{prefix}/nvls"
        ok = self.nvls_supported()
        with torch.cuda.device(self.device):
            if self.rank == 0:
                fd = C.c_int(-1)
                if ok and lib.tfop_nvls_create(self._h, nbytes, C.byref(fd)) == 0:
                    cli.fd_put(f"{base}/fd", fd.value)       # the broker keeps its own duplicate
                    os.close(fd.value)
                else:
                    ok = False
                cli.set(f"{base}/created", b"1" if ok else b"0")
            else:
                if cli.get(f"{base}/created", **sg) != b"1":
                    ok = False
                elif ok:
                    fd = cli.fd_get(f"{base}/fd", **sg)
                    try:
                        ok = lib.tfop_nvls_import(self._h, nbytes, fd) == 0
                    finally:
                        os.close(fd)
            # every rank must have the object before any binds; all or nothing
            votes = cli.all_gather(f"{base}/ok", self.rank, self.world, b"1" if ok else b"0", **sg)
            if any(v != b"1" for v in votes):
                return False
            _lib.check(lib.tfop_nvls_add_device(self._h), "nvls_add_device")
            cli.barrier(f"{base}/added", self.world, **sg)
            _lib.check(lib.tfop_nvls_bind(self._h), "nvls_bind")
            cli.barrier(f"{base}/bound", self.world, **sg)
        self.nvls_bytes = lib.tfop_nvls_bytes(self._h)
        return True

    def nvls_view(self, numel: int, dtype: torch.dtype, byte_off: int = 0) -> torch.Tensor:
        """This rank's copy of the multicast-bound buffer."""
        p = self._lib.tfop_nvls_local_ptr(self._h)
        es = torch.empty((), dtype=dtype).element_size()
        buf = tensor_from_ptr(p, self.nvls_bytes, self.device, owner=self)
        return buf[byte_off: byte_off + numel * es].view(dtype)

    def all_reduce_nvls(self, n: int, *, inp: Optional[torch.Tensor] = None,
                        out: Optional[torch.Tensor] = None, sym_off: int = 0,
                        wire_dtype: torch.dtype = torch.bfloat16, scale: float = 1.0,
                        max_blocks: int = 0, stream=None) -> None:
        """All-reduce reduced inside the NVSwitch (multimem.ld_reduce / multimem.st)."""
        _lib.check(self._lib.tfop_allreduce_nvls(
            self._h, _lib.ptr(inp), _lib.dtype_tag(inp.dtype) if inp is not None else 0,
            _lib.ptr(out), _lib.dtype_tag(out.dtype) if out is not None else 0,
            _lib.dtype_tag(wire_dtype), n, sym_off, float(scale), max_blocks,
            _lib.stream_handle(stream)), "allreduce_nvls")

    # ---- fused GEMM + all-reduce --------------------------------------------------
    @staticmethod
    def gemm_allreduce_workspace_bytes(m: int, n: int, world: int) -> int:
        return _lib.load().tfop_gemm_allreduce_workspace_bytes(m, n, world)

    def gemm_all_reduce(self, a: torch.Tensor, b: torch.Tensor, *, ws_off: int, out_off: int,
                        ldo: Optional[int] = None, scale: float = 1.0, max_ctas: int = 0,
                        stream=None) -> torch.Tensor:
        """out = scale * sum over ranks of (a_r @ b_r.T), computed and exchanged by ONE kernel
        (csrc/gemm_allreduce.cu). Returns this rank's view of the result inside its symmetric
        buffer (bf16 [M, N] at out_off)."""
        assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.shape[1] == b.shape[1]
        a = a if a.stride(1) == 1 and a.stride(0) % 8 == 0 else a.contiguous()
        b = b if b.stride(1) == 1 and b.stride(0) % 8 == 0 else b.contiguous()
        m, k = a.shape
        n = b.shape[0]
        ldo = ldo or n
        _lib.check(self._lib.tfop_gemm_allreduce_bf16_tn(
            self._h, a.data_ptr(), b.data_ptr(), m, n, k, a.stride(0), b.stride(0), ldo,
            float(scale), ws_off, out_off, max_ctas, _lib.stream_handle(stream)), "gemm_allreduce")
        flat = self.view(out_off, (m - 1) * ldo + n, torch.bfloat16)
        return torch.as_strided(flat, (m, n), (ldo, 1))

    def gemm_all_reduce_nvls(self, a: torch.Tensor, b: torch.Tensor, *, ws_off: int, out_off: int = 0,
                             ldo: Optional[int] = None, scale: float = 1.0, max_ctas: int = 0,
                             a_mn: bool = False, b_mn: bool = False, stream=None) -> torch.Tensor:
        """out = scale * sum over ranks of op(a_r) @ op(b_r)^T, ONE kernel: the cta_group::2
        tcgen05 GEMM stores its partial tiles locally and the NVSwitch sums / broadcasts them as
        they complete (csrc/gemm_allreduce.cu, NVLS flavour). ``a`` is [M, K] (or [K, M] with
        ``a_mn``), ``b`` is [N, K] (or [K, N] with ``b_mn``): (True, True) is the weight gradient
        dY^T X from the stored activations. The result is this rank's view of the NVLS buffer at
        ``out_off``; ``ws_off`` names ``gemm_allreduce_nvls_workspace_bytes`` zeroed bytes of the
        symmetric buffer, tied to this (M, N)."""
        assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
        a = a if a.stride(1) == 1 and a.stride(0) % 8 == 0 else a.contiguous()
        b = b if b.stride(1) == 1 and b.stride(0) % 8 == 0 else b.contiguous()
        (k, m) = a.shape if a_mn else a.shape[::-1]
        (kb, n) = b.shape if b_mn else b.shape[::-1]
        assert k == kb, "contraction lengths differ"
        ldo = ldo or n
        _lib.check(self._lib.tfop_gemm_allreduce_nvls_bf16(
            self._h, a.data_ptr(), b.data_ptr(), m, n, k, a.stride(0), b.stride(0), ldo,
            int(a_mn), int(b_mn), float(scale), ws_off, out_off, max_ctas,
            _lib.stream_handle(stream)), "gemm_allreduce_nvls")
        flat = self.nvls_view((m - 1) * ldo + n, torch.bfloat16, out_off)
        return torch.as_strided(flat, (m, n), (ldo, 1))

    @staticmethod
    def gemm_allreduce_nvls_workspace_bytes(m: int, n: int, world: int) -> int:
        return _lib.load().tfop_gemm_allreduce_nvls_workspace_bytes(m, n, world)

    def set_timeout_ms(self, ms: int) -> None:
        """Spin limit of the flag barriers (default 120 s). Process-wide; applies to collectives
        issued or captured afterwards."""
        _lib.check(self._lib.tfop_comm_set_timeout_ms(int(ms)), "comm_set_timeout_ms")

    def error_word(self) -> torch.Tensor:
        """int32 device tensor aliasing the communicator's error word (0 = healthy): copy it to
        pinned host memory next to a step's loss to learn about a timed-out barrier without a
        synchronising call."""
        if "err" not in self._views:
            p = self._lib.tfop_comm_error_ptr(self._h)
            self._views["err"] = tensor_from_ptr(p, 4, self.device, owner=self).view(torch.int32)
        return self._views["err"]

    def raise_if_failed(self) -> None:
        if self.check_error():
            raise CommError(f"rank {self.rank}: a collective timed out waiting for a peer")

    def check_error(self) -> bool:
        flag = C.c_int(0)
        _lib.check(self._lib.tfop_comm_check_error(self._h, C.byref(flag)), "comm_check_error")
        return bool(flag.value)

    def close(self, exit_barrier: Optional[Callable[[], None]] = None) -> bool:
        """Release the buffers. With peers, their kernels may still be reading this rank's memory
        over NVLink, so the ranks must first agree that they are all done: ``exit_barrier`` (or
        the one the communicator was formed with - the launcher's rendezvous store or the
        torch.distributed group) runs after a device synchronise, and only then is the memory
        freed. Without any barrier the buffers stay mapped until the process exits (returns
        False) - never freed under a peer."""
        if not self._h:
            return True
        torch.cuda.synchronize(self.device)
        barrier = exit_barrier or getattr(self, "_exit_barrier", None)
        if self.world > 1 and not self.emulated:
            if barrier is None:
                return False
            barrier()
            self._lib.tfop_comm_quiesce(self._h)
        self._views.clear()
        if getattr(self, "nvls_bytes", 0):
            self._lib.tfop_nvls_destroy(self._h)
        _lib.check(self._lib.tfop_comm_destroy(self._h), "comm_destroy")
        self._h = None
        return True


def all_reduce_emulated(comms: Sequence[SymmetricComm], n: int, *,
                        inps: Optional[Sequence[torch.Tensor]] = None,
                        outs: Optional[Sequence[torch.Tensor]] = None, sym_off: int = 0,
                        wire_dtype: torch.dtype = torch.bfloat16, scale: float = 1.0,
                        algo: int = AUTO, max_blocks: int = 0, stream=None) -> None:
    lib = _lib.load()
    world = len(comms)
    harr = (C.c_void_p * world)(*[c._h for c in comms])
    iarr = (C.c_void_p * world)(*[t.data_ptr() for t in inps]) if inps is not None else None
    oarr = (C.c_void_p * world)(*[t.data_ptr() for t in outs]) if outs is not None else None
    _lib.check(lib.tfop_allreduce_emulated(
        harr, world, iarr, _lib.dtype_tag(inps[0].dtype) if inps is not None else 0, oarr,
        _lib.dtype_tag(outs[0].dtype) if outs is not None else 0, _lib.dtype_tag(wire_dtype), n,
        sym_off, float(scale), algo, max_blocks, _lib.stream_handle(stream)),
        "allreduce_emulated")


def gemm_all_reduce_emulated(comms: Sequence[SymmetricComm], a_list, b_list, *, ws_off: int,
                             out_off: int, scale: float = 1.0, stream=None):
    """All ranks of an emulated group in one cooperative launch (single-GPU tests)."""
    lib = _lib.load()
    world = len(comms)
    m, k = a_list[0].shape
    n = b_list[0].shape[0]
    harr = (C.c_void_p * world)(*[c._h for c in comms])
    aarr = (C.c_void_p * world)(*[t.data_ptr() for t in a_list])
    barr = (C.c_void_p * world)(*[t.data_ptr() for t in b_list])
    _lib.check(lib.tfop_gemm_allreduce_emulated(
        harr, world, aarr, barr, m, n, k, a_list[0].stride(0), b_list[0].stride(0), n,
        float(scale), ws_off, out_off, _lib.stream_handle(stream)), "gemm_allreduce_emulated")
    return [c.view(out_off, m * n, torch.bfloat16).view(m, n) for c in comms]
