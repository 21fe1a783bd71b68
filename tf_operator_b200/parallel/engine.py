"""Data-parallel training engine (K3): the replica-side loop of a PyTorchJob.

    engine = DataParallelEngine(model, comm=SymmetricComm|None, lr=..., ...)
    handle = engine.step(images_u8_pinned, labels_pinned)   # H2D + fwd + bwd + comm + opt
    loss = handle.result()                                   # D2H read

One step is: async H2D of the uint8 batch (copy stream, double-buffered
staging) -> fused u8->bf16 normalise -> forward/backward (cuDNN convs + fused
BN kernels) -> per-bucket multi-tensor pack into symmetric memory and in-place
fused P2P all-reduce on a side stream as the bucket's gradients become ready
(overlapping the rest of backward) -> one fused SGD kernel over the flat
buffers. With ``graph=True`` everything after the H2D copy is one CUDA graph.

Reference parity: the training loop of the reference's example workloads
(examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:
106-120: strategy scope, compile, fit) whose gradient exchange is a NCCL
library call; here it is the hand-written kernels of csrc/allreduce.cu.
"""
from __future__ import annotations

import threading
from typing import Callable, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import optim as fopt
from ..ops import preproc
from ..runtime.symm import CommError, SymmetricComm, TWO_SHOT
from .flat import FlatParams


class StepHandle:
    """Result of an asynchronous step; ``result()`` is the device->host read."""

    def __init__(self, host_loss: torch.Tensor, event: torch.cuda.Event,
                 host_err: Optional[torch.Tensor] = None):
        self._host, self._event, self._err = host_loss, event, host_err

    def result(self) -> float:
        self._event.synchronize()
        if self._err is not None and int(self._err.item()) != 0:
            # the gradient exchange of this step gave up on a peer and reduced nothing: the
            # optimizer then stepped on this rank's own gradient only. Stop here and let the
            # operator restart the gang from the last checkpoint.
            raise CommError("gradient all-reduce timed out waiting for a peer replica")
        return float(self._host.item())


class FlatDataParallel:
    """Model-agnostic core: flat parameter buffers, bucketed gradient exchange overlapped with
    backward, one fused optimizer kernel. ``train_step(closure)`` runs one eager step for any
    model (BERT, wide-and-deep, ...); DataParallelEngine adds the image pipeline + CUDA graph.
    """

    def __init__(self, model: nn.Module, *, comm: Optional[SymmetricComm] = None,
                 optimizer: str = "sgd", lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 1e-4, nesterov: bool = False, betas=(0.9, 0.999),
                 eps: float = 1e-8, bucket_bytes: int = 16 << 20, overlap: bool = True,
                 comm_blocks: int = 32, device: Optional[torch.device] = None,
                 channels_last: bool = False, use_nvls: bool = False, fused_wgrad: bool = False):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.comm = comm
        self.world = comm.world if comm is not None else 1
        self.overlap = overlap and self.world > 1
        self.comm_blocks = comm_blocks
        self.optimizer = optimizer
        self.nesterov = nesterov
        self.use_nvls = use_nvls and comm is not None and getattr(comm, "nvls_bytes", 0) > 0
        model.to(self.device)
        if channels_last:
            model.to(memory_format=torch.channels_last)
        self.model = model
        total = sum((p.numel() + 7) // 8 * 8 for p in model.parameters() if p.requires_grad)
        grad_buf = None
        if comm is not None:
            if self.use_nvls:
                assert comm.nvls_bytes >= total * 2, "NVLS buffer too small for the gradient bucket"
                grad_buf = comm.nvls_view(total, torch.bfloat16)
            else:
                assert comm.nbytes >= total * 2, "symmetric buffer too small for the gradient bucket"
                grad_buf = comm.view(0, total, torch.bfloat16)
        # fused_wgrad: the weight gradient of every once-used TcLinear is computed and summed over
        # the ranks by ONE kernel (GEMM -> in-switch reduction, csrc/gemm_allreduce.cu) straight into
        # its slot of the multicast-bound gradient buffer; those slots belong to no bucket
        fused = []
        if fused_wgrad and self.use_nvls and self.world > 1:
            from ..ops.gemm import TcLinear
            owners = {}                       # a weight tied to another module is used twice
            for m in model.modules():
                for p in m.parameters(recurse=False):
                    owners[id(p)] = owners.get(id(p), 0) + 1
            for m in model.modules():
                if isinstance(m, TcLinear) and owners.get(id(m.weight), 0) == 1 and \
                        m.weight.requires_grad and m.out_features >= 256 and \
                        m.out_features % 8 == 0 and m.in_features % 8 == 0:
                    fused.append(m.weight)
        self.flat = FlatParams(model, self.device, grad_buffer=grad_buf, bucket_bytes=bucket_bytes,
                               fused=fused)
        self.fused_wgrad = bool(fused)
        if fused:
            ws = 0
            for s in self.flat.slots:
                if s.offset >= self.flat.n_fused:
                    continue
                m, n = s.param.shape
                assert s.param.is_contiguous()
                s.param._tfop_grad_target.fused = (comm, ws, 1.0 / self.world,
                                                   comm._lib.tfop_nvls_local_ptr(comm._h))
                ws += SymmetricComm.gemm_allreduce_nvls_workspace_bytes(m, n, self.world)
            assert ws <= comm.nbytes, "symmetric buffer too small for the fused-wgrad counters"
        self.mom = torch.zeros(self.flat.total, dtype=torch.float32, device=self.device)
        self.mom2 = torch.zeros(self.flat.total, dtype=torch.float32, device=self.device) \
            if optimizer == "adamw" else None
        self.hp = torch.zeros(8, dtype=torch.float32, device=self.device)
        # The H2D copy of the hyper-parameters reads host memory when the copy EXECUTES, and an
        # eager host runs steps ahead of the GPU: each upload gets its own pinned buffer out of
        # a ring, reused only after the copy that read it has completed.
        self._hp_host = torch.tensor([lr, momentum, weight_decay, 1.0, betas[0], betas[1], eps, 0.0],
                                     dtype=torch.float32).pin_memory()
        self._hp_ring = [(torch.empty(8, dtype=torch.float32).pin_memory(), torch.cuda.Event())
                         for _ in range(4)]
        self._hp_next = 0
        self.hp.copy_(self._hp_host)
        self.comm_stream = torch.cuda.Stream(device=self.device, priority=-1)
        # eager engines poll the communicator's error word every this many steps (and at every
        # checkpoint); DataParallelEngine reads it back with each step's loss instead
        self.check_comm_every = 50
        self.kernel_launches_per_step = 0
        self._steps = 0
        self._in_backward = False
        self._hook_lock = threading.Lock()
        self._install_hooks()

    # ------------------------------------------------------------------ comm
    def _install_hooks(self):
        if self.world == 1:
            return
        for s in self.flat.slots:
            s.param.register_post_accumulate_grad_hook(self._make_hook())

    def _make_hook(self):
        def hook(p):
            if not self.overlap or not self._in_backward:
                return
            b = self.flat.bucket_of(p)
            if b is None:
                return                        # fused region: already summed by its own kernel
            with self._hook_lock:
                b.pending -= 1
                fire = b.pending == 0
            if fire:
                cur = torch.cuda.current_stream()
                ev = torch.cuda.Event()
                ev.record(cur)
                self.comm_stream.wait_event(ev)
                self._reduce_bucket(b, self.comm_stream)
        return hook

    def _reduce_bucket(self, b, stream):
        # gradients a producer wrote straight into the bucket (GradTarget) need no packing
        todo = [s for s in b.slots if s.param.grad is not None and not self.flat.grad_in_place(s)]
        if todo:
            fopt.multi_pack_bf16([s.param.grad for s in todo], [s.offset for s in todo],
                                 self.flat.grads, 1.0, stream)
        if self.world > 1:
            # in place on the symmetric gradient bucket: sum over ranks * 1/world
            if self.use_nvls:
                # reduced inside the NVSwitch, in place on the multicast-bound gradient bucket
                self.comm.all_reduce_nvls(b.hi - b.lo, sym_off=b.lo * 2, scale=1.0 / self.world,
                                          max_blocks=self.comm_blocks, stream=stream)
            else:
                self.comm.all_reduce(b.hi - b.lo, sym_off=b.lo * 2, wire_dtype=torch.bfloat16,
                                     scale=1.0 / self.world, algo=TWO_SHOT,
                                     max_blocks=self.comm_blocks, stream=stream)

    # ------------------------------------------------------------------ step pieces
    def _begin(self):
        self.flat.step_id += 1
        for p in self.model.parameters():
            p.grad = None
        for b in self.flat.buckets:
            b.pending = len(b.slots)

    def _backward_and_update(self, loss):
        self._in_backward = True
        try:
            loss.backward()
        finally:
            self._in_backward = False
        cur = torch.cuda.current_stream()
        if self.overlap:
            # buckets whose last gradient never arrived (unused parameters) are reduced now
            cur.wait_stream(self.comm_stream)
            for b in self.flat.buckets:
                if b.pending > 0:
                    self._reduce_bucket(b, cur)
        else:
            for b in self.flat.buckets:
                self._reduce_bucket(b, cur)
        if self.optimizer == "adamw":
            fopt.adamw_(self.flat.master, self.mom, self.mom2, self.flat.grads, self.flat.weights,
                        self.flat.n_decay, self.hp, cur)
        else:
            fopt.sgd_momentum_(self.flat.master, self.mom, self.flat.grads, self.flat.weights,
                               self.flat.n_decay, self.hp, self.nesterov, cur)

    def train_step(self, closure) -> torch.Tensor:
        """One eager step: ``closure()`` returns the loss; returns the detached loss tensor."""
        from ..ops import _lib
        l0 = _lib.launches_total()
        if self.optimizer == "adamw":
            self._hp_host[7] = float(self._steps + 1)
            self._upload_hp()
        self._begin()
        loss = closure()
        self._backward_and_update(loss)
        self.kernel_launches_per_step = _lib.launches_total() - l0
        self._steps += 1
        if self.comm is not None and self._steps % self.check_comm_every == 0:
            self.comm.raise_if_failed()          # synchronises: amortised over many steps
        return loss.detach()

    def _upload_hp(self, stream=None):
        buf, ev = self._hp_ring[self._hp_next]
        self._hp_next = (self._hp_next + 1) % len(self._hp_ring)
        ev.synchronize()                      # no-op until the ring wraps around a busy slot
        buf.copy_(self._hp_host)
        st = stream or torch.cuda.current_stream()
        with torch.cuda.stream(st):
            self.hp.copy_(buf, non_blocking=True)
            ev.record(st)

    def set_lr(self, lr: float):
        self._hp_host[0] = lr
        self._upload_hp()

    # ------------------------------------------------------------------ checkpoint / resume
    # The reference operator never touches model state: resume works by
    # convention - stable replica names and a user-mounted path, as in the Keras example
    # (examples/v1/distribution_strategy/keras-API/multi_worker_strategy-with-keras.py:91-110).
    # This is the runtime half of that convention: a replica restarted by the operator under the
    # same name calls load_checkpoint() and continues from the saved step.
    def state_dict(self) -> dict:
        """fp32 master weights, optimizer moments (keyed by parameter name, CPU tensors), module
        buffers, hyper-parameters and the step counter. ``sd["model"]`` loads into the plain
        ``nn.Module`` with ``load_state_dict(..., strict=False)``."""
        torch.cuda.current_stream().synchronize()
        if self.comm is not None:
            self.comm.raise_if_failed()          # never persist weights a dead peer starved
        sd = {"version": 1, "optimizer": self.optimizer, "step": int(self._steps),
              "hp": [float(x) for x in self._hp_host.tolist()],
              "model": self.flat.export_named(self.flat.master),
              "mom": self.flat.export_named(self.mom),
              "buffers": {n: b.detach().to("cpu", copy=True) for n, b in self.model.named_buffers()}}
        if self.mom2 is not None:
            sd["mom2"] = self.flat.export_named(self.mom2)
        return sd

    def load_state_dict(self, sd: dict, strict: bool = True) -> None:
        if sd.get("optimizer", self.optimizer) != self.optimizer:
            raise ValueError(f"checkpoint was written by optimizer {sd['optimizer']!r}, "
                             f"this engine runs {self.optimizer!r}")
        self.flat.import_named(self.flat.master, sd["model"], strict)
        self.flat.refresh_weights()
        self.flat.import_named(self.mom, sd.get("mom", {}), strict)
        if self.mom2 is not None:
            self.flat.import_named(self.mom2, sd.get("mom2", {}), strict)
        with torch.no_grad():
            bufs = dict(self.model.named_buffers())
            for n, t in sd.get("buffers", {}).items():
                if n in bufs:
                    bufs[n].copy_(t.to(bufs[n].device, dtype=bufs[n].dtype))
        self._steps = int(sd.get("step", 0))
        hp = sd.get("hp")
        if hp:
            self._hp_host.copy_(torch.tensor(hp, dtype=torch.float32))
            self.hp.copy_(self._hp_host)
        torch.cuda.current_stream().synchronize()

    def save_checkpoint(self, path: str) -> None:
        """Atomic write (tmp + rename): a replica killed mid-save leaves the previous file. In a
        data-parallel job every rank holds the same state, so rank 0 (chief / master) saves."""
        import os
        tmp = f"{path}.tmp.{os.getpid()}"
        torch.save(self.state_dict(), tmp)
        os.replace(tmp, path)

    def load_checkpoint(self, path: str, strict: bool = True) -> int:
        """Load a checkpoint written by save_checkpoint; returns the step to resume from."""
        sd = torch.load(path, map_location="cpu", weights_only=False)
        self.load_state_dict(sd, strict)
        return self._steps


class DataParallelEngine(FlatDataParallel):
    def __init__(self, model: nn.Module, *, comm: Optional[SymmetricComm] = None,
                 batch_size: int, image_size: int = 224, channels: int = 3,
                 lr: float = 0.1, momentum: float = 0.9, weight_decay: float = 1e-4,
                 nesterov: bool = False, bucket_bytes: int = 16 << 20, graph: bool = True,
                 overlap: bool = True, comm_blocks: int = 32,
                 loss_fn: Optional[Callable] = None, device: Optional[torch.device] = None,
                 label_smoothing: float = 0.0, stem_s2d: bool = False, use_nvls: bool = False):
        super().__init__(model, comm=comm, optimizer="sgd", lr=lr, momentum=momentum,
                         weight_decay=weight_decay, nesterov=nesterov, bucket_bytes=bucket_bytes,
                         overlap=overlap, comm_blocks=comm_blocks, device=device,
                         channels_last=True, use_nvls=use_nvls)
        self.batch_size, self.image_size, self.channels = batch_size, image_size, channels
        self.loss_fn = loss_fn or (lambda out, y: F.cross_entropy(
            out.float(), y, label_smoothing=label_smoothing))

        # static device inputs / outputs (graph-friendly) + double-buffered staging
        shp = (batch_size, image_size, image_size, channels)
        self.x_u8 = torch.zeros(shp, dtype=torch.uint8, device=self.device)
        # the tcgen05 stem kernel tiles the 2x-downsampled image in 8 x 16 pixel boxes: the
        # space-to-depth input form is used for image sides that are multiples of 32, decided
        # here once (other sizes get the plain input form and the stock 7x7 stem)
        self.stem_s2d = stem_s2d and channels == 3 and image_size % 32 == 0
        xshp = (batch_size, image_size // 2 + 3, image_size // 2 + 3, 16) if self.stem_s2d else shp
        self.x_bf16 = torch.zeros(xshp, dtype=torch.bfloat16, device=self.device)
        self.labels = torch.zeros(batch_size, dtype=torch.int64, device=self.device)
        self.loss_dev = torch.zeros((), dtype=torch.float32, device=self.device)
        self._stage_x = [torch.zeros_like(self.x_u8) for _ in range(2)]
        self._stage_y = [torch.zeros_like(self.labels) for _ in range(2)]
        self._stage_ready = [torch.cuda.Event() for _ in range(2)]
        self._stage_free = [torch.cuda.Event() for _ in range(2)]
        self._host_loss = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)]
        self._host_err = [torch.zeros(1, dtype=torch.int32).pin_memory() for _ in range(2)] \
            if comm is not None else None
        self._slot = 0
        self._prefetched = None
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.main_stream = torch.cuda.Stream(device=self.device)
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._use_graph = graph

    def _body(self):
        """fwd + bwd + gradient exchange + optimizer on the current stream."""
        from ..ops import _lib
        l0 = _lib.launches_total()
        if self.stem_s2d:
            x = preproc.stem_space_to_depth(self.x_u8, self.x_bf16)
        else:
            x = preproc.u8_to_bf16_normalize(self.x_u8, self.x_bf16)
        self._begin()
        out = self.model(x)
        loss = self.loss_fn(out, self.labels)
        self.loss_dev.copy_(loss.detach())
        self._backward_and_update(loss)
        self.kernel_launches_per_step = _lib.launches_total() - l0
        if self._steps == 0 and not torch.cuda.is_current_stream_capturing():
            bad = self.flat.check_grad_layouts()
            if bad:
                raise RuntimeError(f"gradient layout differs from parameter layout for {bad[:4]}")

    def _capture(self):
        # warm-up on the main stream (cuDNN autotune, allocator), then capture once
        with torch.cuda.stream(self.main_stream):
            for _ in range(3):
                self._body()
        self.main_stream.synchronize()
        if self.comm is not None:
            torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self.main_stream):
            self._body()
        self._graph = g

    def run_device_step(self):
        """One step on the already-resident device batch (no host traffic)."""
        if self._use_graph:
            if self._graph is None:
                self._capture()
            with torch.cuda.stream(self.main_stream):
                self._graph.replay()
        else:
            with torch.cuda.stream(self.main_stream):
                self._body()
        self._steps += 1

    def prefetch(self, images_u8: torch.Tensor, labels: torch.Tensor):
        """Start the H2D copy of a (pinned) host batch into the free staging slot."""
        slot = self._slot
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self._stage_free[slot])
            self._stage_x[slot].copy_(images_u8, non_blocking=True)
            self._stage_y[slot].copy_(labels, non_blocking=True)
            self._stage_ready[slot].record(self.copy_stream)
        self._prefetched = slot
        self._slot ^= 1

    def step(self, images_u8: Optional[torch.Tensor] = None,
             labels: Optional[torch.Tensor] = None) -> StepHandle:
        """Full step. Pass a host batch, or nothing if ``prefetch`` was already called."""
        if images_u8 is not None:
            self.prefetch(images_u8, labels)
        slot = self._prefetched
        assert slot is not None, "step() needs a batch (argument or prefetch)"
        self._prefetched = None
        ms = self.main_stream
        ms.wait_event(self._stage_ready[slot])
        with torch.cuda.stream(ms):
            self.x_u8.copy_(self._stage_x[slot], non_blocking=True)
            self.labels.copy_(self._stage_y[slot], non_blocking=True)
            self._stage_free[slot].record(ms)
        self.run_device_step()
        host = self._host_loss[self._steps & 1]
        herr = self._host_err[self._steps & 1] if self._host_err is not None else None
        ev = torch.cuda.Event()
        with torch.cuda.stream(ms):
            host.copy_(self.loss_dev, non_blocking=True)
            if herr is not None:
                herr.copy_(self.comm.error_word(), non_blocking=True)
            ev.record(ms)
        return StepHandle(host, ev, herr)

    # ------------------------------------------------------------------ misc
    def set_lr(self, lr: float):
        self._hp_host[0] = lr
        self._upload_hp(self.main_stream)

    def h2d_bytes_per_step(self) -> int:
        return self.x_u8.numel() + self.labels.numel() * 8

    def d2h_bytes_per_step(self) -> int:
        return 4 + (4 if self._host_err is not None else 0)

    def synchronize(self):
        torch.cuda.synchronize(self.device)
