"""GPU inventory and NVLink / NVSwitch topology probe used for replica placement.

The reference leaves placement to kube-scheduler / Volcano; on one 8xB200
box the launcher is the scheduler. Every GPU reaches every peer through NVSwitch at the same
NVLink-5 bandwidth, so any subset is equally good; on boxes without a switch the probe's link
matrix is used to pick the most tightly connected subset. NVLS (multicast) support is probed
through the CUDA driver attribute and only reported, never assumed.
"""
from __future__ import annotations

import itertools
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence


@dataclass
class GpuInfo:
    index: int
    name: str = ""
    uuid: str = ""
    memory_mib: int = 0
    nvlink_peers: Dict[int, int] = field(default_factory=dict)   # peer index -> #links
    nvswitch_links: int = 0


@dataclass
class Topology:
    gpus: List[GpuInfo] = field(default_factory=list)
    has_nvswitch: bool = False
    multicast_supported: Optional[bool] = None
    source: str = "none"

    @property
    def count(self) -> int:
        return len(self.gpus)

    def link_score(self, subset: Sequence[int]) -> int:
        if self.has_nvswitch:
            return len(subset) * (len(subset) - 1)
        by = {g.index: g for g in self.gpus}
        return sum(by[a].nvlink_peers.get(b, 0) for a, b in itertools.permutations(subset, 2))

    def pick(self, n: int, free: Sequence[int]) -> Optional[List[int]]:
        """Choose n GPUs out of ``free`` maximising NVLink connectivity (all-or-nothing)."""
        free = sorted(free)
        if n > len(free):
            return None
        if n == 0:
            return []
        if self.has_nvswitch or n == 1 or len(free) > 12:
            return free[:n]
        return list(max(itertools.combinations(free, n), key=self.link_score))


def _visible_indices(count: int) -> List[int]:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis is None or vis.strip() == "":
        return list(range(count))
    out = []
    for tok in vis.split(","):
        tok = tok.strip()
        if tok.isdigit() and int(tok) < count:
            out.append(int(tok))
    return out


def probe() -> Topology:
    """NVML probe; returns an empty topology on a box without GPUs / NVML."""
    try:
        import pynvml
        pynvml.nvmlInit()
    except Exception:  # noqa: BLE001
        return Topology(source="no-nvml")
    topo = Topology(source="nvml")
    try:
        count = pynvml.nvmlDeviceGetCount()
        handles = {i: pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(count)}
        bus_to_idx = {}
        for i, h in handles.items():
            try:
                bus_to_idx[pynvml.nvmlDeviceGetPciInfo(h).busId] = i
            except Exception:  # noqa: BLE001
                pass
        for i in _visible_indices(count):
            h = handles[i]
            g = GpuInfo(index=i)
            try:
                nm = pynvml.nvmlDeviceGetName(h)
                g.name = nm.decode() if isinstance(nm, bytes) else nm
                uu = pynvml.nvmlDeviceGetUUID(h)
                g.uuid = uu.decode() if isinstance(uu, bytes) else uu
                g.memory_mib = int(pynvml.nvmlDeviceGetMemoryInfo(h).total // (1 << 20))
            except Exception:  # noqa: BLE001
                pass
            for link in range(18):
                try:
                    if pynvml.nvmlDeviceGetNvLinkState(h, link) != 1:
                        continue
                    remote = pynvml.nvmlDeviceGetNvLinkRemotePciInfo(h, link)
                    peer = bus_to_idx.get(remote.busId)
                    if peer is None:
                        g.nvswitch_links += 1       # remote end is not a GPU: a switch
                    else:
                        g.nvlink_peers[peer] = g.nvlink_peers.get(peer, 0) + 1
                except Exception:  # noqa: BLE001
                    break
            topo.gpus.append(g)
        topo.has_nvswitch = any(g.nvswitch_links > 0 for g in topo.gpus)
    finally:
        try:
            pynvml.nvmlShutdown()
        except Exception:  # noqa: BLE001
            pass
    topo.multicast_supported = _probe_multicast()
    return topo


def _probe_multicast() -> Optional[bool]:
    try:
        from cuda import cuda  # cuda-python
        (err,) = cuda.cuInit(0)
        if int(err) != 0:
            return None
        err, dev = cuda.cuDeviceGet(0)
        if int(err) != 0:
            return None
        err, val = cuda.cuDeviceGetAttribute(
            cuda.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return bool(val) if int(err) == 0 else None
    except Exception:  # noqa: BLE001
        return None


def synthetic(n: int, nvswitch: bool = True) -> Topology:
    """A made-up topology for CPU-only tests and launch-latency sweeps."""
    t = Topology(source="synthetic", has_nvswitch=nvswitch)
    for i in range(n):
        t.gpus.append(GpuInfo(index=i, name="synthetic", nvswitch_links=18 if nvswitch else 0))
    return t
