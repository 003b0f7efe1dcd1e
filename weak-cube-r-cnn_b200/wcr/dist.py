"""Image sharding + the one collective of the path.

The hot path is embarrassingly parallel by image: rank r owns a contiguous image
shard and runs every kernel on it with no data-path communication. The only
exchange is the final all-gather of the per-box top-k records (fixed stride,
80 B * k per box), done with ``torch.distributed.all_gather_into_tensor`` (NCCL
over NVLink 5 / NVSwitch on the GPU box; ``gloo`` in the CPU tests). The scoring
kernel writes its records directly into the send buffer (``ScoreWorkspace``'s
``records`` argument), so there is no staging copy.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

RECORD_FLOATS = 20


def shard_range(n_images: int, rank: int, world: int):
    """Contiguous image shard [lo, hi) of rank ``rank``; sizes differ by <= 1."""
    base, rem = divmod(n_images, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def bind_to_gpu_numa(local_rank: int) -> dict:
    """Pin this process (and so its first-touch pinned host buffers) to the CPUs of the NUMA
    node its GPU hangs off: eight ranks that all allocate on node 0 share one socket's memory
    controllers for their H2D copies. Reads sysfs only; a box that does
    not expose the topology (numa_node == -1, one node, a cpuset that excludes the node) is
    left untouched and the reason is reported."""
    import os
    info = {"bound": False}
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        info["pci"] = bdf
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        info["numa_node"] = node
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        info["n_nodes"] = len(nodes)
        allowed = os.sched_getaffinity(0)
        info["cpus_allowed"] = len(allowed)
        if node < 0 or len(nodes) < 2:
            info["why"] = "no NUMA topology exposed for this GPU / single node"
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= allowed
        if not cpus:
            info["why"] = "the process cpuset holds no CPU of the GPU's node"
            return info
        os.sched_setaffinity(0, cpus)
        info["bound"], info["cpus_bound"] = True, len(cpus)
    except (OSError, ValueError, AttributeError, RuntimeError) as e:
        info["why"] = f"{type(e).__name__}: {e}"
    return info


def balanced_assignment(weights, world: int):
    """Blocks -> ranks by weight, longest first (LPT): rank lists of block ids, each ascending.
    Deterministic (ties by block id), so every rank computes the same assignment. Contiguous
    ranges (``shard_range``) leave the slowest rank ~2 % above the mean on the 8192-image job;
    this brings it within a block's granularity of the mean."""
    order = sorted(range(len(weights)), key=lambda b: (-float(weights[b]), b))
    load = [0.0] * world
    out = [[] for _ in range(world)]
    for b in order:
        r = min(range(world), key=lambda i: (load[i], i))
        out[r].append(b)
        load[r] += float(weights[b])
    return [sorted(x) for x in out]


def is_dist() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def gather_counts(n_local: int, device) -> list[int]:
    """Boxes per rank (one tiny all-gather so records can be un-padded)."""
    if not is_dist():
        return [n_local]
    t = torch.tensor([n_local], dtype=torch.int64, device=device)
    out = torch.empty((dist.get_world_size(),), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out, t)
    return [int(x) for x in out.cpu()]


class TopkGather:
    """Pre-allocated send/receive buffers for the top-k all-gather.

    ``send`` (max_boxes, k, 20) is handed to ``ScoreWorkspace(records=...)``;
    ``gather()`` returns the records of every rank, un-padded, in rank order
    (= global image order because shards are contiguous)."""

    def __init__(self, n_local_boxes: int, k: int, device):
        self.k = k
        self.counts = gather_counts(n_local_boxes, device)
        self.max_boxes = max(max(self.counts), 1)
        world = len(self.counts)
        self.send = torch.zeros((self.max_boxes, k, RECORD_FLOATS), dtype=torch.float32, device=device)
        self.recv = torch.empty((world, self.max_boxes, k, RECORD_FLOATS), dtype=torch.float32,
                                device=device)

    def gather_padded(self) -> torch.Tensor:
        if len(self.counts) == 1:
            self.recv[0].copy_(self.send)
        else:
            dist.all_gather_into_tensor(self.recv.view(-1), self.send.view(-1))
        return self.recv

    def gather(self) -> torch.Tensor:
        r = self.gather_padded()
        return torch.cat([r[i, :c] for i, c in enumerate(self.counts)], dim=0)
