"""Rank plumbing for the two multi-GPU modes of the path: batch-sharded sampling (no collective) and
data-parallel training (one all-reduce of a flat gradient buffer per step).  The reference has no distributed code
(SURVEY.md section 2); one process per GPU, launched by torchrun, backend nccl (gloo in the CPU tests)."""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def rank_world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def bind_to_gpu_numa(device_index: int) -> bool:
    """Pin this process to the CPU cores nearest to its GPU (NVML's ideal CPU affinity), so the pinned host buffers of the sampling path
    (1 GiB in, 2 GiB out per step and rank at the bench shape) are first touched on the GPU's own NUMA node: with eight unbound ranks
    the host-memory traffic crosses sockets.  Only under torchrun with more than one rank; returns False (and changes nothing) when NVML
    or the affinity call is unavailable."""
    try:
        import pynvml

        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device_index)
        bus = "%08x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if not cpus:
            return False
        os.sched_setaffinity(0, cpus)
        return True
    except Exception:
        return False


def init_from_env(backend: str = "nccl") -> Tuple[int, int, int]:
    """torchrun contract: RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* in the environment.  Returns (rank, local_rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend == "nccl":
            torch.cuda.set_device(local)
            bind_to_gpu_numa(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend)
    return rank, local, world


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous balanced split of n units: the first n % world ranks own one extra unit (ragged n allowed, empty shards too)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_host_lists(local: List[torch.Tensor], dst: int = 0):
    """Concatenate per-rank lists of HOST tensors along dim 0 on rank ``dst`` (others get None).  Host-side object gather:
    the sampling data path itself has no collective."""
    rank, world = rank_world()
    if world == 1:
        return local
    gathered = [None] * world if rank == dst else None
    dist.gather_object([t.numpy() for t in local], gathered, dst=dst)
    if rank != dst:
        return None
    out = []
    for i in range(len(local)):
        parts = [torch.from_numpy(g[i]) for g in gathered if len(g) > i and g[i].shape[0] > 0]
        out.append(torch.cat(parts, dim=0))
    return out


def allreduce_gradients(params: Sequence[torch.nn.Parameter], flat: torch.Tensor = None) -> None:
    """SUM all-reduce of every ``p.grad`` across ranks (the loss is already scaled by 1/world).  When every gradient is a view
    into ``flat`` (ddim_audio_b200.Model writes its gradients into one flat fp32 buffer) that buffer is reduced in ONE
    collective; otherwise gradients are packed into a temporary flat buffer and unpacked."""
    rank, world = rank_world()
    if world == 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        return
    if flat is not None:
        lo, hi = flat.data_ptr(), flat.data_ptr() + flat.numel() * flat.element_size()
        if all(lo <= g.data_ptr() < hi and g.is_contiguous() for g in grads):
            dist.all_reduce(flat, op=dist.ReduceOp.SUM)
            return
    buf = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    off = 0
    for g in grads:
        g.copy_(buf[off:off + g.numel()].view_as(g))
        off += g.numel()
