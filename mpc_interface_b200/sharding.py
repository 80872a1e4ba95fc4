"""Data parallelism over independent MPC instances.

Instances never interact (nothing in body.py:236-348 couples two of them), so a
batch is split into contiguous shards, one per rank / GPU, and every rank
assembles its shard with no data-path collective.  ``torch.distributed`` is
used only around the path: a barrier for timing and an optional gather of small
per-shard results (checksums, or q / h) for verification -- NCCL on the GPU box,
gloo in the CPU tests.
"""
import numpy as np


def shard_range(total, rank, world):
    """Contiguous shard [start, stop) of ``total`` instances for ``rank``: sizes
    differ by at most one, earlier ranks take the longer shards."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank {} outside a world of {}".format(rank, world))
    base, extra = divmod(int(total), world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_inputs(values, rank, world):
    """Slices every per-instance array of ``values`` ({name: array[B, ...]}) to
    the rank's shard.  The logical batch is generated from ONE seed and sliced,
    so results do not depend on the number of ranks."""
    total = len(next(iter(values.values())))
    start, stop = shard_range(total, rank, world)
    return {name: array[start:stop] for name, array in values.items()}


def checksums(out):
    """[sum(q), sum(h), sum(P), sum(G)] of a shard as a float64 tensor on the
    shard's device (a cheap fingerprint for the optional gather)."""
    import torch

    return torch.stack([out.q.sum(), out.h.sum(), out.P.sum(), out.G.sum()]).to(torch.float64)


def gather_small(tensor, world):
    """All-gather of a small per-shard tensor (same shape on every rank);
    returns the list ordered by rank.  Not on the data path."""
    import torch
    import torch.distributed as dist

    if world == 1 or not dist.is_initialized():
        return [tensor]
    parts = [torch.zeros_like(tensor) for _ in range(world)]
    dist.all_gather(parts, tensor)
    return parts


def gather_rows(local, total, rank, world):
    """All-gather of ragged row blocks (e.g. q or h of each shard) into the
    [total, ...] array of the logical batch, in instance order."""
    import torch
    import torch.distributed as dist

    if world == 1 or not dist.is_initialized():
        return local
    longest = shard_range(total, 0, world)
    width = longest[1] - longest[0]
    padded = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    parts = [torch.zeros_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded)
    rows = []
    for r, part in enumerate(parts):
        start, stop = shard_range(total, r, world)
        rows.append(part[: stop - start])
    return torch.cat(rows)


def max_over_ranks(value, world, device="cpu"):
    """Max of a host float over the ranks (the timing rule of the bench)."""
    import torch
    import torch.distributed as dist

    if world == 1 or not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_host_near_gpu(pci_domain, pci_bus, pci_device, sysfs="/sys"):
    """One process per GPU moves megabytes per QP over its GPU's PCIe link (the
    host-buffer entry points); with several GPUs the pinned host
    buffers have to live on the NUMA node the GPU hangs off, otherwise all ranks
    write through one socket's memory and the inter-socket link.  Call BEFORE
    allocating pinned memory: restricts this process to the CPUs of the GPU's node
    that it is allowed to run on (first touch then places its pages there) and sets
    the memory policy to prefer that node.  Never raises: returns what it did,
    ``{"numa_node": .., "cpus": .., "mempolicy": ..}`` (node -1 = unknown, nothing done)."""
    import ctypes
    import os

    info = {"numa_node": -1, "cpus": 0, "mempolicy": None}
    try:
        dev = "{:04x}:{:02x}:{:02x}.0".format(pci_domain, pci_bus, pci_device)
        with open(os.path.join(sysfs, "bus/pci/devices", dev, "numa_node")) as f:
            node = int(f.read().strip())
        if node < 0:
            return info
        info["numa_node"] = node
        with open(os.path.join(sysfs, "devices/system/node/node{}/cpulist".format(node))) as f:
            local = _parse_cpulist(f.read()) & os.sched_getaffinity(0)
        if local:
            os.sched_setaffinity(0, local)
            info["cpus"] = len(local)
        # set_mempolicy(MPOL_PREFERRED, {node}): x86-64 syscall 238
        if os.uname().machine == "x86_64" and node < 64:
            mask = ctypes.c_ulong(1 << node)
            libc = ctypes.CDLL(None, use_errno=True)
            info["mempolicy"] = int(libc.syscall(238, 1, ctypes.byref(mask), ctypes.c_ulong(65)))
    except (OSError, ValueError):
        pass
    return info
