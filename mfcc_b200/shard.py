"""Multi-GPU sharding of the MFCC path: one process per GPU, clips partitioned, no
collective on the data path.

Every clip is an independent stream after reset and a frame depends only on samples
``170k-1 .. 170k+511`` of its own clip (/root/reference/mfcc/core/preemph.py:18-28,
mfcc/core/frame.py:117-148), so ranks never exchange samples or cepstra while
computing.  ``torch.distributed`` is used for the barriers, for reductions of counters
and timings, and for an optional gather of the results to one rank.  The same code
runs over NCCL (GPUs) and gloo (the CPU tests of the host logic).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) without one."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_clips(n_clips: int, rank: int, world_size: int, layout: str = "block") -> range:
    """Global clip indices owned by `rank`.

    "block": contiguous blocks whose sizes differ by at most one (rows of one host
    buffer stay contiguous per rank); "cyclic": rank r owns r, r+G, r+2G, ...
    (clips ``r::G``)."""
    if not (0 <= rank < world_size):
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    if n_clips < 0:
        raise ValueError("n_clips must be >= 0")
    if layout == "cyclic":
        return range(rank, n_clips, world_size)
    if layout != "block":
        raise ValueError("layout must be 'block' or 'cyclic'")
    base, extra = divmod(n_clips, world_size)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def stream_owner(stream_id: int, world_size: int) -> int:
    """Live streams are pinned to a GPU for their lifetime; their state never migrates."""
    return stream_id % world_size


def split_long_clip(n_samples: int, parts: int, nfft: int = 512, hop: int = 170):
    """Cut ONE long recording into up to `parts` frame-aligned pieces that can be
    processed as independent rows (on one GPU or on several).

    Returns [(f0, f1, s0, s1, skip)]: the piece is samples [s0, s1) of the recording and
    yields the recording's frames [f0, f1) as its own frames [skip, skip + f1 - f0).
    A piece with f0 > 0 starts one hop early (skip = 1): its frame 0 sees a wrong
    pre-emphasis predecessor (the stream-start zero, preemph.py:18) and is dropped, every
    later frame starts at piece sample >= hop and so has its true predecessor.  The
    overlap (nfft samples per cut) is input data every rank can read: nothing is
    exchanged."""
    K = 0 if n_samples < nfft else (n_samples - nfft) // hop + 1
    out = []
    for i in range(parts):
        f0 = K * i // parts
        f1 = K * (i + 1) // parts
        if f1 <= f0:
            continue
        skip = 1 if f0 > 0 else 0
        out.append((f0, f1, hop * (f0 - skip), hop * (f1 - 1) + nfft, skip))
    return out


def bind_to_gpu_numa(device) -> list:
    """Pin this process to the CPU cores NVML reports as local to `device` (its NUMA
    node), so that pinned staging buffers allocated afterwards are first-touched on the
    memory next to the GPU's PCIe root port.  With one process per GPU on a two-socket box
    this keeps host->device copies off the inter-socket link.  Returns the cores used, or
    [] when NVML or the affinity call is unavailable (nothing is changed then)."""
    import os
    try:
        import pynvml
        dev = torch.device(device)
        pr = torch.cuda.get_device_properties(dev)
        pynvml.nvmlInit()
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cores = [64 * w + b for w in range(words) for b in range(64) if (int(mask[w]) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cores = sorted(c for c in cores if c in allowed)
        if not cores:
            return []
        os.sched_setaffinity(0, cores)
        return cores
    except Exception:       # noqa: BLE001 - an optimisation, never a requirement
        return []


def barrier(device=None):
    r, w = world()
    if w > 1:
        dist.barrier()
    if device is not None and torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def reduce_max(value: float, device="cpu") -> float:
    """Max over ranks (timings are reported as the slowest rank's)."""
    t = torch.tensor([value], dtype=torch.float64, device=device)
    if world()[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(value: int, device="cpu") -> int:
    t = torch.tensor([value], dtype=torch.int64, device=device)
    if world()[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return int(t.item())


def gather_cepstra(local: torch.Tensor, n_clips: int, layout: str = "block", dst: int = 0):
    """Assemble the per-rank results ``[n_local, K, C]`` into ``[n_clips, K, C]`` on rank
    `dst` (None elsewhere).  Outside any timed region: the hot path has no collective."""
    r, w = world()
    if w == 1:
        return local
    shapes = [len(shard_clips(n_clips, i, w, layout)) for i in range(w)]
    rows = max(shapes)                       # gather wants equal sizes: pad to the largest shard
    tail = tuple(local.shape[1:])
    send = torch.zeros((rows,) + tail, dtype=local.dtype, device=local.device)
    send[:local.shape[0]] = local
    # int16 is not a collective dtype for NCCL or gloo: move the rows as bytes
    wire = send.reshape(-1).view(torch.uint8)
    raw = [torch.empty_like(wire) for _ in range(w)] if r == dst else None
    dist.gather(wire, raw, dst=dst)
    bufs = [b.view(local.dtype).reshape(send.shape) for b in raw] if r == dst else None
    if r != dst:
        return None
    out = torch.empty((n_clips,) + tail, dtype=local.dtype, device=local.device)
    for i, b in enumerate(bufs):
        idx = shard_clips(n_clips, i, w, layout)
        if len(idx):
            out[idx.start:idx.stop:idx.step] = b[:shapes[i]]
    return out
