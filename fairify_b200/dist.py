"""Multi-GPU plumbing of the pre-solver stage: one process per GPU, contiguous partition range per rank, one
all-gather of packed per-partition records at the end (torch.distributed; NCCL over NVLink on GPUs, gloo in
the CPU tests).  Partitions are independent, so there is no exchange during compute; the Philox counters use
the GLOBAL partition index, which makes the gathered result identical to a single-GPU run.
(The reference has no parallelism at all: its partition loop is sequential, src/AC/Verify-AC.py:109.)
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n_total, rank, world):
    """Contiguous canonical index range [first, first+count) owned by `rank`; sizes differ by at most one."""
    base, rem = divmod(int(n_total), int(world))
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def record_width(NW):
    """uint32 words per partition in the packed record fb_ibp writes (include/fairify_b200.h, fb_prune_out.record):
    verdict flags | dead-mask words | candidate-mask words | first counterexample sample (-1 = none)."""
    return 2 * NW + 2


def pack_records(verdict, dead, cand, ce=None, out=None):
    """The same record assembled from separate tensors (host-side tests; on the device fb_ibp writes it directly)."""
    n, NW = dead.shape
    if out is None:
        out = torch.empty((n, record_width(NW)), dtype=torch.int32, device=dead.device)
    out[:, 0] = verdict
    out[:, 1:1 + NW] = dead
    out[:, 1 + NW:1 + 2 * NW] = cand
    out[:, 1 + 2 * NW] = -1 if ce is None else ce
    return out


def unpack_records(rec, NW):
    """(verdict uint8, dead words, candidate words, first counterexample sample) views of [n, 2*NW+2] records."""
    return rec[:, 0].to(torch.uint8), rec[:, 1:1 + NW], rec[:, 1 + NW:1 + 2 * NW], rec[:, 1 + 2 * NW]


def allgather_records(local, n_total, group=None):
    """Gather every rank's [count_r, W] records into [n_total, W] in canonical order (ranks may own
    counts differing by one: shorter shards are padded for the collective and trimmed afterwards)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    rank = dist.get_rank(group)
    W = local.shape[1]
    cmax = max(shard_range(n_total, r, world)[1] for r in range(world))
    pad = local
    if local.shape[0] < cmax:
        pad = torch.zeros((cmax, W), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
    gathered = torch.empty((world * cmax, W), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, pad.contiguous(), group=group)
    parts = []
    for r in range(world):
        cnt = shard_range(n_total, r, world)[1]
        parts.append(gathered[r * cmax: r * cmax + cnt])
    return torch.cat(parts, dim=0) if any(shard_range(n_total, r, world)[1] != cmax for r in range(world)) else gathered


def prune_grid_sharded(model, grid, S=1000, seed=0, tol=None, n_total=None, group=None, pa_col=None):
    """Every rank prunes (and, given a protected column, screens) its slice of the grid on its own GPU; returns
    (verdict, dead, cand, first counterexample sample) for ALL partitions on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_total = len(grid) if n_total is None else int(n_total)
    first, count = shard_range(n_total, rank, world)
    boxes = grid.boxes(first, count)
    r = model.prune(boxes, S=S, seed=seed, first_index=first, tol=tol, want=("record",))
    rec = r.t["record"]
    if pa_col is not None:
        # (boxes the bounds kernel certified one-signed cannot hold a counterexample pair: not searched)
        certified = (rec[:, 0] & 1).to(torch.uint8)
        model.partition_screen(boxes, pa_col, S=S, seed=seed, first_index=first, record=rec, skip=certified)
    full = allgather_records(rec, n_total, group)
    return unpack_records(full, model.NW)
