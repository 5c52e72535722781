"""Request partitioning across GPUs: independent utterances, full weight replica per GPU, no
collective on the data path.  One process per GPU; `torch.distributed` carries only control-plane traffic (work
assignment check, result gathering as Python objects, max-over-ranks timing).

The reference serialises requests through a single worker thread (tools/../inference.py:748-799,
`launch_thread_safe_queue`); this module is what lets N such workers (one per GPU) share one request list.
"""
from __future__ import annotations

from typing import Any, Callable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def assign(n_requests: int, world_size: int, costs: Sequence[float] | None = None) -> List[List[int]]:
    """Request indices per rank.  Without costs: round-robin (request i -> rank i % world).  With costs (e.g. the
    expected number of frames): longest-processing-time-first greedy bin packing, ties to the lower rank, and every
    rank's list sorted ascending so result order is deterministic."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    if n_requests < 0:
        raise ValueError("n_requests must be >= 0")
    if costs is None:
        return [list(range(r, n_requests, world_size)) for r in range(world_size)]
    if len(costs) != n_requests:
        raise ValueError("costs must have one entry per request")
    load = [0.0] * world_size
    out: List[List[int]] = [[] for _ in range(world_size)]
    for i in sorted(range(n_requests), key=lambda j: (-float(costs[j]), j)):
        r = min(range(world_size), key=lambda q: (load[q], q))
        out[r].append(i)
        load[r] += float(costs[i])
    return [sorted(x) for x in out]


def _world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def run_partitioned(requests: Sequence[Any], fn: Callable[[Any], Any], costs: Sequence[float] | None = None,
                    gather: bool = True) -> List[Any] | None:
    """Run fn(request) for this rank's share; return all results in request order on every rank (gather=True) or
    only this rank's [(index, result)] list.  Results travel as Python objects over the process group (host side);
    tensors should be moved to the CPU by fn."""
    rank, world = _world()
    mine = assign(len(requests), world, costs)[rank]
    local = [(i, fn(requests[i])) for i in mine]
    if not gather:
        return local
    if world == 1:
        parts = [local]
    else:
        parts = [None] * world
        dist.all_gather_object(parts, local)
    out: List[Any] = [None] * len(requests)
    seen = 0
    for p in parts:
        for i, r in p:
            if out[i] is not None:
                raise RuntimeError(f"request {i} was produced by two ranks")
            out[i] = r
            seen += 1
    if seen != len(requests):
        raise RuntimeError(f"{len(requests) - seen} request(s) were not produced by any rank")
    return out


def max_over_ranks(seconds: float, units: float, device: torch.device | str = "cpu") -> Tuple[float, float]:
    """(max time over ranks, total units over ranks) — the bench contract's aggregate: value = units / time."""
    rank, world = _world()
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    u = torch.tensor([units], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(t[0]), float(u[0])
