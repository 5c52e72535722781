"""Request partitioning: host logic + a world_size-2 gloo run on CPU."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fish_speech_b200 import partition as P


def test_assign_round_robin_and_lpt():
    assert P.assign(5, 2) == [[0, 2, 4], [1, 3]]
    assert P.assign(0, 3) == [[], [], []]
    assert P.assign(2, 4) == [[0], [1], [], []]
    a = P.assign(6, 2, costs=[10, 1, 1, 1, 1, 6])
    assert sorted(sum(a, [])) == list(range(6))
    loads = [sum([10, 1, 1, 1, 1, 6][i] for i in r) for r in a]
    assert max(loads) == 10 and min(loads) == 10
    with pytest.raises(ValueError):
        P.assign(3, 0)
    with pytest.raises(ValueError):
        P.assign(3, 2, costs=[1.0])


def test_single_process_run():
    out = P.run_partitioned(list(range(7)), lambda r: r * r)
    assert out == [r * r for r in range(7)]
    assert P.max_over_ranks(1.5, 10.0) == (1.5, 10.0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        reqs = [torch.arange(4) + 10 * i for i in range(5)]
        calls = []

        def fn(x):
            calls.append(int(x[0]) // 10)
            return (x * 2).sum().item()

        out = P.run_partitioned(reqs, fn)
        t, u = P.max_over_ranks(1.0 + rank, float(len(calls)))
        q.put((rank, calls, out, t, u))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = [float(((torch.arange(4) + 10 * i) * 2).sum()) for i in range(5)]
    assert res[0][1] == [0, 2, 4] and res[1][1] == [1, 3]         # disjoint shares, no request run twice
    assert res[0][2] == expect and res[1][2] == expect           # every rank sees all results in request order
    assert res[0][3] == 2.0 and res[1][3] == 2.0                 # max over ranks
    assert res[0][4] == 5.0 and res[1][4] == 5.0                 # units summed over ranks
