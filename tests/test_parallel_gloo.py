"""CPU, world_size 2 over gloo: the N>1 host logic of the emitter path - shard rule, single-producer ray
scatter / result gather (DS/emitters/nerf_op.py:92-166 semantics) and the gradient all-reduce."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerf_emitter_b200 import parallel


def test_shard_sizes_match_reference_rule():
    # get_target_sizes (nerf_op.py:92-95): divmod, the first `rem` ranks take one more
    assert parallel.shard_sizes(10, 4) == [3, 3, 2, 2]
    assert parallel.shard_sizes(3, 8) == [1, 1, 1, 0, 0, 0, 0, 0]
    assert parallel.shard_sizes(0, 2) == [0, 0]
    for n in (1, 7, 4096, 16777216 + 5):
        for w in (1, 2, 4, 8):
            s = parallel.shard_sizes(n, w)
            assert sum(s) == n and max(s) - min(s) <= 1
            ranges = [parallel.shard_range(n, r, w) for r in range(w)]
            assert ranges[0][0] == 0 and ranges[-1][1] == n
            assert all(ranges[i][1] == ranges[i + 1][0] for i in range(w - 1))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        payload = torch.rand(total, 6, generator=g)                     # [o | v] rows, identical on every rank
        mine = parallel.scatter_rays(payload if rank == 0 else None, total, 6, "cpu")
        a, b = parallel.shard_range(total, rank, world)
        ok = torch.equal(mine, payload[a:b])
        # each rank "renders" its shard: a deterministic per-row function
        rgb = mine[:, :3] * 2 + mine[:, 3:]
        full = parallel.gather_rows(rgb, total)
        if rank == 0:
            ok = ok and torch.equal(full, payload[:, :3] * 2 + payload[:, 3:])
        else:
            ok = ok and full is None
        # the one collective of the backward: sum of the flat gradient buffer, unscale folded in
        flat = torch.full((1000,), float(rank + 1))
        parallel.allreduce_grads(flat, scale=0.5)
        ok = ok and torch.equal(flat, torch.full((1000,), 0.5 * sum(range(1, world + 1))))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 64, 1])
def test_world2_scatter_gather_allreduce(total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


def test_single_process_is_identity():
    x = torch.rand(5, 6)
    assert parallel.scatter_rays(x, 5, 6, "cpu") is x
    assert parallel.gather_rows(x, 5) is x
    f = torch.ones(4)
    assert torch.equal(parallel.allreduce_grads(f, scale=2.0), torch.full((4,), 2.0))
