"""world_size-2 run of the multi-GPU host logic on CPU (gloo): the path shards with no data-path
collective — every rank plans the same length-aware shards with zsb_plan_shards(), takes
its own, decodes them (here with the oracle standing in for the GPU, this is a host-logic test), and
only the barrier / max-over-ranks timing / result bookkeeping of bench.py use torch.distributed."""
import os
import socket
import sys
import zlib

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, HERE)
    import gen
    import orclib
    import zsblib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = zsblib.load_product()                       # host-side planner only; no GPU call
    chk = orclib.ref() or orclib.orc()
    sizes = [1000 + 977 * (i % 13) + 40000 * (i % 5 == 0) for i in range(60)]
    plains = [gen.mixed_plain(sz, 300 + i) for i, sz in enumerate(sizes)]
    members = [gen.deflate(p, 6, -15) for p in plains]
    cost = [len(m) + len(p) for m, p in zip(members, plains)]
    shard_of, load = zsblib.plan_shards(lib, cost, world)
    mine = [i for i in range(len(members)) if shard_of[i] == rank]
    out_bytes, crc = 0, 0
    for i in mine:
        r, out = chk.oneshot(members[i], len(plains[i]), -15)
        assert r.ret == 1 and out == plains[i]
        out_bytes += len(out)
        crc ^= zlib.crc32(out)
    dist.barrier()
    t = torch.tensor([float(rank + 1)])                # stands for this rank's step time
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot = torch.tensor([out_bytes, len(mine), load[rank]], dtype=torch.int64)
    gathered = [torch.zeros_like(tot) for _ in range(world)]
    dist.all_gather(gathered, tot)
    x = torch.tensor([crc], dtype=torch.int64)
    xs = [torch.zeros_like(x) for _ in range(world)]
    dist.all_gather(xs, x)
    if rank == 0:
        want = 0
        for p in plains:
            want ^= zlib.crc32(p)
        got = 0
        for v in xs:
            got ^= int(v.item())
        q.put({"max_time": float(t.item()), "bytes": sum(int(g[0]) for g in gathered),
               "members": sum(int(g[1]) for g in gathered), "loads": [int(g[2]) for g in gathered],
               "want_bytes": sum(len(p) for p in plains), "crc_ok": got == want, "n": len(members),
               "total_cost": sum(cost)})
    dist.destroy_process_group()


def test_two_ranks_shard_without_collective():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    res = q.get()
    assert res["max_time"] == 2.0                      # max over ranks
    assert res["members"] == res["n"] and res["bytes"] == res["want_bytes"] and res["crc_ok"]
    assert sum(res["loads"]) == res["total_cost"]
    assert max(res["loads"]) <= 0.55 * res["total_cost"]        # longest-processing-time first balances
