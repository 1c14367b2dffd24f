"""N>1 host path on CPU (gloo, world_size 2): sharding by global agent id and the final statistics
gather that bench.py performs over NCCL on the GPU box."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, agents, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from grl_b200.dist import finalize, shard
    first, count = shard(rank, world, agents)
    returns = torch.arange(first, first + count, dtype=torch.float64) * -1.5   # stands for per-agent returns
    t, c, v = finalize([10.0 + rank, 1.0], [count * 100, count * 7], returns, "cpu")
    if rank == 0:
        torch.save({"t": t, "c": c, "v": v}, out)
    dist.destroy_process_group()


def test_shard_and_gather_world2(tmp_path):
    world, agents = 2, 48
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(world, _free_port(), agents, out), nprocs=world, join=True)
    r = torch.load(out)
    assert r["t"] == [11.0, 1.0]                               # max over ranks
    assert r["c"] == [world * agents * 100, world * agents * 7]   # whole-job sums
    assert torch.equal(r["v"], torch.arange(world * agents, dtype=torch.float64) * -1.5)


def test_shards_are_disjoint_and_cover():
    from grl_b200.dist import shard
    ids = []
    for r in range(8):
        f, n = shard(r, 8, 1000)
        ids += list(range(f, f + n))
    assert ids == list(range(8000))
