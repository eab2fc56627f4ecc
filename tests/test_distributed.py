"""Host-side logic of the batch-sharded data parallelism, world_size 2 over gloo on CPU:
sharding + flat all-reduce of per-shard gradients equals the full-batch gradient."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from msgat_b200.distributed import FlatGradients, shard_bounds
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    X, Y = torch.randn(8, 6), torch.randn(8, 3)
    fg = FlatGradients(model)
    lo, hi = shard_bounds(8, rank, world)
    fg.zero()
    ((model(X[lo:hi]) - Y[lo:hi]) ** 2).mean().backward()
    fg.all_reduce_mean()
    got = torch.cat([fg.flat[o:o + q.numel()] for q, o in zip(fg.params, fg.offsets)]).clone()
    ref_model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    ref_model.load_state_dict(model.state_dict())
    ((ref_model(X) - Y) ** 2).mean().backward()
    want = torch.cat([p.grad.reshape(-1) for p in ref_model.parameters()])
    ret[rank] = float((got - want).abs().max())
    dist.destroy_process_group()


def test_flat_allreduce_matches_full_batch_gradient():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert len(ret) == world and all(v < 1e-6 for v in ret.values()), dict(ret)


def test_shard_covers_batch():
    from msgat_b200.distributed import shard_bounds
    from msgat_b200.synthetic import make_batch, shard
    b = make_batch(8, 2, 1, 5)
    parts = [shard(b, r, 4) for r in range(4)]
    assert torch.equal(torch.cat([p.X for p in parts]), b.X)
    assert [shard_bounds(8, r, 4) for r in range(4)] == [(0, 2), (2, 4), (4, 6), (6, 8)]
