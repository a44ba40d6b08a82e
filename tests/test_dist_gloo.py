"""CPU: multi-rank inference plumbing with world_size 2 over gloo (no GPU needed).

The model is replaced by a per-event reduction so that the test exercises exactly the host-side
logic of the N>1 path: cost-balanced contiguous sharding, sub-batch extraction (edge_index
re-indexing), and the ragged all_gather that restores global event order.
"""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from graphnet_b200.sharding import (allreduce_gradients, predict_sharded, shard_ranges,
                                    take_events)
from graphnet_b200.synthetic import make_pulse_batch
from graphnet_b200.data import Data
import oracle


def _fake_model(d):
    # per-event: [sum of x over nodes (F), number of edges whose endpoints are both inside]
    B = d.n_pulses.shape[0]
    s = torch.zeros(B, d.x.shape[1]).index_add_(0, d.batch, d.x)
    ne = torch.zeros(B).index_add_(0, d.batch[d.edge_index[1]], torch.ones(d.edge_index.shape[1]))
    assert int(d.edge_index.min()) >= 0 and int(d.edge_index.max()) < d.x.shape[0]
    assert bool((d.batch[d.edge_index[0]] == d.batch[d.edge_index[1]]).all())
    return torch.cat([s, ne.unsqueeze(1)], 1)


def _make():
    pb = make_pulse_batch(37, 25.0, seed=3)
    ei = oracle.knn_graph(pb.x[:, :3], 4, pb.batch)
    return Data(x=pb.x, batch=pb.batch, n_pulses=pb.n_pulses, edge_index=ei)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    data = _make()
    out = predict_sharded(_fake_model, data)
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges_balance_and_cover():
    n = torch.tensor([100, 1, 1, 1, 1, 96, 2, 2, 50, 50], dtype=torch.int32)
    r = shard_ranges(n, 3, "pulses")
    assert r[0][0] == 0 and r[-1][1] == 10 and all(a[1] == b[0] for a, b in zip(r, r[1:]))
    loads = [int(n[lo:hi].sum()) for lo, hi in r]
    assert max(loads) <= 1.5 * (int(n.sum()) / 3)
    assert shard_ranges(n, 1) == [(0, 10)]
    assert len(shard_ranges(n[:2], 4)) == 4  # more ranks than events: empty tails are fine
    r2 = shard_ranges(torch.tensor([10, 10, 1000], dtype=torch.int32), 2, "pairs")
    assert r2 == [(0, 2), (2, 3)]


def test_take_events_reindexes_edges():
    data = _make()
    sub = take_events(data, 5, 11)
    assert sub.n_pulses.tolist() == data.n_pulses[5:11].tolist()
    assert int(sub.batch.min()) == 0 and int(sub.batch.max()) == 5
    want = oracle.knn_graph(sub.x[:, :3], 4, sub.batch)
    assert torch.equal(sub.edge_index, want)


def test_predict_sharded_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _fake_model(_make())
    assert torch.allclose(outs[0], want) and torch.allclose(outs[1], want)


def _grad_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.ReLU(), torch.nn.Linear(7, 2))
    frozen = torch.nn.Parameter(torch.ones(3), requires_grad=False)
    x = torch.randn(16, 5, generator=torch.Generator().manual_seed(100 + rank))
    net(x).square().mean().backward()
    if rank == 1:
        net[2].bias.grad = None          # a rank that produced no gradient for a tensor
    local = [None if p.grad is None else p.grad.clone() for p in net.parameters()]
    nbytes = allreduce_gradients(list(net.parameters()) + [frozen])
    q.put((rank, nbytes, [None if t is None else t.tolist() for t in local],
           [p.grad.tolist() for p in net.parameters()]))   # plain lists: no shared-memory handles
    dist.barrier()
    dist.destroy_process_group()


def test_allreduce_gradients_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 200) + 211
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, nb0, loc0, red0), (_, nb1, loc1, red1) = res
    red0 = [torch.tensor(t) for t in red0]
    red1 = [torch.tensor(t) for t in red1]
    nparam = sum(t.numel() for t in red0)
    assert nb0 == nb1 == 4 * nparam
    for a, b, r0, r1 in zip(loc0, loc1, red0, red1):
        a = torch.tensor(a)
        b = torch.zeros_like(a) if b is None else torch.tensor(b)
        assert torch.allclose(r0, (a + b) / 2, atol=1e-7) and torch.equal(r0, r1)
