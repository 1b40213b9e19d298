"""CPU, world_size 2, gloo: the bucketed gradient all-reduce that the multi-GPU path uses (host logic only)."""
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


def _toy(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 32), torch.nn.GELU(),
                               torch.nn.Linear(32, 4), torch.nn.Linear(4, 4))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import proteinbert_b200 as pb
    model = _toy(seed=rank)                       # ranks start different: broadcast must fix that
    unused = torch.nn.Parameter(torch.ones(3))    # a parameter that gets no gradient this step
    params = list(model.parameters()) + [unused]
    red = pb.GradAllReducer(params, bucket_bytes=512)
    assert len(red.buckets) > 2                   # several buckets -> several overlapped collectives
    red.broadcast_parameters(0)
    ref = _toy(seed=0)
    for a, b in zip(model.parameters(), ref.parameters()):
        assert torch.equal(a, b)
    torch.manual_seed(100 + rank)
    x = torch.randn(8, 16)
    for step in range(2):
        for p in params:
            p.grad = None
        red.prepare()
        model(x).square().mean().backward()
        red.finish()
    # expected: mean over ranks of the per-rank gradients
    exp = [torch.zeros_like(p) for p in ref.parameters()]
    for r in range(world):
        torch.manual_seed(100 + r)
        xr = torch.randn(8, 16)
        ref.zero_grad()
        ref(xr).square().mean().backward()
        for e, p in zip(exp, ref.parameters()):
            e += p.grad / world
    ok = all(torch.allclose(p.grad, e, atol=1e-6) for p, e in zip(model.parameters(), exp))
    ok = ok and unused.grad is not None and float(unused.grad.abs().sum()) == 0.0
    out[rank] = ok
    dist.destroy_process_group()


def test_bucketed_allreduce_averages_gradients_world2():
    world = 2
    port = _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}


def test_single_process_reducer_is_a_noop():
    import proteinbert_b200 as pb
    model = _toy(0)
    red = pb.GradAllReducer(model.parameters())
    red.prepare()
    model(torch.randn(2, 16)).sum().backward()
    g = [p.grad.clone() for p in model.parameters()]
    red.finish()
    assert all(torch.equal(a, p.grad) for a, p in zip(g, model.parameters()))
