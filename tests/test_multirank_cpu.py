"""world_size-2 gloo test of the N>1 path's host logic: sample sharding + the single
all-reduce of the packed moment buffer reproduce the one-rank result."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, S, B, Cc, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bayesian_torch_b200 as bt
    torch.manual_seed(0)
    logits = torch.randn(S, B, Cc)           # every rank can form any sample (counter-based RNG in the product)
    first, last = bt.shard_samples(S, rank, world)
    p = torch.softmax(logits[first:last].double(), -1)
    buf = torch.cat([p.sum(0).flatten(), (p * p).sum(0).flatten(), (p * torch.log(p + 1e-15)).sum(-1).sum(0)]).float()
    bt.allreduce_moments(buf)
    if rank == 0:
        ret.put(buf.clone())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_moment_allreduce_matches_single_rank():
    S, B, Cc = 7, 5, 11
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, S, B, Cc, ret)) for r in range(2)]
    [p.start() for p in procs]
    buf = ret.get(timeout=120)
    [p.join(timeout=120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    torch.manual_seed(0)
    logits = torch.randn(S, B, Cc)
    mean, var = oracle.mc_moments(logits)
    bc = B * Cc
    got_mean = buf[:bc].view(B, Cc).double() / S
    got_var = buf[bc:2 * bc].view(B, Cc).double() / S - got_mean ** 2
    torch.testing.assert_close(got_mean, mean, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(got_var, var, rtol=1e-4, atol=1e-6)
    p = torch.softmax(logits.double(), -1)
    mi = oracle.predictive_entropy(p) + buf[2 * bc:].double() / S
    torch.testing.assert_close(mi, oracle.mutual_information(p), rtol=1e-4, atol=1e-5)


def _grad_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bayesian_torch_b200 as bt
    torch.manual_seed(0)                                  # replicated parameters
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    x, y = torch.randn(8, 6), torch.randn(8, 3)
    lo, hi = rank * 4, rank * 4 + 4                       # batch shard of this rank
    torch.nn.functional.mse_loss(model(x[lo:hi]), y[lo:hi]).backward()
    n = bt.allreduce_gradients(model, bucket_mb=1e-4)     # tiny buckets: several collectives
    if rank == 0:
        ret.put((n, [p.grad.clone() for p in model.parameters()]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_allreduce_matches_full_batch():
    """Batch sharded over two ranks + bucketed gradient all-reduce (mean) == the full-batch gradient."""
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, ret)) for r in range(2)]
    [p.start() for p in procs]
    n, grads = ret.get(timeout=120)
    [p.join(timeout=120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert n >= 2
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    x, y = torch.randn(8, 6), torch.randn(8, 3)
    torch.nn.functional.mse_loss(model(x), y).backward()
    for g, p in zip(grads, model.parameters()):
        torch.testing.assert_close(g, p.grad, rtol=1e-5, atol=1e-6)


def _bucket_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bayesian_torch_b200 as bt
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3), torch.nn.Linear(3, 3))
    model[3].weight.requires_grad_(True)
    unused = torch.nn.Parameter(torch.zeros(4))           # a parameter that never gets a gradient: its bucket must still flush
    model.register_parameter("unused", unused)
    buckets = bt.GradientBuckets(model, bucket_mb=1e-4)   # tiny buckets: several collectives, issued from inside backward
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    x, y = torch.randn(8, 6), torch.randn(8, 3)
    lo, hi = rank * 4, rank * 4 + 4
    fired_during_backward = []
    for step in range(2):                                 # two steps: the buckets re-arm
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.mse_loss(model(x[lo:hi]), y[lo:hi]).backward()
        fired_during_backward.append(sum(b["fired"] for b in buckets.buckets))
        n = buckets.finish()
        if step == 0:
            grads = [None if p.grad is None else p.grad.clone().numpy() for p in model.parameters()]   # numpy: pickled by value
        opt.step()
    if rank == 0:
        ret.put((n, fired_during_backward, len(buckets.buckets), grads, [p.detach().clone().numpy() for p in model.parameters()]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_buckets_fire_inside_backward_and_match_full_batch():
    """GradientBuckets: all-reduces are issued by the post-accumulate hooks (before backward() returns), the averaged
    gradients the optimizer sees equal the full-batch ones, and two SGD steps track the single-process run."""
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bucket_worker, args=(r, 2, port, ret)) for r in range(2)]
    [p.start() for p in procs]
    n, fired, nb, grads, params = ret.get(timeout=120)
    [p.join(timeout=120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert nb >= 3 and n >= 3
    assert fired[0] >= nb - 1 and fired[1] >= nb - 1      # every complete bucket went out from inside backward
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3), torch.nn.Linear(3, 3))
    model.register_parameter("unused", torch.nn.Parameter(torch.zeros(4)))
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    x, y = torch.randn(8, 6), torch.randn(8, 3)
    for step in range(2):
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.mse_loss(model(x), y).backward()
        if step == 0:
            for g, p in zip(grads, model.parameters()):
                if p.grad is None:
                    assert g is None
                else:
                    torch.testing.assert_close(torch.from_numpy(g), p.grad, rtol=1e-5, atol=1e-6)
        opt.step()
    for a, p in zip(params, model.parameters()):
        torch.testing.assert_close(torch.from_numpy(a), p.detach(), rtol=1e-5, atol=1e-6)
