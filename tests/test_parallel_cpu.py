"""World-size-2 gloo tests (CPU) of the data-parallel host logic: contiguous clip sharding, the
sum-all-reduce + 1/W gradient average that parallel.allreduce_gradients / Adam.grad_scale implement, the BatchNorm
partial-sum exchange that makes N ranks equal one process on the concatenated batch, and rank-0 broadcast."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from maavss_b200 import nn_ops, parallel
    r, w = parallel.init(backend="gloo", sync_bn=True)
    assert (r, w) == (rank, world) and parallel.world_size() == world and nn_ops.SYNC_BN_GROUP is not None
    torch.manual_seed(0)
    B, C = 6, 5
    x = torch.randn(B, C, 4, 3, dtype=torch.float64)
    wgt = torch.randn(C, dtype=torch.float64)
    lo, hi = parallel.shard_bounds(B, rank, world)
    xs = x[lo:hi]
    # ---- BatchNorm statistics from all-reduced partial sums == statistics of the whole batch
    partial = torch.stack([xs.sum(dim=(0, 2, 3)), (xs * xs).sum(dim=(0, 2, 3))], 1)        # (C, 2) like bn_reduce
    nn_ops._allreduce_partial(partial)
    count = B * 12
    mean = partial[:, 0] / count
    var = partial[:, 1] / count - mean * mean
    ok_bn = torch.allclose(mean, x.mean(dim=(0, 2, 3))) and torch.allclose(var, x.var(dim=(0, 2, 3), unbiased=False))
    # ---- gradient of a mean loss: local grads (local mean) summed over ranks, times 1/W, == global-batch gradient
    class _Opt:
        pass
    opt = _Opt()
    opt.flat = _Opt()
    p = wgt.clone().requires_grad_()
    ((xs * p.view(1, C, 1, 1)) ** 2).mean().backward()
    opt.flat.grad = p.grad.clone()
    parallel.allreduce_gradients(opt)
    g_dp = opt.flat.grad * opt.grad_scale
    pf = wgt.clone().requires_grad_()
    ((x * pf.view(1, C, 1, 1)) ** 2).mean().backward()
    ok_grad = torch.allclose(g_dp, pf.grad) and opt.grad_scale == 1.0 / world
    # ---- broadcast + max all-reduce
    m = torch.nn.Linear(3, 2)
    with torch.no_grad():
        m.weight.fill_(float(rank + 1))
    parallel.broadcast_module(m)
    ok_bc = bool((m.weight == 1.0).all())
    mx = parallel.allreduce_max_(torch.tensor([float(rank)]))
    ok_max = float(mx) == world - 1
    q.put((rank, ok_bn, ok_grad, ok_bc, ok_max))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds():
    from maavss_b200.parallel import shard_bounds
    for n in (0, 1, 7, 16, 256):
        for w in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(120)
def test_two_rank_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=100) for _ in range(2)]
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    for r in res:
        assert all(r[1:]), r
