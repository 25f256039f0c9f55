"""CPU tests of the training driver's host logic (train.py:41-67 drop-in, ModelEMA, data-parallel plumbing over gloo).
The sm_100a kernels are not involved: a small stock torch module stands in for the model."""
import math
import os

import pytest
import torch
from torch import nn

from image_super_resolution_b200 import train as T


def test_model_ema_matches_reference_formula():
    """utils/models.py:31-40: v = d*v + (1-d)*m with d = decay*(1-exp(-n/tau)), over the whole state_dict."""
    torch.manual_seed(0)
    m = nn.Sequential(nn.Conv2d(3, 4, 3), nn.BatchNorm2d(4))
    ema = T.ModelEMA(m, decay=0.9999, tau=50)
    ref = {k: v.clone() for k, v in ema.ema.state_dict().items()}
    for step in range(1, 4):
        with torch.no_grad():
            for p in m.parameters():
                p.add_(torch.randn_like(p) * 0.1)
            m[1].running_mean.add_(0.5)
        ema.update(m)
        d = 0.9999 * (1 - math.exp(-step / 50))
        for k, v in m.state_dict().items():
            if v.dtype.is_floating_point:
                ref[k] = ref[k] * d + (1 - d) * v
    for k, v in ema.ema.state_dict().items():
        if v.dtype.is_floating_point:
            assert torch.allclose(v, ref[k], rtol=1e-6, atol=1e-7), k
    assert ema.updates == 3 and not any(p.requires_grad for p in ema.ema.parameters())


class _Board:
    def __init__(self):
        self.rows = []

    def add_scalar(self, tag, v, step):
        self.rows.append((tag, v, step))


def test_train_loop_order_and_return_value_cpu():
    """Same signature / return as train.py:41-67; every step logs `loss` at epoch*len+idx+1 and steps the schedule and EMA."""
    torch.manual_seed(1)
    model = nn.Sequential(nn.Conv2d(3, 3, 3, padding=1))
    data = [(torch.randn(2, 3, 8, 8), torch.randn(2, 3, 8, 8)) for _ in range(6)]
    opt = torch.optim.Adam(model.parameters(), lr=1e-2)
    sched = torch.optim.lr_scheduler.LinearLR(opt, start_factor=1, end_factor=0.1, total_iters=6)
    ema = T.ModelEMA(model, tau=6)
    board = _Board()
    scaler = torch.amp.GradScaler("cpu", enabled=False)
    losses = T.train(model, ema, data, nn.L1Loss(), opt, scaler, sched, 2, board)
    assert len(losses) == 6 and all(isinstance(v, float) for v in losses)
    assert [r[2] for r in board.rows] == [2 * 6 + i + 1 for i in range(6)] and board.rows[0][0] == "loss"
    assert ema.updates == 6 and abs(opt.param_groups[0]["lr"] - 1e-3) < 1e-9
    assert model.training


def test_shard_of_and_errors():
    b = torch.arange(12).reshape(6, 2)
    assert torch.equal(T.shard_of(b, 1, 3), b[2:4])
    with pytest.raises(ValueError):
        T.shard_of(b, 0, 4)


def _ddp_worker(rank, world, port, q):
    import torch.distributed as dist
    from image_super_resolution_b200.training import allreduce_mean_
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    r, w, dev = T.init_distributed("gloo")
    assert (r, w) == (rank, world) and dev.type == "cpu"
    torch.manual_seed(10 + rank)                       # replicas start different, broadcast makes them equal
    model = nn.Conv2d(3, 2, 3, padding=1)
    T.broadcast_parameters(model)
    g = torch.Generator().manual_seed(3)
    x, y = torch.randn(4, 3, 6, 6, generator=g), torch.randn(4, 2, 6, 6, generator=g)
    xs, ys = T.shard_of(x, rank, world), T.shard_of(y, rank, world)
    loss = (model(xs) - ys).abs().mean()
    loss.backward()
    flat = torch.cat([p.grad.flatten() for p in model.parameters()])
    allreduce_mean_(flat)                              # the one collective of the data-parallel step
    if rank == 0:
        q.put((flat.numpy(), [p.detach().numpy() for p in model.parameters()]))
    dist.barrier()
    dist.destroy_process_group()


def test_data_parallel_gradient_allreduce_world2_gloo_equals_global_batch():
    """Two gloo ranks, each on its shard of the global batch, after the mean all-reduce hold the gradient of the
    global-batch loss (mean of equal-sized shard means), from identical (broadcast) weights."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 1000)
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    flat, params = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    torch.manual_seed(10)
    model = nn.Conv2d(3, 2, 3, padding=1)
    for p, v in zip(model.parameters(), params):
        assert torch.equal(p.detach(), torch.from_numpy(v))          # rank 0's init was broadcast
    g = torch.Generator().manual_seed(3)
    x, y = torch.randn(4, 3, 6, 6, generator=g), torch.randn(4, 2, 6, 6, generator=g)
    (model(x) - y).abs().mean().backward()
    want = torch.cat([p.grad.flatten() for p in model.parameters()])
    assert torch.allclose(torch.from_numpy(flat), want, rtol=1e-5, atol=1e-7)


def test_fused_tail_objects_refuse_cpu_and_fall_through():
    """optim.FlatAdam has no CPU path; clip_grad_norm_ / GradScaler hand anything that is not a FlatAdam to torch's own."""
    import pytest
    import torch
    from image_super_resolution_b200.optim import FlatAdam, GradScaler, clip_grad_norm_
    lin = torch.nn.Linear(4, 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        FlatAdam(lin.parameters(), lr=1e-3)
    lin(torch.ones(2, 4)).sum().backward()
    ref = torch.sqrt(sum(p.grad.pow(2).sum() for p in lin.parameters())).item()
    n = clip_grad_norm_(lin.parameters(), 0.5)
    assert abs(n.item() - ref) < 1e-5
    now = torch.sqrt(sum(p.grad.pow(2).sum() for p in lin.parameters())).item()
    assert abs(now - min(ref, 0.5)) < 1e-4
    sc = GradScaler("cpu", enabled=False)
    opt = torch.optim.SGD(lin.parameters(), lr=0.1)
    sc.unscale_(opt); sc.step(opt); sc.update()
