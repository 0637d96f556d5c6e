"""CPU, world_size 2 over gloo: the data-parallel gradient exchange (convnets_b200/dataparallel.py).

DP oracle: run the reference graph independently on each shard with the same
weights (per-replica BatchNorm statistics), SUM the gradient sets (the loss is
CrossEntropyLoss(reduction='sum'), reference src/basemodel.py:46), apply one Adam step.  The
host-side logic (bucketing, hooks, end-of-backward wait, rank-0 broadcast) is device-agnostic, so
it is exercised here with the un-converted torch.nn mirror on CPU; the CUDA kernels are not
involved (they have no CPU path).
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q, bucket_bytes):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from convnets_b200 import zoo as M
        from convnets_b200.dataparallel import GradSync, sum_scalar
        from oracle import fill, zoo as Z
        torch.manual_seed(100 + rank)  # ranks start from DIFFERENT weights: broadcast must fix it
        s = M.Settings(kind="18", input_size=(3, 32, 32), num_classes=10, dropout_rate=0.0)
        s.device.set_device("cpu")
        model = M.ResNet(s)
        sync = GradSync(model, bucket_bytes=bucket_bytes)
        sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
        model.init_optimizer()
        x, y = fill.batch(8, 32, seed=21)
        xs, ys = x[rank * 4:(rank + 1) * 4], y[rank * 4:(rank + 1) * 4]

        class Loader(list):
            dataset = range(4)

        loss_per_img, _ = model.train_model(Loader([(xs, ys)]))   # the unmodified loop
        total_loss = sum_scalar(loss_per_img * 4)

        # DP oracle on rank 0's view of the weights: per-shard graphs, summed gradients, one Adam
        grads, loss_sum = None, 0.0
        for r in range(world):
            _, l, g, _ = Z.loss_and_grads("ResNet", sd0, x[r * 4:(r + 1) * 4], y[r * 4:(r + 1) * 4], "18")
            loss_sum += float(l)
            grads = g if grads is None else {k: grads[k] + g[k] for k in g}
        want = Z.adam_step(sd0, grads, {}, lr=s.learning_rate, weight_decay=s.weight_decay)
        err = {}
        for k, p in model.named_parameters():
            err[k] = (float((p.grad - grads[k]).norm() / (grads[k].norm() + 1e-12)),
                      float((p.detach() - want[k]).norm() / ((want[k] - sd0[k]).norm() + 1e-12)))
        digest = float(sum(p.detach().double().sum() for p in model.parameters()))
        q.put((rank, max(e[0] for e in err.values()), max(e[1] for e in err.values()),
               abs(total_loss - loss_sum) / abs(loss_sum), digest, len(sync.buckets), sync.launched))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("bucket_bytes", [25 << 20, 1 << 20])
def test_gradsync_matches_dp_oracle(bucket_bytes):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, bucket_bytes)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, g0, u0, l0, d0, nb0, n0), (_, g1, u1, l1, d1, nb1, n1) = res
    assert max(g0, g1) < 1e-4, (g0, g1)        # reduced gradients == sum of shard gradients
    assert max(u0, u1) < 2e-2, (u0, u1)        # Adam update == oracle Adam on the summed gradients
    assert max(l0, l1) < 1e-5
    assert d0 == d1                            # replicas stay bit-identical
    assert nb0 == nb1 and n0 == n1 == nb0      # one all-reduce per bucket per step
    if bucket_bytes == 1 << 20:
        assert nb0 > 5


def _worker_slots(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from convnets_b200.dataparallel import GradSync
        torch.manual_seed(7 + rank)
        net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.BatchNorm2d(8), torch.nn.ReLU(),
                                  torch.nn.Conv2d(8, 16, 3, padding=1), torch.nn.AdaptiveAvgPool2d(1),
                                  torch.nn.Flatten(), torch.nn.Linear(16, 10))
        net = net.to(memory_format=torch.channels_last)      # conv filters KRSC, as convert() leaves them
        sync = GradSync(net, bucket_bytes=2048)
        opt = torch.optim.Adam(net.parameters(), lr=1e-2)
        g = torch.Generator().manual_seed(50 + rank)
        ptrs, ok = None, True
        for _ in range(3):
            # StepGraph's step: gradients dropped (not zeroed) before every backward
            opt.zero_grad(set_to_none=True)
            x = torch.randn(4, 3, 8, 8, generator=g)
            y = torch.randint(0, 10, (4,), generator=g)
            torch.nn.functional.cross_entropy(net(x), y, reduction="sum").backward()
            cur = []
            for p in net.parameters():
                bi, off = sync._slot[p]
                slot = sync._view(sync.buckets[bi], p, off)
                # what optim.FusedAdam's item table needs: the gradient IS the bucket slot, at a fixed
                # address, dense in the parameter's own physical order
                ok &= p.grad.data_ptr() == slot.data_ptr() and p.grad.stride() == p.stride()
                cur.append(p.grad.data_ptr())
            ok &= ptrs is None or ptrs == cur
            ptrs = cur
            opt.step()
        digest = float(sum(p.detach().double().sum() for p in net.parameters()))
        q.put((rank, bool(ok), digest, len(sync.buckets), sync.launched))
    finally:
        dist.destroy_process_group()


def test_gradients_live_in_fixed_bucket_slots_across_steps():
    """The multi-tensor Adam caches one table of (parameter, gradient) addresses per group and a CUDA
    graph replays it: with GradSync every `.grad` must be the same bucket slot on every step even
    though the step drops the gradients (set_to_none) before each backward."""
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_slots, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, ok0, d0, nb0, n0), (_, ok1, d1, nb1, n1) = res
    assert ok0 and ok1
    assert d0 == d1                              # replicas bit-identical after 3 steps
    assert nb0 == nb1 > 1 and n0 == n1 == 3 * nb0
