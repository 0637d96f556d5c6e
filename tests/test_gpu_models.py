"""Full-train-step GPU parity of converted zoo models against the CPU oracle (oracle/zoo.py, which
tests/test_oracle_golden.py pins to the real reference through tests/golden/).

Protocol: identical weights (state_dict copy) and identical seeded inputs; logits,
loss and BatchNorm running statistics must match the fp64 oracle; gradients are compared against
the fp64 oracle next to the noise floor of the reference's own fp32 execution on the same inputs
(whole-network gradients are conditioning-limited, Appendix D), and pass when
ours <= max(tolerance, 2 x noise floor).
  fp32-IO mode : logits rel-L2 <= 1e-4
  bf16 mode    : logits rel-L2 <= 2e-2 against the fp64 oracle run with the library's storage
                 rounding (bf16 feature maps / operand weights, Net.q in oracle/zoo.py) — the same
                 "inputs pre-rounded, only accumulation differs" rule as the per-layer tests;
                 against the un-rounded fp64 oracle the whole-network figure is printed and only
                 bounded loosely (rounding compounds over depth; each LAYER meets 1e-2 on
                 identical inputs in tests/test_gpu_conv.py and test_gpu_ops.py).  2e-2 rather than
                 the per-layer 1e-2: a 1-ulp bf16 rounding flip in one layer (fp32 vs fp64
                 accumulation) is an input perturbation for every later layer.
"""
import pytest
import torch

from oracle import ops as O
from oracle import zoo as Z

CASES = [
    # family, kind, hw, batch
    ("ResNet", "18", 64, 16),
    ("ResNet", "50", 64, 8),
]


def _build(family, kind, hw, dtype):
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, hw, hw), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = getattr(M, family)(s)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    cnb.convert(model, dtype=dtype)
    s.device.move(model)
    return model, sd, s


@pytest.mark.gpu
@pytest.mark.parametrize("family,kind,hw,batch", CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_train_step_parity(family, kind, hw, batch, dtype):
    from convnets_b200 import lib as L
    model, sd, s = _build(family, kind, hw, dtype)
    g = torch.Generator().manual_seed(21)
    x = torch.randn(batch, 3, hw, hw, generator=g)
    y = torch.randint(0, 10, (batch,), generator=g)

    # oracle: fp64 (noise-free) and fp32 (the reference's own precision)
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    lg64, loss64, g64, buf64 = Z.loss_and_grads(family, sd64, x.double(), y, kind)
    lg32, loss32, g32, _ = Z.loss_and_grads(family, sd, x, y, kind)
    lfloor = 0.0
    if dtype == torch.bfloat16:
        exact = lg64
        lg64, loss64, g64, buf64 = Z.loss_and_grads(family, sd64, x.double(), y, kind,
                                                    storage=torch.bfloat16)
        # noise floor of the bf16 computation itself: the same storage-rounded graph accumulated in
        # fp32 instead of fp64 (1-ulp bf16 flips compound through BatchNorm over few samples)
        lg32s, _, g32, _ = Z.loss_and_grads(family, sd, x, y, kind, storage=torch.bfloat16)
        lfloor = O.rel_l2(lg32s, lg64)

    n0 = L.launch_count()
    model.train()
    out = model(x.cuda())
    loss = model.cost_function(out.float(), y.cuda())
    loss.backward()
    torch.cuda.synchronize()
    assert L.launch_count() > n0, "the CUDA library did not run"

    tol = 1e-4 if dtype == torch.float32 else 2e-2
    e_logits = O.rel_l2(out, lg64)
    assert e_logits <= max(tol, 2 * lfloor), f"logits rel-L2 {e_logits} (bf16 noise floor {lfloor})"
    if dtype == torch.bfloat16:
        e_exact = O.rel_l2(out, exact)
        print(f"bf16 logits vs un-rounded fp64 oracle: {e_exact:.3e}")
        # bf16 storage rounding alone (fp64 arithmetic) moves the logits this far from exact:
        e_storage = O.rel_l2(lg64, exact)
        print(f"bf16-storage fp64 oracle vs un-rounded fp64 oracle: {e_storage:.3e}")
        assert e_exact <= max(0.15, 2 * e_storage)
    assert abs(loss.item() - loss64.item()) <= max(tol, 2 * lfloor) * max(1.0, abs(loss64.item()))
    # BatchNorm running statistics and step counters
    new = model.state_dict()
    for k, v in buf64.items():
        if k.endswith("num_batches_tracked"):
            assert int(new[k]) == int(v)
        else:
            assert O.rel_l2(new[k], v) <= max(tol, 1e-4, 2 * lfloor), k
    # gradients vs fp64, next to the reference-fp32 noise floor
    worst = 0.0
    for k, p in model.named_parameters():
        ref = g64[k]
        assert p.grad is not None, k
        ours = O.rel_l2(p.grad, ref)
        floor = O.rel_l2(g32[k], ref)
        gtol = (1e-3 if dtype == torch.float32 else 1e-1)
        small = ref.norm().item() < 1e-6 * max(1.0, sum(v.norm().item() for v in g64.values()))
        if not small:
            assert ours <= max(gtol, 2 * floor), f"{k}: ours {ours:.3e} floor {floor:.3e}"
        worst = max(worst, ours)
    print(f"{family}-{kind} {dtype}: logits {e_logits:.2e} worst grad {worst:.2e}")


@pytest.mark.gpu
def test_state_dict_and_optimizer_see_same_parameters():
    """convert() keeps Parameter objects, names and state_dict keys (reference
    load_state_dict(strict=True), src/basemodel.py:940)."""
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    s = M.Settings(kind="18", input_size=(3, 32, 32), num_classes=10)
    s.device.set_device("cuda:0")
    m = M.ResNet(s)
    keys = list(m.state_dict().keys())
    ids = [id(p) for p in m.parameters()]
    cnb.convert(m)
    assert list(m.state_dict().keys()) == keys
    assert [id(p) for p in m.parameters()] == ids
    m2 = M.ResNet(s)
    m2.load_state_dict(m.state_dict(), strict=True)


@pytest.mark.gpu
def test_train_model_loop_runs_and_learns():
    """The public loop (train_model, reference src/basemodel.py:231-302) over a tiny fixed batch:
    the loss must go down."""
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    torch.manual_seed(0)
    s = M.Settings(kind="18", input_size=(3, 32, 32), num_classes=10, dropout_rate=0.0,
                   learning_rate=1e-3)
    s.device.set_device("cuda:0")
    m = cnb.convert(M.ResNet(s))
    s.device.move(m)
    m.init_optimizer()
    x, y = torch.randn(32, 3, 32, 32), torch.randint(0, 10, (32,))

    class Loader(list):
        dataset = range(32)

    losses = [m.train_model(Loader([(x, y)]))[0] for _ in range(8)]
    assert losses[-1] < losses[0], losses
