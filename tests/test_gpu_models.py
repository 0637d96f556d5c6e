"""Full-train-step GPU parity of converted zoo models against the CPU oracle (oracle/zoo.py, which
tests/test_oracle_golden.py pins to the real reference through tests/golden/).

Protocol: identical weights (state_dict copy) and identical seeded inputs.  Every
bound is `max(tolerance, 2 x measured noise floor)` with tolerance = BASELINE.json's 1e-4 (fp32-IO)
/ 1e-2 (bf16), and every floor is measured in the test and printed next to our error:

  logits / loss / BatchNorm running statistics
      fp32-IO : against the fp64 oracle.
      bf16    : against the fp64 oracle run with the library's storage rounding (bf16 feature maps /
                operand weights, Net.q in oracle/zoo.py) — the same "inputs pre-rounded, only the
                accumulation differs" rule as the per-layer tests.  Floor = the same storage-rounded
                graph accumulated in fp32 (oneDNN and ATen-native convolutions) instead of fp64: a
                1-ulp bf16 rounding flip in one layer is an input perturbation for every later layer.
  gradients (against the fp64 oracle; whole-network gradients are conditioning-limited, Appendix D)
      Floors are OTHER legitimate executions of the same graph against the same fp64 target:
      fp32 oneDNN, fp32 ATen-native, and (fp32-IO) the fp64 oracle on the input perturbed by ~8 fp32
      ulps.  Ours and a floor run are independent draws of the same rounding process, so pairing
      them per parameter is not a test (the worse of two equal draws fails half the time); the
      comparison is quantile against quantile over the parameters of the model: our median, 90th
      percentile and maximum must each be <= max(tolerance, 2 x the floor's same quantile).
"""
import pytest
import torch

from oracle import ops as O
from oracle import zoo as Z

CASES = [
    # family, kind, hw, batch — every family of the zoo (BASELINE.json configs C1..C5 at test size)
    ("ResNet", "18", 64, 16),
    ("ResNet", "50", 64, 8),
    ("VGGNet", "11", 32, 8),
    ("InceptionNetV1", "", 64, 8),
    ("ResNeXt", "50", 64, 8),
    ("SqueezeNet", "1.1", 64, 8),
    ("DenseNet", "121", 64, 8),
    ("MobileNetV1", "", 64, 8),
    ("ShuffleNetV1", "g3", 64, 8),
    ("ShuffleNetV1", "g4", 64, 8),     # 68 / 136 / 272 channels per group: no 16-byte aligned channel slices
    # (at 32 x 32 the last stage is 1 x 1: BatchNorm over 8 values per channel, where two bf16 executions
    # of the reference graph itself differ by O(0.1) on the logits; every layer of that case is held to
    # the oracle on identical inputs by tests/test_gpu_layers_in_network.py)
    ("SEResNet", "50", 64, 8),
    ("SENet", "26", 64, 8),
    ("SKResNet", "26", 64, 8),
    ("SKNet", "50", 64, 8),
]
# BASELINE.json config sizes (224x224), small batch so the fp64 oracle finishes in seconds: every
# layer runs at the real H x W of configs C2 / C3 / C4
CASES_224 = [
    ("ResNet", "50", 224, 4),
    ("MobileNetV1", "", 224, 4),
    ("ShuffleNetV1", "g3", 224, 4),
]


def _build(family, kind, hw, dtype):
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, hw, hw), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[family](s)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    cnb.convert(model, dtype=dtype)
    s.device.move(model)
    return model, sd, s


def _quantiles(v):
    v = sorted(v)
    return v[len(v) // 2], v[min(len(v) - 1, (len(v) * 9) // 10)], v[-1]


def _parity(family, kind, hw, batch, dtype):
    from convnets_b200 import lib as L
    model, sd, s = _build(family, kind, hw, dtype)
    g = torch.Generator().manual_seed(21)
    x = torch.randn(batch, 3, hw, hw, generator=g)
    y = torch.randint(0, 10, (batch,), generator=g)
    bf16 = dtype == torch.bfloat16
    storage = torch.bfloat16 if bf16 else None
    tol = 1e-2 if bf16 else 1e-4

    # target: fp64 oracle (bf16: with the library's storage rounding)
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    lg64, loss64, g64, buf64 = Z.loss_and_grads(family, sd64, x.double(), y, kind, storage=storage)
    # floors: other legitimate executions of the same graph
    floors = []   # list of (logits, grads, buffers)
    lg_a, _, g_a, buf_a = Z.loss_and_grads(family, sd, x, y, kind, storage=storage)       # fp32, oneDNN
    floors.append((lg_a, g_a, buf_a))
    with torch.backends.mkldnn.flags(enabled=False):                                        # fp32, ATen native
        lg_b, _, g_b, buf_b = Z.loss_and_grads(family, sd, x, y, kind, storage=storage)
    floors.append((lg_b, g_b, buf_b))
    if hw <= 64:
        # conditioning of the reference graph itself: ONE pre-activation within a few ulps of zero
        # flips its ReLU mask between two executions and moves everything upstream of it.  The
        # fp64 oracle on the input scaled by (1 + 1e-6 randn) (~8 fp32 ulps; bf16 storage: 2e-3,
        # half a bf16 ulp) measures that spread — two more samples of the floor, which matters for
        # the max over a few hundred parameters (a heavy-tailed statistic estimated from 2 runs).
        for seed in (1, 2):
            gp_ = torch.Generator().manual_seed(seed)
            eps_ = 2e-3 if bf16 else 1e-6
            xp = x.double() * (1 + eps_ * torch.randn(x.shape, generator=gp_, dtype=torch.float64))
            lg_p, _, g_p, buf_p = Z.loss_and_grads(family, sd64, xp, y, kind, storage=storage)
            floors.append((lg_p, g_p, buf_p))
    lfloor = max(O.rel_l2(f[0], lg64) for f in floors[:2])

    n0 = L.launch_count()
    model.train()
    out = model(x.cuda())
    loss = model.cost_function(out.float(), y.cuda())
    loss.backward()
    torch.cuda.synchronize()
    assert L.launch_count() > n0, "the CUDA library did not run"

    tag = f"{family}-{kind} {hw}x{hw} B{batch} {'bf16' if bf16 else 'f32'}"
    e_logits = O.rel_l2(out, lg64)
    print(f"PARITY {tag}: logits ours {e_logits:.2e} floor {lfloor:.2e} bound {max(tol, 2 * lfloor):.2e}")
    assert e_logits <= max(tol, 2 * lfloor), f"logits rel-L2 {e_logits} (floor {lfloor})"
    assert abs(loss.item() - loss64.item()) <= max(tol, 2 * lfloor) * max(1.0, abs(loss64.item()))
    # BatchNorm running statistics and step counters
    new = model.state_dict()
    worst_buf = (0.0, 0.0, "")
    for k, v in buf64.items():
        if k.endswith("num_batches_tracked"):
            assert int(new[k]) == int(v)
            continue
        # a batch mean that is analytically zero (zero-mean input, beta = 0: ShuffleNet's expand
        # convs) has no meaningful relative error: scale by the channel standard deviations instead
        den = v.double().norm().item()
        if k.endswith("running_mean"):
            den = max(den, buf64[k[:-4] + "var"].double().sqrt().norm().item() * 0.1)
        den = max(den, 1e-30)
        err = (new[k].detach().cpu().double() - v.double()).norm().item() / den
        bfloor = max((f[2][k].double() - v.double()).norm().item() / den for f in floors[:2])
        if err > worst_buf[0]:
            worst_buf = (err, bfloor, k)
        assert err <= max(tol, 2 * bfloor, 2 * lfloor), (k, err, bfloor)
    print(f"PARITY {tag}: running stats worst ours {worst_buf[0]:.2e} floor {worst_buf[1]:.2e} ({worst_buf[2]})")
    # gradients: quantile against quantile (see the module docstring)
    gsum = sum(v.norm().item() for v in g64.values() if v is not None)
    ours_all, floor_all, names_all = [], [], []
    for k, p in model.named_parameters():
        ref = g64[k]
        assert p.grad is not None, k
        if ref.norm().item() < 1e-6 * max(1.0, gsum):
            continue   # analytically-zero gradient (ShuffleNet has one): no relative error exists
        names_all.append(k)
        ours_all.append(O.rel_l2(p.grad, ref))
        floor_all.append(max(O.rel_l2(f[1][k], ref) for f in floors))
    oq, fq = _quantiles(ours_all), _quantiles(floor_all)
    worst_i = max(range(len(ours_all)), key=lambda i: ours_all[i])
    print(f"PARITY {tag}: worst gradient {names_all[worst_i]} ours {ours_all[worst_i]:.2e} "
          f"its floor {floor_all[worst_i]:.2e}")
    print(f"PARITY {tag}: grads median/p90/max ours {oq[0]:.2e}/{oq[1]:.2e}/{oq[2]:.2e} "
          f"floor {fq[0]:.2e}/{fq[1]:.2e}/{fq[2]:.2e}")
    for name, o, f in zip(("median", "p90", "max"), oq, fq):
        if name == "max" and f >= 0.5:
            # Two legitimate executions of the reference graph disagree by more than half the
            # gradient's norm on their worst parameter: at this conditioning the worst case is one
            # draw from a heavy tail (a single flipped ReLU mask upstream), not a measurement.  The
            # robust quantiles above still bind, and every layer is held to 1e-2 on identical inputs
            # by tests/test_gpu_layers_in_network.py.
            print(f"PARITY {tag}: max not compared (floor {f:.2e} >= 0.5, ours {o:.2e})")
            continue
        assert o <= max(tol, 2 * f), f"{tag}: gradient {name} {o:.3e} > max({tol}, 2 x floor {f:.3e})"


@pytest.mark.gpu
@pytest.mark.parametrize("family,kind,hw,batch", CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
def test_train_step_parity(family, kind, hw, batch, dtype):
    _parity(family, kind, hw, batch, dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("family,kind,hw,batch", CASES_224)
def test_train_step_parity_config_size(family, kind, hw, batch):
    """One full step at the BASELINE.json input size (224 x 224): every layer of configs C2 / C3 / C4
    at its real H x W against the fp64 oracle, bf16 mode (the mode the configs are quoted in)."""
    _parity(family, kind, hw, batch, torch.bfloat16)


def _randomise_running_stats(sd, seed=4):
    """Non-trivial running statistics (a fresh model has mean 0 / var 1: folding would be a no-op)."""
    g = torch.Generator().manual_seed(seed)
    out = dict(sd)
    for k, v in sd.items():
        if k.endswith("running_mean"):
            out[k] = 0.2 * torch.randn(v.shape, generator=g)
        elif k.endswith("running_var"):
            out[k] = 0.5 + torch.rand(v.shape, generator=g)
        elif k.endswith("bn.weight") or (k.endswith(".weight") and v.dim() == 1):
            out[k] = 0.5 + torch.rand(v.shape, generator=g)
        elif k.endswith(".bias") and v.dim() == 1:
            out[k] = 0.2 * torch.randn(v.shape, generator=g)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("family,kind,hw,batch", CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["f32", "bf16"])
@pytest.mark.parametrize("fold", [True, False], ids=["folded", "bn-kernel"])
def test_eval_parity(family, kind, hw, batch, dtype, fold, monkeypatch):
    """Inference forward (reference valid_model / evaluate / test, src/basemodel.py:304-346,508-576,
    601-722): BatchNorm on running statistics, folded into the producing convolution's filter and
    bias / ReLU epilogue (`folded`) or as a separate cnb_bn_fwd_eval pass, against the fp64 oracle's
    eval-mode logits.  Bound: max(tolerance, 2 x the floor of the same graph in fp32)."""
    from convnets_b200 import layers as LY
    monkeypatch.setattr(LY, "_FOLD_BN", fold)
    model, sd, s = _build(family, kind, hw, dtype)
    sd = _randomise_running_stats(sd)
    model.load_state_dict(sd)
    g = torch.Generator().manual_seed(21)
    x = torch.randn(batch, 3, hw, hw, generator=g)
    bf16 = dtype == torch.bfloat16
    storage = torch.bfloat16 if bf16 else None
    tol = 1e-2 if bf16 else 1e-4
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    with torch.no_grad():
        lg64 = Z.forward(family, sd64, x.double(), kind, training=False, storage=storage)[0]
        lg32 = Z.forward(family, sd, x, kind, training=False, storage=storage)[0]
        exact = Z.forward(family, sd64, x.double(), kind, training=False)[0]
    floor = O.rel_l2(lg32, lg64)
    model.eval()
    before = {k: v.detach().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        out = model(x.cuda())
    torch.cuda.synchronize()
    err = O.rel_l2(out, lg64)
    print(f"EVAL {family}-{kind} {'bf16' if bf16 else 'f32'} {'folded' if fold else 'bn-kernel'}: ours {err:.2e} "
          f"floor {floor:.2e} (vs un-rounded fp64: {O.rel_l2(out, exact):.2e})")
    # bf16 + folding rounds w * gamma / sigma once instead of w and y separately: compare against the
    # un-rounded oracle too, where neither rounding order is privileged
    bound = max(tol, 2 * floor)
    if bf16 and fold:
        assert err <= bound or O.rel_l2(out, exact) <= max(bound, 2 * O.rel_l2(lg64, exact)), (err, floor)
    else:
        assert err <= bound, (err, floor)
    after = model.state_dict()
    for k, v in before.items():     # inference must not touch parameters or running statistics
        assert torch.equal(after[k], v), k


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


@pytest.mark.gpu
def test_packed_weights_follow_fused_adam():
    """torch's fused Adam updates parameters WITHOUT bumping their version counter; the cached bf16
    operand copies must still be rebuilt after every optimizer step (layers._WEIGHT_EPOCH)."""
    from convnets_b200 import layers as L
    model, _, s = _build("ResNeXt", "50", 32, torch.bfloat16)
    model.init_optimizer()
    model.train()
    g = torch.Generator().manual_seed(5)
    convs = [m for m in model.modules() if isinstance(m, L.Conv2d) and not m.cast_input]
    assert any(m.groups > 1 for m in convs) and any(m.groups == 1 for m in convs)
    for step in range(3):
        x = torch.randn(4, 3, 32, 32, generator=g).cuda()
        y = torch.randint(0, 10, (4,), generator=g).cuda()
        model.optimizer.zero_grad()
        loss = model.cost_function(model(x).float(), y)
        for m in convs:  # the copy the forward just used == the current weight, rounded
            wb = m.__dict__["_cnb_pack"][1][0]
            assert torch.equal(wb.float(), m.weight.detach().to(torch.bfloat16).float()), (step, m)
        loss.backward()
        model.optimizer.step()


@pytest.mark.gpu
@pytest.mark.parametrize("family,kind", [("ResNet", "18"), ("MobileNetV1", ""), ("SKResNet", "26")])
def test_step_graph_matches_eager_steps(family, kind):
    """The CUDA-graph replay of the training step (convnets_b200.graph.StepGraph) must walk the
    same trajectory as the eagerly enqueued reference loop — every kernel of the library is
    deterministic, so losses and weights agree to rounding after six Adam steps (three eager, one
    capture + replay, two replays) on changing batches."""
    import convnets_b200 as cnb

    def make():
        model, _, s = _build(family, kind, 32, torch.bfloat16)
        model.init_optimizer()
        model.train()
        return model

    g = torch.Generator().manual_seed(3)
    batches = [(torch.randn(8, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (8,), generator=g).cuda())
               for _ in range(6)]
    eager, graphed = make(), make()
    graphed.load_state_dict(eager.state_dict())
    sg = cnb.StepGraph(graphed, eager_steps=3)
    for i, (x, y) in enumerate(batches):
        eager.optimizer.zero_grad()
        out = eager(x)
        l0 = eager.cost_function(out.float(), y)
        l0.backward()
        eager.optimizer.step()
        l1, _ = sg(x, y)
        assert abs(l0.item() - l1.item()) <= 1e-5 * max(1.0, abs(l0.item())), (i, l0.item(), l1.item())
    assert sg.graph is not None and sg.launches_per_step > 0
    worst = 0.0
    for (k, a), (_, b) in zip(eager.named_parameters(), graphed.named_parameters()):
        worst = max(worst, O.rel_l2(b.detach(), a.detach()))
    assert worst <= 1e-5, worst
    # train_model goes through the graph once it is enabled
    cnb.enable_step_graph(graphed, eager_steps=0)
    loss, acc = graphed.train_model(_Loader(batches[:2]))
    assert loss == loss and 0.0 <= acc <= 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("family,kind", [("ResNet", "18"), ("ResNeXt", "50")])
def test_eval_between_graph_replays_uses_current_weights(family, kind):
    """A replayed step changes every weight without running Python (no optimizer hook, no version
    bump).  An eager forward after it — valid_model, or the ragged last batch of an epoch — must
    re-pack its bf16 operand copies from the CURRENT weights (it reused the copies of
    the first eager validation for the rest of training)."""
    import convnets_b200 as cnb
    from convnets_b200 import layers as L
    graphed, _, _ = _build(family, kind, 32, torch.bfloat16)
    twin, _, _ = _build(family, kind, 32, torch.bfloat16)
    graphed.init_optimizer()
    g = torch.Generator().manual_seed(9)
    batches = [(torch.randn(8, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (8,), generator=g).cuda())
               for _ in range(8)]
    xe = torch.randn(8, 3, 32, 32, generator=g).cuda()
    sg = cnb.StepGraph(graphed, eager_steps=2)
    convs = [m for m in graphed.modules() if isinstance(m, L.Conv2d)]
    seen = []
    for rnd in range(2):
        graphed.train()
        for x, y in batches[rnd * 4:rnd * 4 + 4]:
            sg(x, y)
        assert sg.graph is not None
        graphed.eval()
        with torch.no_grad():
            lg = graphed(xe)
        for m in convs:   # the copy the eval forward just used was built from the current weight
            fold = m.__dict__.get("_cnb_fold")
            if fold is not None:   # conv -> BatchNorm: the filter with the running statistics folded in
                bn = m.__dict__["_cnb_bn"]
                scale = torch.rsqrt(bn.running_var.float() + bn.eps) * bn.weight.detach().float()
                want = m.weight.detach().float() * scale.view(-1, 1, 1, 1)
                assert torch.equal(fold[1][0], want), (rnd, tuple(m.weight.shape))
            else:
                wb = m.__dict__["_cnb_pack"][1][0]
                assert torch.equal(wb.float(), m.weight.detach().to(wb.dtype).float()), (rnd, tuple(m.weight.shape))
        # an eager twin that only ever loads the weights (its optimizer never steps between the checks)
        twin.load_state_dict(graphed.state_dict())
        twin.eval()
        with torch.no_grad():
            lt = twin(xe)
        assert torch.equal(lg, lt), (rnd, O.rel_l2(lg, lt))
        seen.append(lg.float().clone())
    assert not torch.equal(seen[0], seen[1])    # training between the two evaluations moved the weights
    # ragged batch after replays: an eager training step, again on current weights
    graphed.train()
    sg(*batches[0])
    twin.load_state_dict(graphed.state_dict())
    twin.train()
    xr, yr = batches[1][0][:5], batches[1][1][:5]
    want = twin.cost_function(twin(xr).float(), yr)
    got, _ = sg(xr, yr)
    assert abs(got.item() - want.item()) <= 1e-6 * max(1.0, abs(want.item())), (got.item(), want.item())


class _Loader:
    def __init__(self, batches):
        self.batches = batches
        self.dataset = range(sum(int(b[1].numel()) for b in batches))

    def __iter__(self):
        return iter(self.batches)
