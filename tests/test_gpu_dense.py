"""Concat-aware DenseNet blocks (reference src/densenet.py:104-168): the three row-strided BatchNorm
entry points against the dense ones / the float64 oracle, and the whole block function against the
per-layer path (torch.cat by slab copies) and a float64 execution of the same modules."""
import copy
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ops as O  # noqa: E402

CL = torch.channels_last


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("n,h,w,ct,c", [(4, 14, 14, 256, 96), (2, 7, 9, 512, 512), (8, 28, 28, 128, 64)])
def test_prefix_batchnorm_forward_and_backward(n, h, w, ct, c):
    from convnets_b200 import functional as F, lib as L
    lib = L.load()
    g = torch.Generator().manual_seed(ct + c)
    buf = (torch.randn(n, ct, h, w, generator=g) * 1.5 + 0.3).bfloat16().cuda().contiguous(memory_format=CL)
    gam, bet = (torch.rand(c, generator=g) + 0.5).cuda(), torch.randn(c, generator=g).cuda()
    m = n * h * w
    xd = buf[:, :c].contiguous(memory_format=CL)               # the prefix as a dense tensor
    # dense reference call: also yields the statistics the prefix call is given
    rm0, rv0 = torch.zeros(c).cuda(), torch.ones(c).cuda()
    ctx = type("Ctx", (), {"save_for_backward": lambda self, *t: setattr(self, "saved_tensors", t),
                           "needs_input_grad": (True,) * 8})()
    y_ref = F._BatchNormFn.forward(ctx, xd, None, gam, bet, rm0, rv0, True, 0.1, 1e-5, True, None)
    _, _, mask_ref, _, mean, rstd = ctx.saved_tensors
    # forward, prefix form
    rm1, rv1 = torch.zeros(c).cuda(), torch.ones(c).cuda()
    nbt = torch.zeros((), dtype=torch.int64).cuda()
    d = L.BnDesc(m, c, L.CNB_BF16, L.CNB_BN_RELU, 1e-5, 0.1, nbt.data_ptr())
    ws = torch.empty(lib.cnb_bn_workspace_bytes(C.byref(d)), dtype=torch.uint8).cuda()
    y = torch.empty_like(xd)
    mask = torch.empty(m * c // 8, dtype=torch.uint8).cuda()
    L.check(lib.cnb_bn_fwd_train_prefix(C.byref(d), _p(buf), ct, _p(mean), _p(rstd), _p(gam), _p(bet), _p(rm1),
                                        _p(rv1), _p(y), _p(mask), _p(ws), ws.numel(), _st()), "fwd prefix")
    torch.cuda.synchronize()
    assert torch.equal(y, y_ref) and torch.equal(mask, mask_ref)
    assert int(nbt) == 1
    assert torch.allclose(rm1, rm0, rtol=1e-6, atol=1e-7) and torch.allclose(rv1, rv0, rtol=1e-5, atol=1e-6)
    # backward, prefix form: accumulate onto an existing gradient buffer
    dy = torch.randn(n, c, h, w, generator=g).bfloat16().cuda().contiguous(memory_format=CL)
    acc0 = torch.randn(n, ct, h, w, generator=g).bfloat16().cuda().contiguous(memory_format=CL)
    acc = acc0.clone(memory_format=CL)
    dgam, dbet = torch.empty(c).cuda(), torch.empty(c).cuda()
    d = L.BnDesc(m, c, L.CNB_BF16, L.CNB_BN_RELU, 1e-5, 0.1, None)
    L.check(lib.cnb_bn_bwd_prefix(C.byref(d), _p(buf), ct, None, _p(mask), _p(dy), _p(mean), _p(rstd), _p(gam),
                                  _p(acc), ct, _p(dgam), _p(dbet), _p(ws), ws.numel(), _st()), "bwd prefix")
    torch.cuda.synchronize()
    dx64, dg64, db64, _ = O.batch_norm_train_grads(xd.float().cpu(), gam.cpu(), bet.cpu(), dy.float().cpu(), 1e-5,
                                                   True, None, storage=torch.bfloat16)
    want = acc0[:, :c].double().cpu() + dx64.double()
    floor = O.rel_l2(want.bfloat16(), want)
    assert O.rel_l2(acc[:, :c], want) <= max(1e-2, 2 * floor)
    assert torch.equal(acc[:, c:], acc0[:, c:])                 # channels outside the prefix untouched
    assert O.rel_l2(dgam, dg64) <= 1e-3 and O.rel_l2(dbet, db64) <= 1e-3


def test_stats_from_conv_partials():
    from convnets_b200 import functional as F, lib as L
    lib = L.load()
    g = torch.Generator().manual_seed(1)
    x = torch.randn(4, 128, 14, 14, generator=g).bfloat16().cuda().contiguous(memory_format=CL)
    wt = (torch.randn(32, 128, 3, 3, generator=g) * 0.05).cuda().contiguous(memory_format=CL)
    y, part = F.conv2d(x, wt, None, 1, 1, 1, 1, bn_stats=True)
    assert part is not None
    m = 4 * 14 * 14
    d = L.BnDesc(m, 32, L.CNB_BF16, 0, 1e-5, 0.1, None)
    ws = torch.empty(lib.cnb_bn_workspace_bytes(C.byref(d)), dtype=torch.uint8).cuda()
    table_m, table_r = torch.zeros(64).cuda(), torch.zeros(64).cuda()
    L.check(lib.cnb_bn_stats_from_partials(C.byref(d), _p(part), part.shape[0],
                                           C.c_void_p(table_m.data_ptr() + 4 * 16),
                                           C.c_void_p(table_r.data_ptr() + 4 * 16), _p(ws), ws.numel(), _st()),
            "stats")
    torch.cuda.synchronize()
    yd = y.double()
    mean = yd.mean(dim=(0, 2, 3))
    rstd = 1.0 / torch.sqrt(yd.var(dim=(0, 2, 3), unbiased=False) + 1e-5)
    assert torch.allclose(table_m[16:48].double(), mean, rtol=1e-4, atol=1e-5)
    assert torch.allclose(table_r[16:48].double(), rstd, rtol=1e-4)
    assert float(table_m[:16].abs().sum() + table_m[48:].abs().sum()) == 0.0


@pytest.mark.parametrize("layers,c0,hw,batch,p", [(6, 64, 16, 4, 0.0), (4, 128, 8, 8, 0.0), (3, 64, 8, 4, 0.5)])
def test_dense_block_buffer_vs_per_layer_and_fp64(layers, c0, hw, batch, p, monkeypatch):
    import convnets_b200 as cnb
    from convnets_b200 import dense, zoo as M
    from convnets_b200.zoo.densenet import DenseBlock
    torch.manual_seed(3)
    s = M.Settings(kind="121", input_size=(3, 32, 32), num_classes=10, dropout_rate=p)
    s.device.set_device("cuda:0")

    class Net(torch.nn.Module):       # the factory interface DenseBlock needs
        def __init__(self):
            super().__init__()
            self.setting = s
            self.in_channels = c0

        def conv2d(self, in_channels=None, num_filters=None, set_output=True, **kw):
            cin = self.in_channels if in_channels is None else in_channels
            layer = torch.nn.Conv2d(cin, num_filters, **kw)
            if set_output:
                self.in_channels = num_filters
            return layer

    def build():
        torch.manual_seed(11)
        net = Net()
        blk = DenseBlock(net, layers, 32, 4)
        for mod in blk.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                torch.nn.init.uniform_(mod.weight, 0.5, 1.5)
                torch.nn.init.normal_(mod.bias, 0, 0.2)
        return blk

    ref64 = build().double()
    a, b = build(), build()
    wrap_a, wrap_b = torch.nn.Sequential(a), torch.nn.Sequential(b)
    cnb.convert(wrap_a)
    cnb.convert(wrap_b)
    wrap_a.cuda().train()
    wrap_b.cuda().train()
    g = torch.Generator().manual_seed(5)
    x = torch.randn(batch, c0, hw, hw, generator=g)
    dout = torch.randn(batch, c0 + 32 * layers, hw, hw, generator=g)
    xa = x.bfloat16().cuda().contiguous(memory_format=CL).requires_grad_(True)
    xb = x.bfloat16().cuda().contiguous(memory_format=CL).requires_grad_(True)
    monkeypatch.setattr(dense, "_ENABLED", True)
    assert dense.eligible(a, xa)
    ya = wrap_a(xa)
    ya.backward(dout.bfloat16().cuda())
    monkeypatch.setattr(dense, "_ENABLED", False)
    yb = wrap_b(xb)
    yb.backward(dout.bfloat16().cuda())
    torch.cuda.synchronize()
    assert ya.shape == yb.shape == (batch, c0 + 32 * layers, hw, hw)
    if p == 0.0:
        # same kernels for every conv, BatchNorm statistics by a different (equally exact) route
        assert O.rel_l2(ya, yb) <= 1e-2
        assert O.rel_l2(xa.grad, xb.grad) <= 3e-2
        x64 = x.bfloat16().double().requires_grad_(True)
        y64 = ref64.train()(x64)
        y64.backward(dout.bfloat16().double())
        ea, eb = O.rel_l2(ya, y64), O.rel_l2(yb, y64)
        print(f"DENSE block L{layers} C{c0}: output buffer {ea:.2e} per-layer {eb:.2e}; "
              f"dx buffer {O.rel_l2(xa.grad, x64.grad):.2e} per-layer {O.rel_l2(xb.grad, x64.grad):.2e}")
        assert ea <= max(2e-2, 2 * eb)
        assert O.rel_l2(xa.grad, x64.grad) <= max(5e-2, 2 * O.rel_l2(xb.grad, x64.grad))
        for (k, pa), (_, pb), (_, p64) in zip(a.named_parameters(), b.named_parameters(), ref64.named_parameters()):
            assert pa.grad is not None, k
            fa, fb = O.rel_l2(pa.grad, p64.grad), O.rel_l2(pb.grad, p64.grad)
            assert fa <= max(5e-2, 2 * fb), (k, fa, fb)
    else:
        # dropout draws differ between the two runs: structure only (every parameter has a gradient,
        # running statistics moved, the input channels pass through unchanged)
        assert torch.equal(ya[:, :c0], xa.detach())
        for k, pa in a.named_parameters():
            assert pa.grad is not None and torch.isfinite(pa.grad).all(), k
    for (k, va), (_, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        if k.endswith("num_batches_tracked"):
            assert int(va) == int(vb) == 1, k
        elif (k.endswith("running_mean") or k.endswith("running_var")) and p == 0.0:
            assert O.rel_l2(va, vb) <= 1e-2, k
