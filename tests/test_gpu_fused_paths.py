"""GPU parity of the fused / batched entry points that sit beside the plain per-layer ones:

  * cnb_conv2d_dgrad_add  — data-grad whose epilogue adds the skip gradient of an identity-shortcut
                            residual block (reference src/resnet.py:139-156,176-196: `output += identity`
                            makes the block input receive two gradients)
  * cnb_pack_filters      — one launch re-packing every tensor-core filter of a model
                            (reference src/basemodel.py:284: every optimizer.step() changes them all)
  * cnb_rowim2col_nchw    — RGB stem as an R x 1 convolution over the row-im2col of the batch
                            (reference src/resnet.py:66, src/vggnet.py:62, src/mobilenet_v1.py:57)

each against the validated plain path (bit for bit where the arithmetic is the same) and against the
float64 oracle.  Also: the train-step parity suite of tests/test_gpu_models.py runs with these
paths on (they are the defaults) — test_switches_off_same_result here holds the "off" paths to the
same answers so neither side can rot.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ops as O  # noqa: E402

CL = torch.channels_last


def _lib():
    from convnets_b200 import lib as L
    return L, L.load()


def _p(t):
    return C.c_void_p(t.data_ptr())


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# (N, C_in, H, K_out, R, pad): conv1 of identity blocks (1x1 reduce), basic-block 3x3s, ragged tiles
DGRAD_ADD = [(8, 256, 56, 64, 1, 0), (8, 512, 28, 128, 1, 0), (8, 1024, 14, 256, 1, 0),
             (16, 2048, 7, 512, 1, 0), (8, 64, 56, 64, 3, 1), (3, 72, 9, 40, 3, 1), (2, 128, 5, 24, 3, 1)]


@pytest.mark.parametrize("n,c,h,k,r,pad", DGRAD_ADD)
def test_dgrad_add_epilogue(n, c, h, k, r, pad):
    from convnets_b200 import functional as F_
    L, lib = _lib()
    torch.manual_seed(5)
    dy = torch.randn(n, k, h, h).bfloat16()
    w = torch.randn(k, c, r, r) / (c * r * r) ** 0.5
    skip = torch.randn(n, c, h, h).bfloat16()
    dyd = dy.cuda().contiguous(memory_format=CL)
    wd = w.cuda().contiguous(memory_format=CL)
    sk = skip.cuda().contiguous(memory_format=CL)
    wb, wt = F_.pack_filter(wd, transposed=True)
    d = L.ConvDesc(n, h, h, c, k, r, r, 1, 1, pad, pad, 1, 1, 1, L.CNB_BF16, 0, L.CNB_ALGO_TC)
    ws = torch.empty(max(lib.cnb_conv2d_workspace_bytes(C.byref(d), L.CNB_DGRAD), 16), dtype=torch.uint8,
                     device="cuda")
    dx0, dx1 = torch.empty_like(sk), torch.empty_like(sk)
    L.check(lib.cnb_conv2d_dgrad(C.byref(d), _p(dyd), _p(wb), _p(wt), _p(dx0), _p(ws), ws.numel(), _st()),
            "dgrad")
    L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), _p(dyd), _p(wb), _p(wt), _p(sk), None, _p(dx1), _p(ws),
                                     ws.numel(), _st()), "dgrad_add")
    torch.cuda.synchronize()
    # the fp32 accumulator is not observable: the fused result (ONE rounding of acc + skip) must lie
    # within one bf16 ulp of the two-rounding result round(round(acc) + skip) ...
    two = dx0.float() + sk.float()
    err = (dx1.float() - two).abs()
    ulp = two.abs() * 2.0 ** -7 + dx0.float().abs() * 2.0 ** -8 + 1e-30
    assert bool((err <= ulp).all()), float((err / ulp).max())
    # ... and match the float64 oracle of dgrad + skip to the bf16 bar
    dx_ref, _, _ = O.conv2d_grads(torch.zeros(n, c, h, h).bfloat16(), w.bfloat16(), None, dy, 1, pad)
    assert O.rel_l2(dx1, dx_ref.double() + skip.double()) <= 1e-2
    # in-place form (dx aliases the addend) is what autograd uses
    inpl = sk.clone()
    L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), _p(dyd), _p(wb), _p(wt), _p(inpl), None, _p(inpl), _p(ws),
                                     ws.numel(), _st()), "dgrad_add(in place)")
    torch.cuda.synchronize()
    assert torch.equal(inpl, dx1)
    # masked addend (the lazy skip gradient): bit for bit the plain call on the pre-masked tensor
    if c % 32 == 0:
        bits = torch.rand(n, h, h, c) < 0.6                       # NHWC order, like the ReLU mask
        packed = (bits.view(-1, 8).to(torch.uint8) << torch.arange(8, dtype=torch.uint8)).sum(1).to(torch.uint8)
        pre = (skip.float() * bits.permute(0, 3, 1, 2).float()).bfloat16().cuda().contiguous(memory_format=CL)
        dx2, dx3 = torch.empty_like(sk), torch.empty_like(sk)
        L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), _p(dyd), _p(wb), _p(wt), _p(pre), None, _p(dx2), _p(ws),
                                         ws.numel(), _st()), "dgrad_add(pre-masked)")
        L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), _p(dyd), _p(wb), _p(wt), _p(sk), _p(packed.cuda()), _p(dx3),
                                         _p(ws), ws.numel(), _st()), "dgrad_add(mask)")
        torch.cuda.synchronize()
        assert torch.equal(dx2, dx3)
        assert torch.equal(F_.mask_apply(sk, packed.cuda()), pre)


@pytest.mark.parametrize("family,kind", [("ResNet", "18"), ("ResNet", "50"), ("DenseNet", "121"),
                                         ("InceptionNetV1", "")])
def test_batched_pack_equals_per_layer_pack(family, kind):
    import convnets_b200 as cnb
    from convnets_b200 import functional as F_
    from convnets_b200 import zoo as M
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, 64, 64), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[family](s)
    cnb.convert(model)
    s.device.move(model)
    group = next(m.__dict__["_cnb_group"] for m in model.modules() if "_cnb_group" in m.__dict__)
    group.refresh()
    torch.cuda.synchronize()
    assert len(group.layers) > 4
    for m, (wb, wt) in zip(group.layers, group.bufs):
        rb, rt = F_.pack_filter(m.weight, transposed=wt is not None)
        assert torch.equal(wb, rb), tuple(m.weight.shape)
        assert wt is None or torch.equal(wt, rt), tuple(m.weight.shape)
        assert torch.equal(wb.float(), m.weight.detach().bfloat16().float())


# (N, C, H, W, K, R, stride, pad, bias): ResNet, SqueezeNet 1.0, VGG / MobileNet, Inception stems
STEMS = [(8, 3, 64, 64, 64, 7, 2, 3, False), (4, 3, 61, 59, 96, 7, 2, 0, True),
         (6, 3, 32, 32, 64, 3, 1, 1, True), (5, 3, 33, 33, 32, 3, 2, 1, False),
         (2, 1, 20, 20, 16, 5, 1, 2, False), (2, 3, 224, 224, 64, 7, 2, 3, False)]


@pytest.mark.parametrize("n,c,h,w,k,r,st,pd,hb", STEMS)
@pytest.mark.parametrize("rows", [True, False], ids=["row-im2col", "full-im2col"])
def test_rgb_stem_vs_oracle(n, c, h, w, k, r, st, pd, hb, rows):
    """Both stem paths against the float64 oracle: output, BatchNorm partial sums, dW, db."""
    from convnets_b200 import functional as F_
    torch.manual_seed(7)
    x = torch.randn(n, c, h, w)
    wt = torch.randn(k, c, r, r) / (c * r * r) ** 0.5
    b = torch.randn(k) if hb else None
    fn = F_.conv2d_input_rows if rows else F_.conv2d_input
    wd = wt.cuda().contiguous(memory_format=CL).requires_grad_(True)
    bd = None if b is None else b.cuda().requires_grad_(True)
    y = fn(x.cuda(), wd, bd, st, pd, 1, torch.bfloat16, bn_stats=not hb)
    part = None
    if isinstance(y, tuple):
        y, part = y
    dy = torch.randn(y.shape).bfloat16()
    y.backward(dy.cuda().contiguous(memory_format=CL))
    torch.cuda.synchronize()
    xq, wq = x.bfloat16(), wt.bfloat16()       # the operand values the kernels see
    y_ref = O.conv2d(xq, wq, b, st, pd)
    _, dw_ref, db_ref = O.conv2d_grads(xq, wq, b, dy, st, pd)
    assert y.shape == y_ref.shape
    assert O.rel_l2(y, y_ref) <= 1e-2
    assert wd.grad.shape == wt.shape and O.rel_l2(wd.grad, dw_ref) <= 1e-2
    if hb:
        assert O.rel_l2(bd.grad, db_ref) <= 1e-2
    if part is not None:
        # column sums / sums of squares of the STORED (bf16) output
        ys = y.detach().float().cpu().permute(0, 2, 3, 1).reshape(-1, k).double()
        tot = part.double().sum(0).cpu()
        assert O.rel_l2(tot[0], ys.sum(0)) <= 1e-3 and O.rel_l2(tot[1], (ys * ys).sum(0)) <= 1e-3


def _train_steps(monkeypatch, family, kind, on, steps=3):
    import convnets_b200 as cnb
    import sys
    from convnets_b200 import layers as LY
    CV = sys.modules["convnets_b200.convert"]     # the module (the package re-exports the function)
    from convnets_b200 import zoo as M
    monkeypatch.setattr(CV, "_FUSED_SKIP", on)
    monkeypatch.setattr(LY, "_BATCHED_PACK", on)
    monkeypatch.setattr(LY, "_STEM_ROWS", on)
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, 64, 64), num_classes=10, dropout_rate=0.0, learning_rate=1e-3)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[family](s)
    cnb.convert(model)
    s.device.move(model)
    model.init_optimizer()
    model.train()
    g = torch.Generator().manual_seed(3)
    out = []
    for _ in range(steps):
        x = torch.randn(8, 3, 64, 64, generator=g).cuda()
        y = torch.randint(0, 10, (8,), generator=g).cuda()
        model.optimizer.zero_grad()
        logits = model(x)
        loss = model.cost_function(logits.float(), y)
        loss.backward()
        grads = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
        model.optimizer.step()
        out.append((logits.detach().float().clone(), loss.item(), grads))
    return out


@pytest.mark.parametrize("family,kind", [("ResNet", "18"), ("ResNet", "50"), ("SEResNet", "50")])
def test_switches_off_same_result(monkeypatch, family, kind):
    """Fused skip gradient + batched pack + row-im2col stem (the defaults) against the plain paths:
    same weights, same batches.  The packs are bit-identical and the skip-gradient fold only removes
    one bf16 rounding, so the forward agrees to the accumulation-order noise of the stem.  This is an
    A/B consistency check between two bf16 executions (parity against the oracle, with these paths on,
    is tests/test_gpu_models.py): whole-network gradients of two legitimate bf16 runs differ by their
    conditioning, so the gradient bar only has to catch a structural error —
    a dropped skip gradient moves every upstream parameter by O(1)."""
    on = _train_steps(monkeypatch, family, kind, True)
    off = _train_steps(monkeypatch, family, kind, False)
    for (l1, s1, g1), (l0, s0, g0) in zip(on[:1], off[:1]):
        assert O.rel_l2(l1, l0) <= 1e-2, O.rel_l2(l1, l0)
        assert abs(s1 - s0) <= 1e-2 * max(1.0, abs(s0))
        gsum = sum(v.norm().item() for v in g0.values())
        errs = sorted(O.rel_l2(g1[k], g0[k]) for k in g0 if g0[k].norm().item() > 1e-6 * gsum)
        print(f"{family}-{kind}: on-vs-off logits {O.rel_l2(l1, l0):.2e}, grads median {errs[len(errs) // 2]:.2e} "
              f"max {errs[-1]:.2e}")
        assert errs[len(errs) // 2] <= 1e-1, errs[len(errs) // 2]


# ------------------------------------------------------------------ depthwise conv -> BatchNorm partials
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("n,c,h,w,stride", [(4, 64, 28, 28, 1), (3, 128, 14, 14, 2), (2, 32, 9, 11, 1),
                                            (2, 68, 7, 7, 2), (8, 512, 4, 4, 1), (2, 1024, 1, 1, 1)])
def test_depthwise_emits_batchnorm_partials(n, c, h, w, stride, dtype, monkeypatch):
    """The 3x3 depthwise forward kernels emit the [rows][2][C] table of per-channel sums / sums of
    squares of the STORED output (Conv2dBlockDW, reference src/mobilenet_v1.py:90-125): the table must
    reproduce the sums of y, the output itself must be bit-identical to the plain launch, and
    BatchNorm fed with the table must agree with BatchNorm that reads y."""
    from convnets_b200 import functional as F
    monkeypatch.setenv("CNB_DW_STATS", "1")      # opt-in path (measured slower on MobileNet-v1, see dwconv.cu)
    g = torch.Generator().manual_seed(c + h)
    x = torch.randn(n, c, h, w, generator=g).to(dtype).cuda().contiguous(memory_format=CL)
    wt = (torch.randn(c, 1, 3, 3, generator=g) * 0.4).cuda()
    y0 = F.conv2d(x, wt, None, stride, 1, 1, c)
    y1, part = F.conv2d(x, wt, None, stride, 1, 1, c, bn_stats=True)
    assert torch.equal(y0, y1)
    if (c // 4) % min(32, 1 << ((c // 4).bit_length() - 1)) != 0:
        assert part is None        # a partly filled channel block: the kernel keeps to the plain form
        return
    assert part is not None and part.shape[1:] == (2, c) and part.shape[0] >= 1
    yd = y1.double()
    s_ref, q_ref = yd.sum(dim=(0, 2, 3)), (yd * yd).sum(dim=(0, 2, 3))
    s, q = part[:, 0].double().sum(0), part[:, 1].double().sum(0)
    scale = yd.abs().sum(dim=(0, 2, 3)).clamp_min(1e-30)
    assert ((s - s_ref).abs() / scale).max().item() < 1e-5
    assert ((q - q_ref).abs() / q_ref.clamp_min(1e-30)).max().item() < 1e-5
    if n * y1.shape[2] * y1.shape[3] > 1:
        gam, bet = torch.rand(c).cuda() + 0.5, torch.randn(c).cuda()
        a = F.batch_norm(y1, gam, bet, None, None, True, 0.1, 1e-5, relu=True, partials=part)
        b = F.batch_norm(y1, gam, bet, None, None, True, 0.1, 1e-5, relu=True)
        assert O.rel_l2(a, b) <= (1e-2 if dtype == torch.bfloat16 else 1e-5)


# ------------------------------------------------------------------ BatchNorm backward sums / lazy skip gradient
def _one_step_grads(monkeypatch, family, kind, sums, lazy, hw=64, batch=8):
    import convnets_b200 as cnb
    from convnets_b200 import functional as F_
    from convnets_b200 import zoo as M
    monkeypatch.setattr(F_, "_BN_BWD_SUMS", sums)
    monkeypatch.setattr(F_, "_LAZY_SKIP", lazy)
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, hw, hw), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = cnb.convert(M.FAMILIES[family](s))
    s.device.move(model)
    model.train()
    g = torch.Generator().manual_seed(3)
    x = torch.randn(batch, 3, hw, hw, generator=g).cuda()
    y = torch.randint(0, 10, (batch,), generator=g).cuda()
    from convnets_b200 import lib as L
    L.trace_begin()
    logits = model(x)
    model.cost_function(logits.float(), y).backward()
    calls = [c[0] for c in L.trace_end()]
    return logits.detach().float(), {k: p.grad.detach().clone() for k, p in model.named_parameters()}, calls


@pytest.mark.parametrize("family,kind,hw", [("ResNet", "18", 64), ("ResNet", "50", 64), ("ResNet", "50", 224),
                                            ("ResNeXt", "50", 64)])
def test_bn_backward_sums_and_lazy_skip_gradient_vs_plain(monkeypatch, family, kind, hw):
    """cnb_conv2d_dgrad_bn + cnb_bn_bwd_from_partials (BatchNorm backward sums formed by the data-grad
    epilogue) and the masked-addend skip gradient against the plain kernels on the same step: the
    forward is the same code, so the logits are bit-identical, and the gradients may differ only by
    the order of fp32 sums and a handful of bf16 roundings downstream of them."""
    batch = 4 if hw == 224 else 8
    l0, g0, n0 = _one_step_grads(monkeypatch, family, kind, False, False, hw, batch)
    for sums, lazy in ((True, False), (False, True), (True, True)):
        l1, g1, n1 = _one_step_grads(monkeypatch, family, kind, sums, lazy, hw, batch)
        assert torch.equal(l0, l1)
        gsum = sum(v.norm().item() for v in g0.values())
        errs = sorted((O.rel_l2(g1[k], g0[k]), k) for k in g0 if g0[k].norm().item() > 1e-6 * gsum)
        med, worst = errs[len(errs) // 2], errs[-1]
        used = n1.count("cnb_bn_bwd_from_partials")
        print(f"{family}-{kind} {hw}: sums={sums} lazy={lazy}: {used} of {used + n1.count('cnb_bn_bwd')} BatchNorm "
              f"backward calls took the data-grad's sums, grads vs plain median {med[0]:.2e} max {worst[0]:.2e} "
              f"({worst[1]})")
        assert (used > 0) == sums and "cnb_bn_bwd_from_partials" not in n0
        assert med[0] <= 2e-2 and worst[0] <= 2e-1, (med, worst)


def test_tensor_core_conv_from_a_thread_that_never_touched_cuda():
    """The tensor-map encodes are driver-API calls and need a context current on the calling thread.
    A thread whose FIRST CUDA-related call is cnb_conv2d_fwd (a C-ABI caller's worker thread; the
    autograd device thread when no gradient had to be materialised before a convolution's backward)
    used to get CUDA_ERROR_INVALID_CONTEXT from cuTensorMapEncodeTiled; the library now binds the
    primary context of the operand's device and retries."""
    import threading
    from convnets_b200 import lib as L
    lib = L.load()
    n, h, w, c, k = 2, 16, 16, 64, 128
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(n, h, w, c, device="cuda", generator=g).bfloat16()
    wt = (torch.randn(k, c, device="cuda", generator=g) / c ** 0.5).bfloat16()
    y = torch.zeros(n, h, w, k, device="cuda", dtype=torch.bfloat16)
    d = L.ConvDesc(n, h, w, c, k, 1, 1, 1, 1, 0, 0, 1, 1, 1, L.CNB_BF16, 0, L.CNB_ALGO_TC)
    ws = torch.empty(max(1, lib.cnb_conv2d_workspace_bytes(C.byref(d), L.CNB_FWD)), dtype=torch.uint8, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = {}

    def worker():
        out["status"] = lib.cnb_conv2d_fwd(C.byref(d), C.c_void_p(x.data_ptr()), C.c_void_p(wt.data_ptr()), None,
                                           C.c_void_p(y.data_ptr()), None, C.c_void_p(ws.data_ptr()), ws.numel(),
                                           stream)
        out["error"] = L.last_error()     # thread-local message

    t = threading.Thread(target=worker)
    t.start()
    t.join()
    torch.cuda.synchronize()
    assert out["status"] == 0, out["error"]
    ref = (x.float().view(-1, c) @ wt.float().t()).view(n, h, w, k)
    assert O.rel_l2(y.float(), ref) < 1e-2
