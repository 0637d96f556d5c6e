"""GPU parity of the bandwidth-bound ops (BatchNorm, pooling, element-wise, data movement,
gating) through the Python wrappers -> C ABI, against oracle/ops.py.

bf16 mode: rel-L2 <= 1e-2; fp32-IO mode: rel-L2 <= 1e-4; max-pool argmax indices and channel
shuffle bit-exact (BASELINE.json north_star).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ops as O  # noqa: E402

DT = [torch.float32, torch.bfloat16]
TOL = {torch.float32: 1e-4, torch.bfloat16: 1e-2}
CL = torch.channels_last


def F_():
    from convnets_b200 import functional
    return functional


def dev(t, requires_grad=False):
    t = t.cuda()
    if t.dim() == 4:
        t = t.contiguous(memory_format=CL)
    return t.requires_grad_(requires_grad)


def rnd(*shape, dtype=torch.float32, scale=1.0, shift=0.0):
    return (torch.randn(*shape) * scale + shift).to(dtype)


# ------------------------------------------------------------------ BatchNorm
@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("shape", [(4, 64, 14, 14), (3, 60, 7, 5), (2, 10, 9, 9), (8, 32, 1, 1),
                                   (2, 256, 28, 28), (1, 2048, 7, 7), (8, 64, 56, 56), (4, 128, 37, 41)])
@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True), (False, True)])
@pytest.mark.parametrize("fused_bwd", [True, False], ids=["bwd1pass", "bwd2pass"])
def test_batch_norm_train(dtype, shape, relu, res, fused_bwd, monkeypatch):
    # backward: the single-read cooperative kernel (tensors that fit L2) / the reduce + dx pair
    monkeypatch.setenv("CNB_BN_BWD_FUSED_MB", "64" if fused_bwd else "0")
    if fused_bwd and dtype != torch.bfloat16:
        pytest.skip("the single-read backward is the bf16 hot path")
    torch.manual_seed(1)
    n, c, h, w = shape
    x = rnd(*shape, dtype=dtype, scale=2.0, shift=3.0)  # mean >> 0: exercises the shifted sums
    r = rnd(*shape, dtype=dtype) if res else None
    gamma, beta = torch.rand(c) + 0.5, torch.randn(c)
    rm, rv = torch.randn(c), torch.rand(c) + 0.5
    dy = rnd(*shape, dtype=dtype)
    y_ref, rm_ref, rv_ref, mean_ref, rstd_ref = O.batch_norm_train(x, gamma, beta, rm, rv, 0.1, 1e-5, relu, r, storage=dtype)
    dx_ref, dg_ref, db_ref, dr_ref = O.batch_norm_train_grads(x, gamma, beta, dy, 1e-5, relu, r, storage=dtype)

    xd, rd = dev(x, True), (dev(r, True) if res else None)
    gd, bd = dev(gamma, True), dev(beta, True)
    rmd, rvd = rm.cuda(), rv.cuda()
    y = F_().batch_norm(xd, gd, bd, rmd, rvd, True, 0.1, 1e-5, relu=relu, residual=rd)
    y.backward(dev(dy))
    tol = TOL[dtype]
    assert y.shape == x.shape and y.is_contiguous(memory_format=CL)
    assert O.rel_l2(y, y_ref) <= tol
    assert O.rel_l2(rmd, rm_ref) <= 1e-5 and O.rel_l2(rvd, rv_ref) <= 1e-4
    # bf16: the only error source is the bf16 rounding of the stored dx (measured floor: the fp64
    # result rounded to bf16, ~1.7e-3); the bound is north_star's 1e-2 or twice that floor
    dx_floor = O.rel_l2(dx_ref.to(dtype), dx_ref)
    e_dx = O.rel_l2(xd.grad, dx_ref)
    assert e_dx <= max(tol, 2 * dx_floor), (e_dx, dx_floor)
    assert O.rel_l2(gd.grad, dg_ref) <= tol and O.rel_l2(bd.grad, db_ref) <= tol
    if res:
        assert O.rel_l2(rd.grad, dr_ref) <= tol


@pytest.mark.parametrize("dtype", DT)
def test_batch_norm_eval_and_errors(dtype):
    torch.manual_seed(2)
    x = rnd(2, 24, 6, 6, dtype=dtype)
    gamma, beta, rm, rv = torch.rand(24) + 0.5, torch.randn(24), torch.randn(24), torch.rand(24) + 0.5
    y = F_().batch_norm(dev(x), gamma.cuda(), beta.cuda(), rm.cuda(), rv.cuda(), False, 0.1, 1e-5, relu=True)
    assert O.rel_l2(y, O.batch_norm_eval(x, gamma, beta, rm, rv, 1e-5, relu=True)) <= TOL[dtype]
    with pytest.raises(ValueError):  # torch: "Expected more than 1 value per channel when training"
        F_().batch_norm(dev(rnd(1, 8, 1, 1, dtype=dtype)), None, None, None, None, True)


def test_batch_norm_module_state():
    from convnets_b200 import layers
    torch.manual_seed(3)
    m = layers.BatchNorm2d(16).cuda()
    ref = torch.nn.BatchNorm2d(16)
    x = rnd(4, 16, 5, 5)
    for _ in range(3):
        m(dev(x))
        ref(x)
    assert int(m.num_batches_tracked) == 3
    assert O.rel_l2(m.running_mean, ref.running_mean) <= 1e-5
    assert O.rel_l2(m.running_var, ref.running_var) <= 1e-5
    assert list(m.state_dict().keys()) == list(ref.state_dict().keys())


# ------------------------------------------------------------------ pooling
POOLS = [(3, 2, 1), (2, 2, 0), (3, 2, 0), (3, 1, 1)]


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("k,s,p", POOLS)
@pytest.mark.parametrize("shape", [(2, 64, 17, 17), (2, 24, 8, 9), (1, 6, 5, 5)])
def test_max_pool(dtype, k, s, p, shape):
    torch.manual_seed(4)
    x = rnd(*shape, dtype=dtype)
    x = (x * 4).round() / 4  # many exact ties -> the first-maximum rule matters
    y_ref, idx_ref = O.max_pool2d(x, k, s, p)
    dy = rnd(*y_ref.shape, dtype=dtype)
    xd = dev(x, True)
    y, idx = F_().max_pool2d(xd, k, s, p, return_indices=True)
    y.backward(dev(dy))
    assert torch.equal(y.float().cpu(), y_ref.float())                    # values exact
    assert torch.equal(idx.cpu().long(), idx_ref)                         # argmax bit-exact
    assert O.rel_l2(xd.grad, O.max_pool2d_grad(x, dy, k, s, p)) <= TOL[dtype]


def test_max_pool_nan_and_padding():
    x = torch.full((1, 2, 4, 4), -1.0)
    x[0, 0, 1, 1] = float("nan")
    y_ref, idx_ref = O.max_pool2d(x, 3, 2, 1)
    y, idx = F_().max_pool2d(dev(x), 3, 2, 1, return_indices=True)
    assert torch.equal(torch.isnan(y.cpu()), torch.isnan(y_ref))
    assert torch.equal(idx.cpu().long(), idx_ref)
    assert float(y[0, 1].max()) == -1.0  # padding (-inf) never wins, 0 never appears


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("k,s,p", [(3, 2, 1), (2, 2, 0)])
def test_avg_pool(dtype, k, s, p):
    torch.manual_seed(5)
    x = rnd(2, 40, 14, 14, dtype=dtype)
    y_ref = O.avg_pool2d(x, k, s, p)
    dy = rnd(*y_ref.shape, dtype=dtype)
    xd = dev(x, True)
    y = F_().avg_pool2d(xd, k, s, p)
    y.backward(dev(dy))
    assert O.rel_l2(y, y_ref) <= TOL[dtype]
    assert O.rel_l2(xd.grad, O.avg_pool2d_grad(x, dy, k, s, p)) <= TOL[dtype]


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("shape", [(4, 2048, 7, 7), (3, 60, 5, 4), (2, 256, 56, 56)])
def test_global_avg_pool(dtype, shape):
    torch.manual_seed(6)
    x = rnd(*shape, dtype=dtype, shift=1.0)
    xd = dev(x, True)
    y = F_().global_avg_pool2d(xd)
    assert y.dtype == torch.float32 and tuple(y.shape) == (shape[0], shape[1], 1, 1)
    assert O.rel_l2(y, O.global_avg_pool2d(x)) <= 1e-5
    dy = torch.randn(shape[0], shape[1], 1, 1)
    y.backward(dy.cuda())
    ref = (dy.double() / (shape[2] * shape[3])).expand(shape)
    assert O.rel_l2(xd.grad, ref) <= TOL[dtype]


# ------------------------------------------------------------------ element-wise
@pytest.mark.parametrize("dtype", DT)
def test_relu_add_relu(dtype):
    torch.manual_seed(7)
    a, b = rnd(2, 30, 7, 9, dtype=dtype), rnd(2, 30, 7, 9, dtype=dtype)
    ad, bd = dev(a, True), dev(b, True)
    y = F_().relu(ad)
    assert torch.equal(y.float().cpu(), O.relu(a).float())
    y.backward(dev(b))
    assert torch.equal(ad.grad.float().cpu(), (b.float() * (a.float() > 0)))
    ad.grad = None
    z = F_().add_relu(ad, bd)
    z_ref = O.add_relu(a, b)
    assert O.rel_l2(z, z_ref) <= TOL[dtype]
    g = rnd(2, 30, 7, 9, dtype=dtype)
    z.backward(dev(g))
    mask = (z.float().cpu() > 0)
    assert torch.equal(ad.grad.float().cpu(), g.float() * mask)
    assert torch.equal(bd.grad.float().cpu(), g.float() * mask)
    # 2-D classifier activations use the same kernel
    v = rnd(5, 4096, dtype=dtype)
    assert torch.equal(F_().relu(v.cuda()).float().cpu(), O.relu(v).float())


# ------------------------------------------------------------------ data movement
@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c,g", [(60, 3), (240, 3), (24, 1), (272, 4), (384, 8), (30, 3), (21, 3),
                                 (960, 3), (136, 2)])
def test_channel_shuffle_bit_exact(dtype, c, g):
    torch.manual_seed(8)
    x = rnd(2, c, 7, 7, dtype=dtype)
    xd = dev(x, True)
    y = F_().channel_shuffle(xd, g)
    ref = O.channel_shuffle(x, g)
    assert torch.equal(y.cpu(), ref)                                                  # bit-exact
    assert np.array_equal(O.channel_shuffle_numpy(x.float().numpy(), g), ref.float().numpy())
    dy = rnd(2, c, 7, 7, dtype=dtype)
    y.backward(dev(dy))
    xr = x.clone().float().requires_grad_(True)
    O.channel_shuffle(xr, g)  # torch view path for the gradient
    yr = xr.view(2, g, c // g, 7, 7).transpose(1, 2).contiguous().view(2, -1, 7, 7)
    yr.backward(dy.float())
    assert torch.equal(xd.grad.float().cpu(), xr.grad)


@pytest.mark.parametrize("dtype", DT)
def test_cat_channels(dtype):
    torch.manual_seed(9)
    xs = [rnd(2, c, 6, 6, dtype=dtype) for c in (64, 32, 20, 3)]
    xds = [dev(t, True) for t in xs]
    y = F_().cat_channels(xds)
    assert torch.equal(y.cpu(), torch.cat(xs, dim=1))
    dy = rnd(*y.shape, dtype=dtype)
    y.backward(dev(dy))
    off = 0
    for t, td in zip(xs, xds):
        assert torch.equal(td.grad.cpu(), dy[:, off:off + t.shape[1]])
        off += t.shape[1]


@pytest.mark.parametrize("dtype", DT)
def test_to_nhwc_and_cast(dtype):
    torch.manual_seed(10)
    x = torch.randn(3, 3, 33, 31)
    y = F_().to_nhwc(x.cuda(), dtype)
    assert y.is_contiguous(memory_format=CL) and y.dtype == dtype
    assert torch.equal(y.cpu(), x.to(dtype))
    w = torch.randn(64, 16, 3, 3).cuda().contiguous(memory_format=CL)
    p = F_().pack_weight(w, dtype)
    assert torch.equal(p.cpu(), w.cpu().to(dtype)) and p.is_contiguous(memory_format=CL)


# ------------------------------------------------------------------ gating
@pytest.mark.parametrize("dtype", DT)
def test_se_scale(dtype):
    torch.manual_seed(11)
    x, s = rnd(3, 64, 9, 9, dtype=dtype), torch.rand(3, 64, 1, 1)
    xd, sd = dev(x, True), s.cuda().requires_grad_(True)
    y = F_().se_scale(xd, sd)
    s_q = s.to(dtype)  # the gate is applied in the storage dtype, as under the reference's autocast
    assert O.rel_l2(y, O.se_scale(x, s_q)) <= TOL[dtype]
    dy = rnd(3, 64, 9, 9, dtype=dtype)
    y.backward(dev(dy))
    assert O.rel_l2(xd.grad, dy.double() * s_q.double()) <= TOL[dtype]
    assert O.rel_l2(sd.grad, (dy.double() * x.double()).sum(dim=(2, 3), keepdim=True)) <= TOL[dtype]


@pytest.mark.parametrize("dtype", DT)
def test_sk_fuse_select(dtype):
    torch.manual_seed(12)
    n, c, h, w = 3, 128, 7, 7
    u0, u1 = rnd(n, c, h, w, dtype=dtype), rnd(n, c, h, w, dtype=dtype)
    logits = torch.randn(n, 2 * c, 1, 1)
    u0d, u1d, ld = dev(u0, True), dev(u1, True), logits.cuda().requires_grad_(True)
    g = F_().sk_fuse_gap(u0d, u1d)
    assert O.rel_l2(g, O.sk_fuse_gap(u0, u1)) <= 1e-5
    y = F_().sk_select(u0d, u1d, ld)
    assert O.rel_l2(y, O.sk_select(u0, u1, logits)) <= TOL[dtype]
    dy = rnd(n, c, h, w, dtype=dtype)
    dg = torch.randn(n, c, 1, 1)
    (y.float() * dev(dy).float()).sum().backward(retain_graph=True)
    a, b, l = (t.double().requires_grad_(True) for t in (u0, u1, logits))
    fm = torch.stack([a, b], dim=1)
    att = torch.softmax(l.reshape(n, 2, c, 1, 1), dim=1)
    (torch.sum(fm * att, dim=1) * dy.double()).sum().backward()
    assert O.rel_l2(u0d.grad, a.grad) <= TOL[dtype] and O.rel_l2(u1d.grad, b.grad) <= TOL[dtype]
    assert O.rel_l2(ld.grad, l.grad) <= TOL[dtype]
    u0d.grad = u1d.grad = None
    g.backward(dg.cuda())
    ref = (dg.double() / (h * w)).expand(n, c, h, w)
    assert O.rel_l2(u0d.grad, ref) <= TOL[dtype] and O.rel_l2(u1d.grad, ref) <= TOL[dtype]


# ------------------------------------------------------------------ error behaviour
def test_errors_are_loud():
    from convnets_b200 import lib
    import ctypes as C
    L = lib.load()
    d = lib.ConvDesc(1, 8, 8, 6, 8, 3, 3, 1, 1, 1, 1, 1, 1, 4, lib.CNB_F32, 0, 0)  # 6 % 4 != 0
    assert L.cnb_conv2d_workspace_bytes(C.byref(d), 0) == 0
    P, Q = C.c_int32(), C.c_int32()
    assert L.cnb_conv2d_out_dims(C.byref(d), C.byref(P), C.byref(Q)) == -1
    assert "divisible by groups" in lib.last_error()
    # CNB_ALGO_TC on an fp32 problem must be refused, never silently rerouted
    d = lib.ConvDesc(1, 8, 8, 64, 64, 3, 3, 1, 1, 1, 1, 1, 1, 1, lib.CNB_F32, 0, lib.CNB_ALGO_TC)
    assert L.cnb_conv2d_uses_tc(C.byref(d), 0) == -2
    with pytest.raises(RuntimeError):
        F_().relu(torch.randn(4))  # CPU tensor: no fallback


# ------------------------------------------------------------------ network-input convolution
@pytest.mark.parametrize("case", [
    dict(n=3, c=3, hw=33, k=64, r=7, stride=2, pad=3, bias=False),   # ResNet-family stem
    dict(n=2, c=3, hw=18, k=24, r=3, stride=1, pad=1, bias=True),    # VGG first layer, no BN
    dict(n=2, c=1, hw=12, k=16, r=5, stride=1, pad=2, bias=False),   # single-channel input
])
def test_conv2d_input_im2col(case):
    """cnb_im2col_nchw + 1x1-descriptor GEMM == conv2d of the fp32 NCHW batch (bf16 mode)."""
    torch.manual_seed(5)
    n, c, hw, k, r = case["n"], case["c"], case["hw"], case["k"], case["r"]
    x = torch.randn(n, c, hw, hw)
    w = torch.randn(k, c, r, r) / (c * r * r) ** 0.5
    b = torch.randn(k) if case["bias"] else None
    xq, wq = x.bfloat16(), w.bfloat16()
    y_ref = O.conv2d(xq, wq, b, case["stride"], case["pad"])
    dy = torch.randn(y_ref.shape).bfloat16()
    _, dw_ref, db_ref = O.conv2d_grads(xq, wq, b, dy, case["stride"], case["pad"])
    wd = w.cuda().contiguous(memory_format=CL).requires_grad_(True)
    bd = None if b is None else b.cuda().requires_grad_(True)
    y = F_().conv2d_input(x.cuda(), wd, bd, case["stride"], case["pad"], 1, torch.bfloat16)
    assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=CL)
    y.backward(dy.cuda().contiguous(memory_format=CL))
    assert O.rel_l2(y, y_ref) <= 1e-2
    assert O.rel_l2(wd.grad, dw_ref) <= 1e-2
    if b is not None:
        assert O.rel_l2(bd.grad, db_ref) <= 1e-2


@pytest.mark.parametrize("dtype", DT)
@pytest.mark.parametrize("c,h,w,r,stride,pad,dil", [
    (3, 33, 37, 7, 2, 3, 1),     # ResNet-family stem (output-staged kernel, odd column-block count)
    (3, 29, 29, 7, 2, 0, 1),     # SqueezeNet 1.0 stem, no padding
    (3, 18, 20, 3, 1, 1, 1),     # VGG / MobileNet first layer (even column-block count: padded rows)
    (1, 12, 12, 5, 1, 2, 1),     # single-channel input
    (5, 11, 13, 3, 2, 2, 2),     # dilated taps
])
def test_im2col_nchw_bit_exact(dtype, c, h, w, r, stride, pad, dil):
    """cnb_im2col_nchw through the C ABI: column (r*S + s)*C + c of row (n, p, q) holds the rounded
    input pixel (or zero in the padding / alignment columns) — bit-exact against torch unfold."""
    import ctypes as C
    from convnets_b200 import lib
    L = lib.load()
    torch.manual_seed(13)
    n = 3
    x = torch.randn(n, c, h, w)
    code = lib.CNB_BF16 if dtype == torch.bfloat16 else lib.CNB_F32
    v = 8 if dtype == torch.bfloat16 else 4
    rsc = r * r * c
    kd = (rsc + v - 1) // v * v
    d = lib.ConvDesc(n, h, w, c, 8, r, r, stride, stride, pad, pad, dil, dil, 1, code, 0, lib.CNB_ALGO_AUTO)
    P, Q = C.c_int32(), C.c_int32()
    assert L.cnb_conv2d_out_dims(C.byref(d), C.byref(P), C.byref(Q)) == 0
    P, Q = P.value, Q.value
    xd = x.cuda()
    col = torch.full((n * P * Q, kd), float("nan"), dtype=dtype, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = L.cnb_im2col_nchw(C.byref(d), C.c_void_p(xd.data_ptr()), C.c_void_p(col.data_ptr()), kd,
                           C.c_void_p(st))
    assert rc == 0, lib.last_error()
    torch.cuda.synchronize()
    u = torch.nn.functional.unfold(x, r, dilation=dil, padding=pad, stride=stride)   # [n, c*r*r, P*Q]
    u = u.view(n, c, r * r, P * Q).permute(0, 3, 2, 1).reshape(n * P * Q, rsc).to(dtype)
    ref = torch.zeros(n * P * Q, kd, dtype=dtype)
    ref[:, :rsc] = u
    assert torch.equal(col.cpu(), ref)


# ------------------------------------------------------------------ conv-epilogue statistics
@pytest.mark.parametrize("shape,k,ks,stride", [((4, 64, 14, 14), 256, 1, 1), ((3, 128, 9, 9), 96, 3, 1),
                                              ((2, 64, 15, 15), 64, 3, 2), ((8, 256, 7, 7), 512, 1, 1)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_conv_bn_partials_and_relu_mask(shape, k, ks, stride, relu, res):
    """conv2d(bn_stats=True) emits per-tile sums of the stored output; BatchNorm fed with them
    (and using the 1-bit ReLU mask in backward) must agree with the oracle on the same conv
    output, forward and backward."""
    f = F_()
    dt = torch.bfloat16
    n, c, h, w = shape
    x = rnd(*shape, dtype=dt)
    wt = rnd(k, c, ks, ks, scale=(2.0 / (c * ks * ks)) ** 0.5)
    xd, wd = dev(x), dev(wt)
    y, part = f.conv2d(xd, wd, None, stride, ks // 2, 1, 1, wpack=f.pack_weight(wd, dt), bn_stats=True)
    assert part is not None and part.shape[1:] == (2, k)
    yf = y.float()
    assert O.rel_l2(part[:, 0].sum(0), yf.sum(dim=(0, 2, 3))) <= 1e-4
    assert O.rel_l2(part[:, 1].sum(0), (yf * yf).sum(dim=(0, 2, 3))) <= 1e-4
    g, b = 1.0 + 0.2 * torch.randn(k), 0.1 * torch.randn(k)
    rm, rv = torch.zeros(k), torch.ones(k)
    r = rnd(*y.shape, dtype=dt) if res else None
    yin = y.detach().clone().requires_grad_(True)
    gd, bd = dev(g, True), dev(b, True)
    rd = dev(r, True) if res else None
    rmd, rvd = rm.cuda(), rv.cuda()
    out = f.batch_norm(yin, gd, bd, rmd, rvd, True, 0.1, 1e-5, relu=relu, residual=rd, partials=part)
    want, rm2, rv2, _, _ = O.batch_norm_train(y, g, b, rm, rv, relu=relu, residual=r, storage=dt)
    assert O.rel_l2(out, want) <= 1e-2
    assert O.rel_l2(rmd, rm2) <= 1e-3 and O.rel_l2(rvd, rv2) <= 1e-3
    dy = rnd(*y.shape, dtype=dt)
    out.backward(dev(dy))
    dx, dg, db, dr = O.batch_norm_train_grads(y, g, b, dy, relu=relu, residual=r, storage=dt)
    assert O.rel_l2(yin.grad, dx) <= 1e-2
    assert O.rel_l2(gd.grad, dg) <= 1e-2 and O.rel_l2(bd.grad, db) <= 1e-2
    if res:
        assert O.rel_l2(rd.grad, dr) <= 1e-2


# ------------------------------------------------------------------ uint8 input pipeline (f4)
CINIC_MEAN = (0.47889522, 0.47227842, 0.43047404)
CINIC_STD = (0.24205776, 0.23828046, 0.25874835)


@pytest.mark.parametrize("shape", [(4, 32, 32, 3), (2, 224, 224, 3), (3, 17, 23, 3), (2, 9, 10, 1)])
@pytest.mark.parametrize("nhwc", [True, False])
def test_normalize_u8_bit_exact(shape, nhwc):
    """cnb_normalize_u8 == ToTensor + Normalize (src/mngrdata.py:151-153), every bit."""
    g = torch.Generator().manual_seed(5)
    img = torch.randint(0, 256, shape, generator=g, dtype=torch.uint8)
    c = shape[3]
    ref = O.to_tensor_normalize(img, CINIC_MEAN[:c], CINIC_STD[:c])
    src = img if nhwc else img.permute(0, 3, 1, 2).contiguous()
    got = F_().normalize_u8(src.cuda(), CINIC_MEAN[:c], CINIC_STD[:c], nhwc=nhwc).cpu()
    assert got.shape == ref.shape and got.dtype == torch.float32
    assert torch.equal(got.view(torch.int32), ref.view(torch.int32))


@pytest.mark.parametrize("n,hw,size,pad", [(16, 32, 32, 4), (5, 32, 28, 0), (3, 30, 32, 4)])
def test_normalize_u8_crop_flip_bit_exact(n, hw, size, pad):
    """RandomCrop(32, padding=4) + RandomHorizontalFlip (src/mngrdata.py:146-147) as index arithmetic."""
    from convnets_b200.loader import PrefetchLoader
    g = torch.Generator().manual_seed(9)
    img = torch.randint(0, 256, (n, hw, hw, 3), generator=g, dtype=torch.uint8)
    ld = PrefetchLoader([], "cuda", crop=size, padding=pad, flip_p=0.5, generator=g)
    aug = ld.draw_augmentation(n, hw, hw)
    assert int(aug[:, 0].max()) <= hw + 2 * pad - size and int(aug[:, 2].max()) <= 1
    ref = O.to_tensor_normalize(O.random_crop_flip(img.numpy(), size, pad, aug.numpy()), CINIC_MEAN, CINIC_STD)
    got = F_().normalize_u8(img.cuda(), CINIC_MEAN, CINIC_STD, nhwc=True, crop=size, padding=pad,
                            aug=aug.cuda()).cpu()
    assert torch.equal(got.view(torch.int32), ref.view(torch.int32))


def test_prefetch_loader_normalizes_uint8_batches():
    from convnets_b200.loader import PrefetchLoader
    g = torch.Generator().manual_seed(3)
    batches = [(torch.randint(0, 256, (8, 32, 32, 3), generator=g, dtype=torch.uint8).pin_memory(),
                torch.randint(0, 10, (8,), generator=g).pin_memory()) for _ in range(5)]
    out = list(PrefetchLoader(batches, "cuda", normalize=(CINIC_MEAN, CINIC_STD), nhwc=True))
    torch.cuda.synchronize()
    assert len(out) == 5
    for (x, y), (xh, yh) in zip(out, batches):
        assert x.is_cuda and x.dtype == torch.float32 and tuple(x.shape) == (8, 3, 32, 32)
        assert torch.equal(x.cpu(), O.to_tensor_normalize(xh, CINIC_MEAN, CINIC_STD))
        assert torch.equal(y.cpu(), yh)
