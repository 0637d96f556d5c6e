"""Full-size properties of the CUDA path (BASELINE.json configs: batch 256 per GPU, 224x224 nets).

The oracle (CPU, float64) cannot be run at these sizes in seconds, so parity at full size is held
through size-independent properties of the operators, all through the Python wrappers -> C ABI:

* convolution: exact homogeneity (scaling an operand by 2 scales every floating-point partial sum
  by 2, so forward / data-grad / weight-grad must be BIT-identical up to that factor), batch
  sharding (a sample's output does not depend on which other samples share the launch — the data-
  parallel partitioning of SURVEY.md section 8e), run-to-run determinism of the split-K weight grad;
* BatchNorm: the normalised output has zero mean / unit variance per channel, dx is orthogonal to
  1 and to x-hat per channel, dbeta = sum(dy), the residual gradient is the masked dy bit for bit;
* max-pool: y is the value AT the returned argmax, the gradient is conserved;
* channel shuffle: the backward permutation is the inverse of the forward one, bit for bit, and
  both match the view/transpose definition of reference src/shufflenet_v1.py:173-190.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

CL = torch.channels_last
B = 256  # BASELINE.json: batch 256 per GPU


def F_():
    from convnets_b200 import functional
    return functional


def cl(t):
    return t.contiguous(memory_format=CL)


def conv_all(x, w, dy, stride, pad, groups=1):
    """(y, dx, dw) of one convolution on fresh leaves (so every call has its own autograd graph)."""
    f = F_()
    xd = cl(x.detach().clone()).requires_grad_(True)
    wd = cl(w.detach().clone()).requires_grad_(True)
    y = f.conv2d(xd, wd, None, stride, pad, 1, groups, wpack=f.pack_weight(wd, x.dtype))
    y.backward(cl(dy))
    torch.cuda.synchronize()
    return y.detach(), xd.grad.detach(), wd.grad.detach()


# ResNet-50 @224 layer signatures and one grouped ResNeXt-50 layer
CONV_CASES = [
    dict(c=64, k=64, r=3, hw=56, stride=1, pad=1),      # stage-1 3x3 (im2col TMA, N tile 64)
    dict(c=256, k=64, r=1, hw=56, stride=1, pad=0),     # stage-1 1x1 reduce (HBM-bound GEMM)
    dict(c=128, k=128, r=3, hw=56, stride=2, pad=1),    # strided 3x3: dgrad by parity classes
    dict(c=512, k=2048, r=1, hw=7, stride=1, pad=0),    # stage-4 expand (N tiles of 256)
    dict(c=128, k=128, r=3, hw=56, stride=1, pad=1, groups=32),   # ResNeXt-50 32x4d
]


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda d: "C{c}K{k}R{r}H{hw}s{stride}".format(**d))
def test_conv_full_size_homogeneity_sharding_determinism(case):
    torch.manual_seed(31)
    c, k, r, hw, stride, pad = (case[n] for n in ("c", "k", "r", "hw", "stride", "pad"))
    g = case.get("groups", 1)
    dev = torch.device("cuda:0")
    x = cl(torch.randn(B, c, hw, hw, device=dev).bfloat16())
    w = cl(torch.randn(k, c // g, r, r, device=dev) / (c // g * r * r) ** 0.5)
    ho = (hw + 2 * pad - r) // stride + 1
    dy = cl(torch.randn(B, k, ho, ho, device=dev).bfloat16())

    y, dx, dw = conv_all(x, w, dy, stride, pad, g)
    assert y.shape == (B, k, ho, ho) and torch.isfinite(y.float()).all()
    assert torch.isfinite(dx.float()).all() and torch.isfinite(dw).all()

    # determinism: same inputs, same bits (fixed-order split-K and BatchNorm-free epilogues)
    y_b, dx_b, dw_b = conv_all(x, w, dy, stride, pad, g)
    assert torch.equal(y, y_b) and torch.equal(dx, dx_b) and torch.equal(dw, dw_b)

    # exact homogeneity in the activations ...
    y2, _, dw2 = conv_all(x * 2, w, dy, stride, pad, g)
    assert torch.equal(y2.float(), y.float() * 2)
    assert torch.equal(dw2, dw * 2)
    # ... and in the upstream gradient
    _, dx3, dw3 = conv_all(x, w, dy * 2, stride, pad, g)
    assert torch.equal(dx3.float(), dx.float() * 2)
    assert torch.equal(dw3, dw * 2)

    # batch sharding: the second half of the batch alone gives the second half of y and dx
    h = B // 2
    y_h, dx_h, dw_h = conv_all(x[h:], w, dy[h:], stride, pad, g)
    assert torch.equal(y_h, y[h:])
    assert torch.equal(dx_h, dx[h:])
    # weight gradients of the two shards add up to the full-batch one (different split-K
    # partitions: equal to fp32 summation noise, not bit for bit)
    _, _, dw_l = conv_all(x[:h], w, dy[:h], stride, pad, g)
    num = (dw_l.double() + dw_h.double() - dw.double()).norm()
    assert (num / dw.double().norm()).item() <= 1e-4


@pytest.mark.parametrize("shape", [(B, 64, 112, 112), (B, 256, 56, 56), (B, 2048, 7, 7)])
@pytest.mark.parametrize("relu,res", [(False, False), (True, True)])
def test_batch_norm_full_size_properties(shape, relu, res):
    torch.manual_seed(32)
    f = F_()
    dev = torch.device("cuda:0")
    n, c, h, w = shape
    m = n * h * w
    x = cl((torch.randn(shape, device=dev) * 1.5 + 0.7).bfloat16())
    r = cl(torch.randn(shape, device=dev).bfloat16()) if res else None
    gamma = (torch.rand(c, device=dev) + 0.5).requires_grad_(True)
    beta = torch.randn(c, device=dev).requires_grad_(True)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    xd = x.clone().requires_grad_(True)
    rd = r.clone().requires_grad_(True) if res else None
    y = f.batch_norm(xd, gamma, beta, rm, rv, True, 0.1, 1e-5, relu=relu, residual=rd)
    dy = cl(torch.randn(shape, device=dev).bfloat16())
    y.backward(dy)
    torch.cuda.synchronize()

    xf = x.float()
    mean = xf.mean(dim=(0, 2, 3))
    var = xf.var(dim=(0, 2, 3), unbiased=False)
    # running statistics: momentum 0.1 from (0, 1); unbiased variance
    assert (rm - 0.1 * mean).abs().max().item() <= 1e-4
    assert (rv - (0.9 + 0.1 * var * m / (m - 1))).abs().max().item() <= 1e-3
    xhat = (xf - mean.view(1, -1, 1, 1)) * torch.rsqrt(var + 1e-5).view(1, -1, 1, 1)

    if not relu and not res:
        # forward: (y - beta) / gamma is x-hat up to the bf16 rounding of y
        yn = (y.detach().float() - beta.detach().view(1, -1, 1, 1)) / gamma.detach().view(1, -1, 1, 1)
        assert yn.mean(dim=(0, 2, 3)).abs().max().item() <= 5e-3
        assert (yn.var(dim=(0, 2, 3), unbiased=False) - 1).abs().max().item() <= 2e-2
        mask = None
    else:
        mask = (y.detach() > 0)
        # the residual branch receives the ReLU-masked upstream gradient, bit for bit
        assert torch.equal(rd.grad, torch.where(mask, dy, torch.zeros_like(dy)))

    dyf = dy.float() if mask is None else torch.where(mask, dy, torch.zeros_like(dy)).float()
    dxf = xd.grad.float()
    # dx is orthogonal to the constant and to x-hat in every channel
    scale = dxf.abs().sum(dim=(0, 2, 3)).clamp_min(1e-30)
    assert (dxf.sum(dim=(0, 2, 3)).abs() / scale).max().item() <= 2e-3
    assert ((dxf * xhat).sum(dim=(0, 2, 3)).abs() / scale).max().item() <= 5e-3
    # dbeta = sum dy', dgamma = sum dy' * x-hat
    dscale = dyf.abs().sum(dim=(0, 2, 3))
    assert ((beta.grad - dyf.sum(dim=(0, 2, 3))).abs() / dscale).max().item() <= 1e-4
    assert ((gamma.grad - (dyf * xhat).sum(dim=(0, 2, 3))).abs() / dscale).max().item() <= 1e-3


def test_max_pool_full_size_argmax_and_conservation():
    torch.manual_seed(33)
    f = F_()
    dev = torch.device("cuda:0")
    n, c, hw = B, 64, 112   # the stem pool of the ResNet families at 224x224
    x = cl(((torch.randn(n, c, hw, hw, device=dev) * 4).round() / 4).bfloat16())   # many ties
    xd = x.clone().requires_grad_(True)
    y, idx = f.max_pool2d(xd, 3, 2, 1, return_indices=True)
    assert y.shape == (n, c, 56, 56)
    # y is the input value at the returned argmax (h*W + w inside the (n, c) plane)
    picked = x.reshape(n, c, hw * hw).gather(2, idx.reshape(n, c, -1).long()).reshape(y.shape)
    assert torch.equal(picked, y.detach())
    # ... and no element of the window is larger (checked against a shifted-max of the padded plane)
    xp = torch.nn.functional.pad(x.float(), (1, 1, 1, 1), value=float("-inf"))
    best = torch.full_like(y, float("-inf"), dtype=torch.float32)
    for dh in range(3):
        for dw_ in range(3):
            best = torch.maximum(best, xp[:, :, dh:dh + 112:2, dw_:dw_ + 112:2])
    assert torch.equal(best, y.detach().float())
    dy = cl(torch.randn(y.shape, device=dev).bfloat16())
    y.backward(dy)
    torch.cuda.synchronize()
    dx = xd.grad.float()
    # every dy lands on exactly one input position: sums agree up to the bf16 rounding of dx
    num = (dx.sum(dim=(2, 3)) - dy.float().sum(dim=(2, 3))).abs()
    den = dy.float().abs().sum(dim=(2, 3))
    assert (num / den).max().item() <= 2e-3
    # positions that are nobody's argmax get exactly zero
    hit = torch.zeros(n, c, hw * hw, device=dev, dtype=torch.float32)
    hit.scatter_(2, idx.reshape(n, c, -1).long(), 1.0)
    assert (dx.reshape(n, c, -1) * (hit == 0)).abs().sum().item() == 0.0


@pytest.mark.parametrize("c,g,hw", [(60, 3, 28), (240, 3, 28), (480, 3, 14), (272, 4, 28), (384, 8, 14)])
def test_channel_shuffle_full_size_inverse(c, g, hw):
    torch.manual_seed(34)
    f = F_()
    dev = torch.device("cuda:0")
    x = cl(torch.randn(B, c, hw, hw, device=dev).bfloat16())
    xd = x.clone().requires_grad_(True)
    y = f.channel_shuffle(xd, g)
    # definition (reference src/shufflenet_v1.py:173-190): view (g, C/g) -> transpose -> flatten
    ref = x.view(B, g, c // g, hw, hw).transpose(1, 2).reshape(B, c, hw, hw)
    assert torch.equal(y.detach(), ref)
    # applying the inverse permutation restores x; autograd's backward IS that inverse
    assert torch.equal(f.channel_shuffle(y.detach(), c // g), x)
    dy = cl(torch.randn(B, c, hw, hw, device=dev).bfloat16())
    y.backward(dy)
    torch.cuda.synchronize()
    ref_dx = dy.view(B, c // g, g, hw, hw).transpose(1, 2).reshape(B, c, hw, hw)
    assert torch.equal(xd.grad, ref_dx)
