"""Per-op CPU oracle (test infrastructure; see oracle/__init__.py).

Each function restates one op of the reference's hot path.  The arithmetic is
the torch CPU op the reference itself executes, evaluated in float64 so that it is a noise-free
target for both the fp32-IO and the bf16 mode of the CUDA library.  Index-exact ops also have a
plain numpy restatement that fixes the tie-breaking rules independently of torch.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def _d(t):
    return None if t is None else t.detach().to("cpu", torch.float64)


# --- a7..a12: nn.Conv2d as built by reference src/basemodel.py:966-997 ------------------------
def conv2d(x, w, b=None, stride=1, padding=0, dilation=1, groups=1):
    return F.conv2d(_d(x), _d(w), _d(b), stride, padding, dilation, groups)


def conv2d_grads(x, w, b, dy, stride=1, padding=0, dilation=1, groups=1):
    """(dx, dw, db) of sum(conv2d(x, w, b) * dy) — the convolution_backward the autograd engine
    runs from loss.backward() (reference src/basemodel.py:258,278)."""
    x = _d(x).requires_grad_(True)
    w = _d(w).requires_grad_(True)
    b = None if b is None else _d(b).requires_grad_(True)
    y = F.conv2d(x, w, b, stride, padding, dilation, groups)
    y.backward(_d(dy))
    return x.grad, w.grad, (None if b is None else b.grad)


# --- a13: nn.BatchNorm2d (reference src/basemodel.py:1008,1061) -------------------------------
def _materialise(y, storage):
    """The reference's module chain stores bn(x) in the activation dtype before `output +=
    identity` (src/resnet.py:151,191).  In bf16 mode that rounding decides the sign of
    near-zero sums and hence the ReLU mask, so the oracle models it (value only; the gradient
    passes straight through, as in the reference's autograd graph)."""
    if storage is None or storage == torch.float64:
        return y
    return y + (y.detach().to(storage).to(torch.float64) - y.detach())


def batch_norm_train(x, gamma, beta, running_mean, running_var, momentum=0.1, eps=1e-5,
                     relu=False, residual=None, storage=None):
    """Returns (y, new_running_mean, new_running_var, batch_mean, batch_rstd)."""
    x = _d(x)
    rm, rv = _d(running_mean).clone(), _d(running_var).clone()
    y = F.batch_norm(x, rm, rv, _d(gamma), _d(beta), True, momentum, eps)
    if residual is not None:
        y = _materialise(y, storage) + _d(residual)
    if relu:
        y = F.relu(y)
    mean = x.mean(dim=(0, 2, 3))
    var = x.var(dim=(0, 2, 3), unbiased=False)
    return y, rm, rv, mean, 1.0 / torch.sqrt(var + eps)


def batch_norm_eval(x, gamma, beta, running_mean, running_var, eps=1e-5, relu=False, residual=None,
                    storage=None):
    y = F.batch_norm(_d(x), _d(running_mean), _d(running_var), _d(gamma), _d(beta), False, 0.0, eps)
    if residual is not None:
        y = _materialise(y, storage) + _d(residual)
    return F.relu(y) if relu else y


def batch_norm_train_grads(x, gamma, beta, dy, eps=1e-5, relu=False, residual=None, storage=None):
    """(dx, dgamma, dbeta, dresidual) of the training-mode forward above."""
    x = _d(x).requires_grad_(True)
    g = _d(gamma).requires_grad_(True)
    b = _d(beta).requires_grad_(True)
    r = None if residual is None else _d(residual).requires_grad_(True)
    y = F.batch_norm(x, None, None, g, b, True, 0.1, eps)
    if r is not None:
        y = _materialise(y, storage) + r
    if relu:
        y = F.relu(y)
    y.backward(_d(dy))
    return x.grad, g.grad, b.grad, (None if r is None else r.grad)


# --- a16: nn.MaxPool2d (reference src/basemodel.py:1017-1024) ----------------------------------
def max_pool2d(x, kernel_size, stride=None, padding=0):
    """torch's own result: (values, int64 indices h*W+w)."""
    return F.max_pool2d(_d(x), kernel_size, stride, padding, return_indices=True)


def max_pool2d_numpy(x, kernel_size, stride=None, padding=0):
    """Independent restatement of the window scan: row-major over in-bounds positions, update on
    `v > best or isnan(v)`, index = h*W + w; padding is never selected."""
    x = np.asarray(x, dtype=np.float64)
    k = (kernel_size, kernel_size) if isinstance(kernel_size, int) else kernel_size
    s = k if stride is None else ((stride, stride) if isinstance(stride, int) else stride)
    p = (padding, padding) if isinstance(padding, int) else padding
    N, C, H, W = x.shape
    P = (H + 2 * p[0] - k[0]) // s[0] + 1
    Q = (W + 2 * p[1] - k[1]) // s[1] + 1
    y = np.empty((N, C, P, Q))
    idx = np.empty((N, C, P, Q), dtype=np.int64)
    for n in range(N):
        for c in range(C):
            for oh in range(P):
                for ow in range(Q):
                    h0, w0 = oh * s[0] - p[0], ow * s[1] - p[1]
                    h1, w1 = min(h0 + k[0], H), min(w0 + k[1], W)
                    h0, w0 = max(h0, 0), max(w0, 0)
                    best, bi = -np.inf, h0 * W + w0
                    for h in range(h0, h1):
                        for w in range(w0, w1):
                            v = x[n, c, h, w]
                            if v > best or np.isnan(v):
                                best, bi = v, h * W + w
                    y[n, c, oh, ow] = best
                    idx[n, c, oh, ow] = bi
    return y, idx


def max_pool2d_grad(x, dy, kernel_size, stride=None, padding=0):
    x = _d(x).requires_grad_(True)
    F.max_pool2d(x, kernel_size, stride, padding).backward(_d(dy))
    return x.grad


# --- a17: nn.AvgPool2d, count_include_pad=True (reference src/basemodel.py:1026-1033) ----------
def avg_pool2d(x, kernel_size, stride=None, padding=0):
    return F.avg_pool2d(_d(x), kernel_size, stride, padding, ceil_mode=False, count_include_pad=True)


def avg_pool2d_grad(x, dy, kernel_size, stride=None, padding=0):
    x = _d(x).requires_grad_(True)
    F.avg_pool2d(x, kernel_size, stride, padding).backward(_d(dy))
    return x.grad


# --- a18: nn.AdaptiveAvgPool2d(1) (reference src/basemodel.py:1035-1042) -----------------------
def global_avg_pool2d(x):
    return F.adaptive_avg_pool2d(_d(x), 1)


# --- a19: ShuffleUnit.channel_shuffle (reference src/shufflenet_v1.py:173-190) -----------------
def channel_shuffle(x, groups):
    x = x.detach().cpu()
    b, c, h, w = x.shape
    return x.view(b, groups, c // groups, h, w).transpose(1, 2).contiguous().view(b, -1, h, w)


def channel_shuffle_numpy(x, groups):
    """out[:, j*g + i] = in[:, i*(C/g) + j]"""
    x = np.asarray(x)
    c = x.shape[1]
    cpg = c // groups
    out = np.empty_like(x)
    for i in range(groups):
        for j in range(cpg):
            out[:, j * groups + i] = x[:, i * cpg + j]
    return out


# --- a14/a15: ReLU, residual add ----------------------------------------------------------------
def relu(x):
    return F.relu(_d(x))


def add_relu(a, b):
    return F.relu(_d(a) + _d(b))


# --- a21: SEUnit scale (reference src/se_resnet.py:253) -----------------------------------------
def se_scale(x, s):
    x, s = _d(x), _d(s)
    return x * s.reshape(x.shape[0], x.shape[1], 1, 1).expand_as(x)


# --- a22: SKConv fuse / select (reference src/sknet.py:226-240) ---------------------------------
def sk_fuse_gap(u0, u1):
    fm = torch.stack([_d(u0), _d(u1)], dim=1)
    return F.adaptive_avg_pool2d(torch.sum(fm, dim=1), 1)


def sk_select(u0, u1, logits):
    u0, u1 = _d(u0), _d(u1)
    n, c = u0.shape[:2]
    fm = torch.stack([u0, u1], dim=1)
    att = torch.softmax(_d(logits).reshape(n, 2, c, 1, 1), dim=1)
    return torch.sum(fm * att, dim=1)


def rel_l2(a, b) -> float:
    """||a - b|| / ||b|| in float64 (b is the oracle); 0/0 -> 0."""
    a = torch.as_tensor(a).detach().to("cpu", torch.float64)
    b = torch.as_tensor(b).detach().to("cpu", torch.float64)
    den = torch.linalg.vector_norm(b).item()
    num = torch.linalg.vector_norm(a - b).item()
    if den == 0.0:
        return 0.0 if num == 0.0 else float("inf")
    return num / den


# --- f4: the DataLoader transforms of reference src/mngrdata.py:139-165 -------------------------
def to_tensor_normalize(img_u8_nhwc, mean, std):
    """transforms.ToTensor + transforms.Normalize on a batch of decoded images [N, H, W, C] uint8,
    restated from their published definitions (torchvision is absent here: ToTensor is
    `img.permute(2, 0, 1).contiguous().to(float32).div(255)`, Normalize is
    `tensor.sub_(mean[:, None, None]).div_(std[:, None, None])` with fp32 mean / std tensors).
    Stays in fp32 on purpose: the CUDA kernel is held to it bit for bit."""
    x = torch.as_tensor(img_u8_nhwc).permute(0, 3, 1, 2).contiguous().to(torch.float32).div(255)
    m = torch.as_tensor(mean, dtype=torch.float32)[None, :, None, None]
    s = torch.as_tensor(std, dtype=torch.float32)[None, :, None, None]
    return x.sub_(m).div_(s)


def random_crop_flip(img_u8_nhwc, size, padding, aug):
    """RandomCrop(size, padding) (zero fill, src/mngrdata.py:146) then RandomHorizontalFlip (:147) with
    the draws already made: aug[n] = (top, left, flip).  Plain numpy indexing on the uint8 image."""
    a = np.asarray(img_u8_nhwc)
    n, h, w, c = a.shape
    padded = np.zeros((n, h + 2 * padding, w + 2 * padding, c), dtype=a.dtype)
    padded[:, padding:padding + h, padding:padding + w] = a
    out = np.empty((n, size, size, c), dtype=a.dtype)
    for i in range(n):
        top, left, flip = (int(v) for v in aug[i])
        crop = padded[i, top:top + size, left:left + size]
        out[i] = crop[:, ::-1] if flip else crop
    return out


# --- a4 / f2: optim.Adam as the reference steps it (src/basemodel.py:58-66,284) -------------------
def adam_step(p, g, m, v, step, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0):
    """One update of torch's single-tensor Adam (torch/optim/adam.py `_single_tensor_adam`, the
    path a CPU model takes) restated on numpy float32 arrays, operation by operation; `step` is
    the 1-based count of this update.  Returns (p, m, v) as new float32 arrays.  Pinned against
    torch.optim.Adam on CPU in tests/test_optim.py."""
    f, d = np.float32, np.float64
    p, g, m, v = (np.asarray(a, dtype=f).copy() for a in (p, g, m, v))

    def fma(a, b, c):   # the CPU vector kernels fuse these (product of two fp32 is exact in fp64)
        return (np.asarray(a, dtype=f).astype(d) * np.asarray(b, dtype=f).astype(d) + c.astype(d)).astype(f)

    if weight_decay != 0:
        g = fma(f(weight_decay), p, g)                           # grad.add(param, alpha=wd)
    m = fma(f(1 - beta1), g - m, m)                              # exp_avg.lerp_(grad, 1 - beta1)
    v = fma(f(1 - beta2) * g, g, v * f(beta2))                   # mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2_sqrt = (1 - beta2 ** step) ** 0.5
    denom = np.sqrt(v) / f(bc2_sqrt) + f(eps)                    # (sqrt / bc2_sqrt).add_(eps)
    p = p + (f(-(lr / bc1)) * m) / denom                         # addcdiv_(exp_avg, denom, value=-step_size)
    return p.astype(f), m.astype(f), v.astype(f)
