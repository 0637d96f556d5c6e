"""Functional CPU restatement of the reference's model families and of one training step.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Each network is a plain function over a
reference-style state_dict (same keys as the reference's nn.Module tree); the arithmetic is the
torch CPU op the reference itself reaches, in whatever dtype the state_dict carries (fp32 = the
reference's own precision; fp64 = a noise-free target).  Architecture constants that are not in
the state_dict (strides, paddings, stage layout) are restated here with the reference line they
come from.  Pinned against the real reference by oracle/make_golden.py -> tests/golden/.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F

EPS, MOMENTUM = 1e-5, 0.1  # nn.BatchNorm2d defaults used at reference src/basemodel.py:1008


class Net:
    """Execution context: parameters/buffers by state_dict key, train/eval mode, and the BN
    running-stat updates produced by a training forward."""

    def __init__(self, sd: Dict[str, torch.Tensor], training: bool = True, storage=None):
        self.sd = sd
        self.training = training
        self.storage = storage  # None = exact; torch.bfloat16 = model the library's bf16 mode
        self.new_buffers: Dict[str, torch.Tensor] = {}

    def q(self, t):
        """Storage rounding of the library's bf16 mode: feature maps and the conv operand copy of
        the weights live in bf16, arithmetic and accumulation are wider.  Value-only (the gradient
        passes straight through), so the oracle stays differentiable."""
        if self.storage is None:
            return t
        return t + (t.detach().to(self.storage).to(t.dtype) - t.detach())

    def has(self, key: str) -> bool:
        return key in self.sd

    # nn.Conv2d (reference src/basemodel.py:966-997): bias present iff batch_norm=False
    def conv(self, key, x, stride=1, padding=0, dilation=1, groups=1):
        return self.q(F.conv2d(x, self.q(self.sd[key + ".weight"]), self.sd.get(key + ".bias"),
                               stride, padding, dilation, groups))

    # nn.BatchNorm2d (reference src/basemodel.py:1008,1061)
    def bn(self, key, x):
        g, b = self.sd[key + ".weight"], self.sd[key + ".bias"]
        rm, rv = self.sd[key + ".running_mean"], self.sd[key + ".running_var"]
        if not self.training:
            return F.batch_norm(x, rm, rv, g, b, False, 0.0, EPS)
        rm, rv = rm.clone(), rv.clone()
        y = F.batch_norm(x, rm, rv, g, b, True, MOMENTUM, EPS)
        self.new_buffers[key + ".running_mean"] = rm
        self.new_buffers[key + ".running_var"] = rv
        self.new_buffers[key + ".num_batches_tracked"] = self.sd[key + ".num_batches_tracked"] + 1
        return y

    # conv2d_block (reference src/basemodel.py:999-1015): Sequential(conv[, bn][, relu]) with
    # integer child names 0, 1, 2
    def block(self, key, x, activation=True, residual=None, **conv_kw):
        """`residual`: the block's output is the last term of `output += identity; relu`
        (src/resnet.py:151-154).  In bf16 mode bn(x) is materialised in the storage dtype before
        the add, as the reference's module chain would do, and the sum is rounded after the ReLU."""
        y = self.conv(key + ".0", x, **conv_kw)
        if self.has(key + ".1.weight"):
            y = self.bn(key + ".1", y)
        if residual is not None:
            return self.q(F.relu(self.q(y) + residual))
        return self.q(F.relu(y) if activation else y)

    def linear(self, key, x):
        return F.linear(x, self.sd[key + ".weight"], self.sd.get(key + ".bias"))


# ------------------------------------------------------------------------------------------------
# ResNet (reference src/resnet.py:17-196)
_RESNET = {  # kind -> (bottleneck, blocks per stage)   src/resnet.py:29-36
    "18": (False, (2, 2, 2, 2)), "34": (False, (3, 4, 6, 3)), "26": (True, (2, 2, 2, 2)),
    "50": (True, (3, 4, 6, 3)), "101": (True, (3, 4, 23, 3)), "152": (True, (3, 8, 36, 3)),
}


def resnet(net: Net, x, kind="50"):
    bottleneck, depths = _RESNET[str(kind)]
    x = net.block("features.0", net.q(x), stride=2, padding=3)     # 7x7 s2 stem, src/resnet.py:66
    x = F.max_pool2d(x, 3, 2, 1)                                   # src/resnet.py:67
    idx = 2
    for stage, depth in enumerate(depths):
        for i in range(depth):
            stride = 2 if (i == 0 and stage > 0) else 1           # src/resnet.py:70-79
            p = f"features.{idx}"
            identity = x
            if net.has(p + ".dim_synch.0.weight"):                 # src/resnet.py:90-93
                identity = net.block(p + ".dim_synch", x, activation=False, stride=stride)
            if bottleneck:                                         # src/resnet.py:158-196
                b = p + ".res_bottleneck"
                y = net.block(b + ".0", x)
                y = net.block(b + ".1", y, stride=stride, padding=1)
                x = net.block(b + ".2", y, residual=identity)      # += identity; relu  (:191-194)
            else:                                                  # src/resnet.py:119-156
                b = p + ".res_block"
                y = net.block(b + ".0", x, stride=stride, padding=1)
                x = net.block(b + ".1", y, residual=identity, padding=1)
            idx += 1
    x = F.adaptive_avg_pool2d(x, 1)                                # src/resnet.py:81
    x = torch.flatten(x, 1)
    return net.linear("classifier.1", x)                           # Dropout(p) is identity at p=0


MODELS = {"ResNet": resnet}


def forward(name: str, sd, x, kind, training=True, storage=None):
    net = Net(sd, training, storage)
    return MODELS[name](net, x, kind), net


# ------------------------------------------------------------------------------------------------
# one optimisation step of MultiClassBaseModel.train_model (reference src/basemodel.py:245-296)
def is_param(key: str) -> bool:
    return not key.endswith(("running_mean", "running_var", "num_batches_tracked"))


def loss_and_grads(name, sd, x, y, kind, storage=None):
    """CrossEntropyLoss(reduction='sum') (src/basemodel.py:46) and its gradients with respect to
    every parameter.  Returns (logits, loss, grads by key, new BN buffers)."""
    leaf = {k: (v.detach().clone().requires_grad_(True) if is_param(k) else v) for k, v in sd.items()}
    logits, net = forward(name, leaf, x, kind, training=True, storage=storage)
    loss = F.cross_entropy(logits, y, reduction="sum")
    keys = [k for k in leaf if is_param(k)]
    grads = torch.autograd.grad(loss, [leaf[k] for k in keys], allow_unused=True)
    return logits.detach(), loss.detach(), dict(zip(keys, grads)), net.new_buffers


def adam_step(sd, grads, state, lr=0.01, weight_decay=1e-4, betas=(0.9, 0.999), eps=1e-8):
    """torch.optim.Adam as configured at reference src/basemodel.py:58-66: weight decay is an L2
    term added to the gradient; bias-corrected moments; eps outside the square root."""
    state["t"] = t = state.get("t", 0) + 1
    out = dict(sd)
    for k, g in grads.items():
        if g is None:
            continue
        p = sd[k]
        g = g + weight_decay * p
        m = state.get(("m", k), torch.zeros_like(p)) * betas[0] + (1 - betas[0]) * g
        v = state.get(("v", k), torch.zeros_like(p)) * betas[1] + (1 - betas[1]) * g * g
        state[("m", k)], state[("v", k)] = m, v
        denom = (v.sqrt() / math.sqrt(1 - betas[1] ** t)) + eps
        out[k] = p - (lr / (1 - betas[0] ** t)) * (m / denom)
    return out


def train_step(name, sd, x, y, kind, state, lr=0.01, weight_decay=1e-4):
    logits, loss, grads, newbuf = loss_and_grads(name, sd, x, y, kind)
    out = adam_step(sd, grads, state, lr, weight_decay)
    out.update(newbuf)
    return out, logits, loss, grads
