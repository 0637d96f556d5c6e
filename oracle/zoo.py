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

    def cb(self, conv_key, bn_key, x, activation=True, residual=None, **conv_kw):
        """conv [-> bn] [-> += residual] [-> relu].  `residual`: the block's output is the last
        term of `output += identity; relu` (src/resnet.py:151-154).  In bf16 mode bn(x) is
        materialised in the storage dtype before the add, as the reference's module chain would
        do, and the sum is rounded after the ReLU."""
        y = self.conv(conv_key, x, **conv_kw)
        if self.has(bn_key + ".weight"):
            y = self.bn(bn_key, y)
        if residual is not None:
            return self.q(F.relu(self.q(y) + residual))
        return self.q(F.relu(y) if activation else y)

    # conv2d_block (reference src/basemodel.py:999-1015): Sequential(conv[, bn][, relu]) with
    # integer child names 0, 1, 2
    def block(self, key, x, activation=True, residual=None, **conv_kw):
        return self.cb(key + ".0", key + ".1", x, activation, residual, **conv_kw)

    # class Conv2dBlock (reference src/basemodel.py:1045-1067): children conv, bn, relu
    def nblock(self, key, x, activation=True, residual=None, **conv_kw):
        return self.cb(key + ".conv", key + ".bn", x, activation, residual, **conv_kw)

    def bn_relu(self, key, x):
        """BatchNorm2d -> ReLU(inplace) pre-activation pair (src/densenet.py:117-118,124-125)."""
        return self.q(F.relu(self.bn(key, x)))

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


def _head(net, x, key="classifier.1"):
    """GAP output -> flatten -> Dropout(p=0) -> Linear (e.g. src/resnet.py:104-116)."""
    return net.linear(key, torch.flatten(x, 1))


# ------------------------------------------------------------------------------------------------
# VGGNet (reference src/vggnet.py:17-101): 3x3 p1 conv blocks and 2x2 max-pools from the config
# table (:30-35), GAP (:70), Dropout/Linear(4096)/ReLU x2 + Linear (:78-92)
_VGG = {
    "11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M",
           512, 512, 512, 512, "M"],
}


def vggnet(net: Net, x, kind="11"):
    x = net.q(x)
    for i, e in enumerate(_VGG[str(kind)]):
        x = F.max_pool2d(x, 2, 2) if e == "M" else net.block(f"features.{i}", x, padding=1)
    x = torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)
    x = F.relu(net.linear("classifier.1", x))
    x = F.relu(net.linear("classifier.4", x))
    return net.linear("classifier.7", x)


# ------------------------------------------------------------------------------------------------
# InceptionNet-v1 (reference src/inceptionnet_v1.py:17-143): stem (:45-49), nine InceptionBlocks
# with max-pools between the stages (:52-66), four branches concatenated (:104-142)
def _inception_block(net, p, x):
    b1 = net.block(p + ".branch1", x)
    b2 = net.block(p + ".branch2.1", net.block(p + ".branch2.0", x), padding=1)
    b3 = net.block(p + ".branch3.1", net.block(p + ".branch3.0", x), padding=2)      # 5x5 p2
    b4 = net.block(p + ".branch4.1", F.max_pool2d(x, 3, 1, 1))
    return torch.cat([b1, b2, b3, b4], dim=1)


def inceptionnet_v1(net: Net, x, kind=""):
    x = net.block("features.0", net.q(x), stride=2, padding=3)
    x = F.max_pool2d(x, 3, 2, 1)
    x = net.block("features.2", x)
    x = net.block("features.3", x, padding=1)
    x = F.max_pool2d(x, 3, 2, 1)
    for i in (5, 6):
        x = _inception_block(net, f"features.{i}", x)
    x = F.max_pool2d(x, 3, 2, 1)
    for i in (8, 9, 10, 11, 12):
        x = _inception_block(net, f"features.{i}", x)
    x = F.max_pool2d(x, 3, 2, 1)
    for i in (14, 15):
        x = _inception_block(net, f"features.{i}", x)
    return _head(net, F.adaptive_avg_pool2d(x, 1))


# ------------------------------------------------------------------------------------------------
# ResNeXt (reference src/resnext.py:17-190): widths 128..1024, expansion 2, cardinality 32 on the
# 3x3 (:84-95,:168); NO global average pool before the classifier (:66-80) -> NCHW flatten
_RESNEXT = {"26": (2, 2, 2, 2), "50": (3, 4, 6, 3), "101": (3, 4, 23, 3), "152": (3, 8, 36, 3)}


def resnext(net: Net, x, kind="50"):
    x = net.block("features.0", net.q(x), stride=2, padding=3)
    x = F.max_pool2d(x, 3, 2, 1)
    idx = 2
    for stage, depth in enumerate(_RESNEXT[str(kind)]):
        for i in range(depth):
            stride = 2 if (i == 0 and stage > 0) else 1
            p = f"features.{idx}"
            identity = x
            if net.has(p + ".dim_synch.0.weight"):
                identity = net.block(p + ".dim_synch", x, activation=False, stride=stride)
            b = p + ".res_bottleneck"
            y = net.block(b + ".0", x)
            y = net.block(b + ".1", y, stride=stride, padding=1, groups=32)
            x = net.block(b + ".2", y, residual=identity)
            idx += 1
    return _head(net, x)


# ------------------------------------------------------------------------------------------------
# SqueezeNet (reference src/squeezenet.py:17-154): Fire = squeeze 1x1 -> cat(expand 1x1, expand
# 3x3 p1) (:127-153); un-padded max-pools; classifier = 1x1 conv block + GAP (:104-121)
_SQUEEZE = {
    "1.0": [("conv", 7, 2), ("pool",), ("fire",), ("fire",), ("fire",), ("pool",), ("fire",),
            ("fire",), ("fire",), ("fire",), ("pool",), ("fire",)],
    "1.1": [("conv", 3, 2), ("pool",), ("fire",), ("fire",), ("pool",), ("fire",), ("fire",),
            ("pool",), ("fire",), ("fire",), ("fire",), ("fire",)],
}


def squeezenet(net: Net, x, kind="1.1"):
    x = net.q(x)
    for i, cfg in enumerate(_SQUEEZE[str(kind)]):
        p = f"features.{i}"
        if cfg[0] == "conv":
            x = net.block(p, x, stride=cfg[2])                      # padding 0 (:96)
        elif cfg[0] == "pool":
            x = F.max_pool2d(x, 3, 2)
        else:
            s = net.block(p + ".squeeze", x)
            x = torch.cat([net.block(p + ".expand_1x1", s),
                           net.block(p + ".expand_3x3", s, padding=1)], dim=1)
    x = net.block("classifier.1", x)                                # Dropout(p=0) before it
    return torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)


# ------------------------------------------------------------------------------------------------
# DenseNet (reference src/densenet.py:17-190): pre-activation BN-ReLU-conv1x1(4k) / BN-ReLU-
# conv3x3(k) layers whose output is concatenated to their input (:104-140); transitions BN-ReLU-
# conv1x1(C/2)-AvgPool 2/2 (:171-190); final BN-ReLU-GAP (:80-82).  Dropout layers are p=0.
_DENSE = {"121": (6, 12, 24, 16), "169": (6, 12, 32, 32), "201": (6, 12, 48, 32),
          "264": (6, 12, 64, 48), "161": (6, 12, 36, 24)}


def densenet(net: Net, x, kind="121"):
    blocks = _DENSE[str(kind)]
    x = net.block("features.0", net.q(x), stride=2, padding=3)
    x = F.max_pool2d(x, 3, 2, 1)
    idx = 2
    for bi, depth in enumerate(blocks):
        for j in range(depth):
            p = f"features.{idx}.dense_block.{j}"
            y = net.conv(p + ".bottleneck.3", net.bn_relu(p + ".bottleneck.0", x))
            y = net.conv(p + ".conv_block.3", net.bn_relu(p + ".conv_block.0", y), padding=1)
            x = torch.cat([x, y], dim=1)
        idx += 1
        if bi != len(blocks) - 1:
            p = f"features.{idx}"
            y = net.conv(p + ".conv", net.bn_relu(p + ".batch_norm", x))
            x = net.q(F.avg_pool2d(y, 2, 2))
            idx += 1
    x = net.bn_relu(f"features.{idx}", x)
    return _head(net, F.adaptive_avg_pool2d(x, 1))


# ------------------------------------------------------------------------------------------------
# MobileNet-v1 (reference src/mobilenet_v1.py:17-125): 3x3 s2 stem Conv2dBlock, thirteen
# depthwise-separable blocks conv_dw/bn_dw/relu_dw/conv_pw/bn_pw/relu_pw (:90-125), GAP
_MOBILE_STRIDES = (1, 2, 1, 2, 1, 2, 1, 1, 1, 1, 1, 2, 1)                           # :49-62


def mobilenet_v1(net: Net, x, kind=""):
    x = net.nblock("features.0", net.q(x), stride=2, padding=1)
    for i, stride in enumerate(_MOBILE_STRIDES, start=1):
        p = f"features.{i}"
        c = x.shape[1]
        x = net.cb(p + ".conv_dw", p + ".bn_dw", x, stride=stride, padding=1, groups=c)
        x = net.cb(p + ".conv_pw", p + ".bn_pw", x)
    return _head(net, F.adaptive_avg_pool2d(x, 1))


# ------------------------------------------------------------------------------------------------
# ShuffleNet-v1 (reference src/shufflenet_v1.py:17-190): 3x3 s2 stem (24), max-pool, units of
# grouped 1x1 compress -> channel shuffle -> depthwise 3x3 (no ReLU) -> grouped 1x1 expand (no
# ReLU) (:96-171); downsample units concatenate the 3x3/2 avg-pooled input FIRST (:148,:167) and
# emit C_out - C_in channels (:115-116); the very first unit's compress conv is ungrouped (:111)
# but the shuffle still uses `groups` (:162,:179)
_SHUFFLE_OUT = {"g1": 144, "g2": 200, "g3": 240, "g4": 272, "g8": 384}                # :30-36


def channel_shuffle(x, groups):
    """out[:, j*g + i] = in[:, i*(C/g) + j]  (src/shufflenet_v1.py:173-190)"""
    b, c, h, w = x.shape
    return x.reshape(b, groups, c // groups, h, w).transpose(1, 2).reshape(b, c, h, w)


def shufflenet_v1(net: Net, x, kind="g3"):
    g = int(str(kind)[1:])
    x = net.block("features.0", net.q(x), stride=2, padding=1)
    x = F.max_pool2d(x, 3, 2, 1)
    idx = 2
    for stage, repeat in enumerate((3, 7, 3)):
        for j in range(1 + repeat):
            down = j == 0
            p = f"features.{idx}"
            first = stage == 0 and j == 0
            y = net.nblock(p + ".conv1x1_group_compress", x, groups=1 if first else g)
            y = channel_shuffle(y, g)
            c = y.shape[1]
            y = net.nblock(p + ".conv3x3_depthwise", y, activation=False, stride=2 if down else 1,
                           padding=1, groups=c)
            if down:
                y = net.nblock(p + ".conv1x1_group_expand", y, activation=False, groups=g)
                identity = net.q(F.avg_pool2d(x, 3, 2, 1))        # count_include_pad=True
                x = net.q(F.relu(torch.cat([identity, y], dim=1)))
            else:
                x = net.nblock(p + ".conv1x1_group_expand", y, residual=x, groups=g)
            idx += 1
    return _head(net, F.adaptive_avg_pool2d(x, 1))


# ------------------------------------------------------------------------------------------------
# Squeeze-and-excitation (reference src/se_resnet.py:216-255 = src/senet.py:216-255): GAP ->
# Linear(C, C/16, no bias) -> ReLU -> Linear(C/16, C, no bias) -> sigmoid -> x * s
def _se_unit(net, p, x):
    s = torch.flatten(F.adaptive_avg_pool2d(x, 1), 1)
    s = F.relu(F.linear(s, net.sd[p + ".excitation.0.weight"]))
    s = torch.sigmoid(F.linear(s, net.sd[p + ".excitation.2.weight"]))
    return net.q(x * s[:, :, None, None])


def _se_family(net, x, depths, body, bottleneck, groups):
    """Shared by SEResNet (src/se_resnet.py:17-214) and SENet (src/senet.py:17-214): ResNet /
    ResNeXt layout with an SEUnit as the last child of the block body; `output += identity;
    relu` follows the SE scaling (:206-211)."""
    x = net.block("features.0", net.q(x), stride=2, padding=3)
    x = F.max_pool2d(x, 3, 2, 1)
    idx = 2
    for stage, depth in enumerate(depths):
        for i in range(depth):
            stride = 2 if (i == 0 and stage > 0) else 1
            p = f"features.{idx}"
            identity = x
            if net.has(p + ".dim_synch.0.weight"):
                identity = net.block(p + ".dim_synch", x, activation=False, stride=stride)
            b = f"{p}.{body}"
            if bottleneck:
                y = net.nblock(b + ".0", x)
                y = net.nblock(b + ".1", y, stride=stride, padding=1, groups=groups)
                y = net.nblock(b + ".2", y, activation=False)
                y = _se_unit(net, b + ".3", y)
            else:
                y = net.nblock(b + ".0", x, stride=stride, padding=1)
                y = net.nblock(b + ".1", y, activation=False, padding=1)
                y = _se_unit(net, b + ".2", y)
            x = net.q(F.relu(y + identity))
            idx += 1
    return _head(net, F.adaptive_avg_pool2d(x, 1))


def se_resnet(net: Net, x, kind="50"):
    bottleneck, depths = _RESNET[str(kind)]
    return _se_family(net, x, depths, "res_bottleneck" if bottleneck else "res_block", bottleneck, 1)


def senet(net: Net, x, kind="50"):
    return _se_family(net, x, _RESNEXT[str(kind)], "se_bottleneck", True, 32)


# ------------------------------------------------------------------------------------------------
# Selective kernel (reference src/sknet.py:173-242 = src/sk_resnet.py:171-240): two grouped 3x3
# paths (dilation 1 and 2, padding = dilation, g = 32, shared stride) -> sum -> GAP -> 1x1 conv +
# BN + ReLU descriptor (d = max(C/16, 32)) -> two 1x1 convs WITH bias -> softmax across the paths
# -> weighted sum
def _sk_conv(net, p, x, stride):
    u = [net.nblock(f"{p}.kernels.{i}", x, stride=stride, padding=1 + i, dilation=1 + i, groups=32)
         for i in range(2)]
    s = F.adaptive_avg_pool2d(u[0] + u[1], 1)
    z = net.cb(p + ".descriptor.conv", p + ".descriptor.bn", s) if net.storage is None else \
        F.relu(net.bn(p + ".descriptor.bn", F.conv2d(s, net.sd[p + ".descriptor.conv.weight"])))
    a = torch.stack([F.conv2d(z, net.sd[f"{p}.attentions.{i}.0.weight"],
                              net.sd[f"{p}.attentions.{i}.0.bias"]) for i in range(2)], dim=1)
    a = torch.softmax(a, dim=1)
    return net.q(u[0] * a[:, 0] + u[1] * a[:, 1])


def _sk_family(net, x, depths, body):
    x = net.block("features.0", net.q(x), stride=2, padding=3)
    x = F.max_pool2d(x, 3, 2, 1)
    idx = 2
    for stage, depth in enumerate(depths):
        for i in range(depth):
            stride = 2 if (i == 0 and stage > 0) else 1
            p = f"features.{idx}"
            identity = x
            if net.has(p + ".dim_synch.0.weight"):
                identity = net.block(p + ".dim_synch", x, activation=False, stride=stride)
            b = f"{p}.{body}"
            y = net.nblock(b + ".0", x)
            y = _sk_conv(net, b + ".1", y, stride)
            x = net.nblock(b + ".2", y, residual=identity)
            idx += 1
    return _head(net, F.adaptive_avg_pool2d(x, 1))


def sk_resnet(net: Net, x, kind="50"):
    return _sk_family(net, x, _RESNEXT[str(kind)], "res_bottleneck")   # src/sk_resnet.py:29-34


def sknet(net: Net, x, kind="50"):
    return _sk_family(net, x, _RESNEXT[str(kind)], "sk_bottleneck")    # src/sknet.py:29-34


MODELS = {
    "VGGNet": vggnet, "InceptionNetV1": inceptionnet_v1, "ResNet": resnet, "ResNeXt": resnext,
    "SqueezeNet": squeezenet, "DenseNet": densenet, "MobileNetV1": mobilenet_v1,
    "ShuffleNetV1": shufflenet_v1, "SEResNet": se_resnet, "SENet": senet, "SKResNet": sk_resnet,
    "SKNet": sknet,
}


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
