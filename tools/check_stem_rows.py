#!/usr/bin/env python
"""GPU check of the row-im2col stem (functional.conv2d_input_rows, CNB_STEM_ROWS=1) against the
validated full-im2col stem (functional.conv2d_input): outputs, BatchNorm partial sums and weight
gradients on the stems of the zoo, then their timing at batch 256.  Exit status 0 = all within 1e-2
rel-L2 of each other (both are bf16 tensor-core paths with different summation orders)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

CL = torch.channels_last


def rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-300)).item()


def main():
    from convnets_b200 import functional as F_
    dev = torch.device("cuda:0")
    bad = 0
    # (N, C, H, W, K, R, stride, pad, bias): ResNet, SqueezeNet 1.0, VGG / MobileNet, Inception stems
    cases = [(8, 3, 64, 64, 64, 7, 2, 3, False), (4, 3, 61, 59, 96, 7, 2, 0, True),
             (6, 3, 32, 32, 64, 3, 1, 1, True), (5, 3, 33, 33, 32, 3, 2, 1, False),
             (2, 1, 20, 20, 16, 5, 1, 2, False)]
    for (n, c, h, w, k, r, st, pd, hb) in cases:
        torch.manual_seed(7)
        x = torch.randn(n, c, h, w, device=dev)
        wt = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).contiguous(memory_format=CL)
        b = torch.randn(k, device=dev) if hb else None
        outs = []
        for fn in (F_.conv2d_input, F_.conv2d_input_rows):
            wd = wt.clone().requires_grad_(True)
            bd = None if b is None else b.clone().requires_grad_(True)
            y = fn(x, wd, bd, st, pd, 1, torch.bfloat16, bn_stats=not hb)
            part = None
            if isinstance(y, tuple):
                y, part = y
            torch.manual_seed(8)
            dy = torch.randn(y.shape, device=dev).bfloat16().contiguous(memory_format=CL)
            y.backward(dy)
            torch.cuda.synchronize()
            outs.append((y.detach(), None if part is None else part.sum(0), wd.grad, None if bd is None else bd.grad))
        (y0, p0, g0, b0), (y1, p1, g1, b1) = outs
        errs = {"y": rel(y1, y0), "dw": rel(g1, g0)}
        if p0 is not None and p1 is not None:
            errs["bn sums"] = rel(p1, p0)
        if b0 is not None:
            errs["db"] = rel(b1, b0)
        ok = all(v <= 1e-2 for v in errs.values()) and y0.shape == y1.shape and g0.shape == g1.shape
        bad += not ok
        print((n, c, h, w, k, r, st, pd), {k_: f"{v:.2e}" for k_, v in errs.items()}, "OK" if ok else "MISMATCH")
    # timing, ResNet stem at batch 256
    x = torch.randn(256, 3, 224, 224, device=dev)
    wt = (torch.randn(64, 3, 7, 7, device=dev) / 147 ** 0.5).contiguous(memory_format=CL).requires_grad_(True)
    for name, fn in (("full im2col", F_.conv2d_input), ("row im2col", F_.conv2d_input_rows)):
        for it in range(3):
            if it == 1:
                torch.cuda.synchronize()
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record()
            y, _ = fn(x, wt, None, 2, 3, 1, torch.bfloat16, bn_stats=True)
            y.backward(torch.ones_like(y))
        e1.record()
        torch.cuda.synchronize()
        print(f"{name}: {e0.elapsed_time(e1) / 2:.3f} ms per fwd+wgrad (includes the ones_like fill)")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
