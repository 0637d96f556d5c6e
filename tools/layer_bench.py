#!/usr/bin/env python
"""Per-layer micro-benchmark of the C-ABI kernels on ResNet-50 (B=256, 224x224) signatures.

Times each entry point with CUDA events (L2 flushed between repetitions) and prints achieved
TFLOP/s (convs) or GB/s of algorithmic bytes (BatchNorm).  Short enough to be the target of
`ncu --set full -k regex:...` captures (tools/ is measurement tooling, not part of the product).
"""
import argparse
import ctypes as C
import json
import sys, os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from convnets_b200 import functional as F_, lib as L

CL = torch.channels_last
# (C, H, K, R, stride)  — ResNet-50 @224, per-GPU batch 256
CONVS = [
    (64, 56, 64, 1, 1), (64, 56, 64, 3, 1), (64, 56, 256, 1, 1), (256, 56, 64, 1, 1),
    (256, 56, 128, 1, 1), (128, 56, 128, 3, 2), (128, 28, 512, 1, 1), (512, 28, 128, 1, 1),
    (128, 28, 128, 3, 1), (256, 14, 256, 3, 1), (256, 14, 1024, 1, 1), (1024, 14, 256, 1, 1),
    (512, 7, 512, 3, 1), (512, 7, 2048, 1, 1), (2048, 7, 512, 1, 1), (256, 56, 512, 1, 2),
]
BNS = [(64, 112), (64, 56), (256, 56), (128, 28), (512, 28), (256, 14), (1024, 14), (512, 7), (2048, 7)]


def flush(buf):
    buf.zero_()


def timeit(fn, reps, buf):
    ms = []
    for _ in range(reps):
        flush(buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return min(ms), sorted(ms)[len(ms) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="", help="comma list: conv,bn")
    ap.add_argument("--convs", default="", help="indices into CONVS, comma separated")
    ap.add_argument("--bns", default="", help="indices into BNS")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    L.load()
    N = a.batch
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    rows = []
    only = set(a.only.split(",")) if a.only else {"conv", "bn"}
    convs = [CONVS[int(i)] for i in a.convs.split(",")] if a.convs else CONVS
    bns = [BNS[int(i)] for i in a.bns.split(",")] if a.bns else BNS
    if "conv" in only:
        for (c, h, k, r, s) in convs:
            pad = r // 2
            x = torch.randn(N, c, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=CL).requires_grad_(True)
            w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).contiguous(memory_format=CL).requires_grad_(True)
            wp = F_.pack_weight(w, torch.bfloat16)
            y = F_.conv2d(x, w, None, s, pad, 1, 1, wpack=wp)
            dy = torch.randn_like(y)
            P = y.shape[2]
            fl = 2.0 * N * P * P * k * c * r * r
            lib = L.load()
            d = L.ConvDesc(N, h, h, c, k, r, r, s, s, pad, pad, 1, 1, 1, L.CNB_BF16, 0, 0)
            ws = torch.empty(max(lib.cnb_conv2d_workspace_bytes(C.byref(d), i) for i in range(3)) + 1024, dtype=torch.uint8, device=dev)
            dx = torch.empty_like(x); dw = torch.empty_like(w)
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            p = lambda t: C.c_void_p(t.data_ptr())
            fns = {
                "fwd": lambda: L.check(lib.cnb_conv2d_fwd(C.byref(d), p(x), p(wp), None, p(y), None, p(ws), ws.numel(), st), "fwd"),
                "dgrad": lambda: L.check(lib.cnb_conv2d_dgrad(C.byref(d), p(dy), p(wp), None, p(dx), p(ws), ws.numel(), st), "dgrad"),
                "wgrad": lambda: L.check(lib.cnb_conv2d_wgrad(C.byref(d), p(x), p(dy), p(dw), None, p(ws), ws.numel(), st), "wgrad"),
            }
            byts = (x.numel() + y.numel() + w.numel()) * 2
            for name, fn in fns.items():
                fn(); torch.cuda.synchronize()
                best, med = timeit(fn, a.reps, buf)
                rows.append(dict(op=f"conv_{name}", C=c, H=h, K=k, R=r, stride=s, ms=best, ms_med=med,
                                 tflops=fl / best / 1e9, min_ms_hbm=byts / 6.5e9, min_ms_tc=fl / 1.37e12))
                print(f"conv {name:5s} C{c:4d} H{h:3d} K{k:4d} R{r} s{s}: {best:7.3f} ms {fl / best / 1e9:7.1f} TF/s "
                      f"(hbm floor {byts / 6.5e9:.3f} ms, tc floor {fl / 1.37e12:.3f} ms)", flush=True)
            del x, w, y, dy, dx, dw, ws
    if "bn" in only:
        for (c, h) in bns:
            x = torch.randn(N, c, h, h, device=dev).to(torch.bfloat16).contiguous(memory_format=CL).requires_grad_(True)
            g = torch.ones(c, device=dev, requires_grad=True); b = torch.zeros(c, device=dev, requires_grad=True)
            rm = torch.zeros(c, device=dev); rv = torch.ones(c, device=dev)
            res = torch.randn_like(x)
            for relu, add in ((True, False), (True, True)):
                holder = {}

                def fwd():
                    holder["y"] = F_.batch_norm(x, g, b, rm, rv, True, 0.1, 1e-5, relu=relu, residual=res if add else None)

                fwd(); dy = torch.randn_like(holder["y"]); torch.cuda.synchronize()
                best, med = timeit(fwd, a.reps, buf)
                A = x.numel() * 2
                nf = 2 + (1 if add else 0)
                print(f"bn fwd  C{c:4d} H{h:3d} relu{int(relu)} add{int(add)}: {best:7.3f} ms  {nf * A / best / 1e6:7.0f} GB/s algorithmic", flush=True)
                rows.append(dict(op="bn_fwd", C=c, H=h, relu=relu, add=add, ms=best, gbs=nf * A / best / 1e6))

                def bwd():
                    x.grad = None
                    holder["y"].backward(dy, retain_graph=True)

                bwd(); torch.cuda.synchronize()
                best, med = timeit(bwd, a.reps, buf)
                nb = 3 + (1 if add else 0)
                print(f"bn bwd  C{c:4d} H{h:3d} relu{int(relu)} add{int(add)}: {best:7.3f} ms  {nb * A / best / 1e6:7.0f} GB/s algorithmic", flush=True)
                rows.append(dict(op="bn_bwd", C=c, H=h, relu=relu, add=add, ms=best, gbs=nb * A / best / 1e6))
            del x, res
    if a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
