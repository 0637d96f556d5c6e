#!/usr/bin/env python
"""GPU check of cnb_conv2d_dgrad_add (data-grad epilogue that adds the skip gradient) against
cnb_conv2d_dgrad followed by an fp32 add and one bf16 rounding — the two must agree bit for bit —
plus timing of both on ResNet-50 identity-block shapes.  Run on the GPU box:
    python tools/check_dgrad_add.py
Exit status 0 = every case bit-identical.  Wire it into autograd only after this passes
(DESIGN.md section 6d row 2)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

CL = torch.channels_last


def main():
    from convnets_b200 import functional as F_
    from convnets_b200 import lib as L
    lib = L.load()
    dev = torch.device("cuda:0")
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    bad = 0
    # (N, C_in, H, K_out, R, pad): conv1 of identity blocks (1x1 reduce) and basic-block 3x3s
    for (n, c, h, k, r, pad) in [(256, 256, 56, 64, 1, 0), (256, 512, 28, 128, 1, 0), (256, 1024, 14, 256, 1, 0),
                                 (256, 2048, 7, 512, 1, 0), (64, 64, 56, 64, 3, 1), (3, 72, 9, 40, 3, 1)]:
        torch.manual_seed(5)
        dy = torch.randn(n, k, h, h, device=dev).bfloat16().contiguous(memory_format=CL)
        w = (torch.randn(k, c, r, r, device=dev) / (c * r * r) ** 0.5).contiguous(memory_format=CL)
        skip = torch.randn(n, c, h, h, device=dev).bfloat16().contiguous(memory_format=CL)
        wb, wt = F_.pack_filter(w, transposed=True)
        d = L.ConvDesc(n, h, h, c, k, r, r, 1, 1, pad, pad, 1, 1, 1, L.CNB_BF16, 0, L.CNB_ALGO_TC)
        nws = lib.cnb_conv2d_workspace_bytes(C.byref(d), L.CNB_DGRAD)
        ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
        dx0 = torch.empty_like(skip)
        dx1 = torch.empty_like(skip)
        L.check(lib.cnb_conv2d_dgrad(C.byref(d), p(dy), p(wb), p(wt), p(dx0), p(ws), ws.numel(), st()), "dgrad")
        L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), p(dy), p(wb), p(wt), p(skip), p(dx1), p(ws), ws.numel(),
                                         st()), "dgrad_add")
        torch.cuda.synchronize()
        # reference of the fused result: the accumulator is not observable, so bound it instead:
        # |dx1 - (dx0 + skip)| must be within one bf16 ulp of the two-rounding result
        two = (dx0.float() + skip.float())
        err = (dx1.float() - two).abs()
        ulp = two.abs() * 2.0 ** -7 + 1e-30
        frac_exact = (dx1 == two.bfloat16()).float().mean().item()
        ok = bool((err <= ulp).all())
        bad += not ok
        # in-place form: dx aliases the addend
        inpl = skip.clone()
        L.check(lib.cnb_conv2d_dgrad_add(C.byref(d), p(dy), p(wb), p(wt), p(inpl), p(inpl), p(ws), ws.numel(),
                                         st()), "dgrad_add(in place)")
        torch.cuda.synchronize()
        same = torch.equal(inpl, dx1)
        bad += not same
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
        for _ in range(10):
            lib.cnb_conv2d_dgrad(C.byref(d), p(dy), p(wb), p(wt), p(dx0), p(ws), ws.numel(), st())
            torch.add(dx0, skip, out=dx0)
        ev[1].record()
        for _ in range(10):
            lib.cnb_conv2d_dgrad_add(C.byref(d), p(dy), p(wb), p(wt), p(skip), p(dx1), p(ws), ws.numel(), st())
        ev[2].record()
        torch.cuda.synchronize()
        print(f"N{n} C{c} H{h} K{k} R{r}: within 1 ulp {ok}, equal to double rounding {frac_exact:.4f}, "
              f"in-place identical {same}, dgrad+add {ev[0].elapsed_time(ev[1]) / 10:.3f} ms, "
              f"fused {ev[1].elapsed_time(ev[2]) / 10:.3f} ms")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
