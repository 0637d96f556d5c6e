"""Stand-alone probe of one convolution case through the C ABI (run in a subprocess so that a
faulting or hanging kernel cannot take the whole test session down).

usage: python tests/tc_probe.py <case-json>
prints one JSON line: {"case": ..., "fwd": rel_l2, "dgrad": ..., "wgrad": ..., "uses_tc": [...]}
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from convnets_b200 import functional as F_  # noqa: E402
from convnets_b200 import lib as L  # noqa: E402
from oracle import ops as O  # noqa: E402


def diag(name, got, ref):
    """Error structure hints: where (rows/channels) the mismatch concentrates."""
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu()
    err = (got - ref).abs()
    out = {"rel_l2": O.rel_l2(got, ref), "max_abs": err.max().item(), "ref_max": ref.abs().max().item()}
    if out["rel_l2"] > 2e-2 and got.dim() == 4:
        n, c, h, w = got.shape
        e = err.permute(0, 2, 3, 1).reshape(-1, c)  # [pixels, channels]
        thr = 0.05 * ref.abs().max().item() + 1e-6
        bad = e > thr
        out["bad_frac"] = bad.double().mean().item()
        out["bad_per_chan8"] = [round(bad[:, i::8].double().mean().item(), 3) for i in range(min(8, c))]
        out["bad_per_row8"] = [round(bad[i::8].double().mean().item(), 3) for i in range(8)]
        out["bad_by_chan_block16"] = [round(bad[:, i:i + 16].double().mean().item(), 3) for i in range(0, min(c, 128), 16)]
        out["bad_by_row_block32"] = [round(bad[i:i + 32].double().mean().item(), 3) for i in range(0, min(e.shape[0], 256), 32)]
        out["got_zero_frac"] = (got == 0).double().mean().item()
        out["nan_frac"] = torch.isnan(got).double().mean().item()
    return out


def main():
    cases = json.loads(sys.argv[1])
    for case in (cases if isinstance(cases, list) else [cases]):
        probe(case)


def probe(case):
    torch.manual_seed(case.get("seed", 0))
    dev = torch.device("cuda:0")
    dtype = torch.bfloat16 if case.get("dtype", "bf16") == "bf16" else torch.float32
    N, C, H, W, K, R, S = (case[k] for k in ("N", "C", "H", "W", "K", "R", "S"))
    st, pd, dl, g = case.get("stride", 1), case.get("pad", 0), case.get("dil", 1), case.get("groups", 1)
    algo = {"auto": L.CNB_ALGO_AUTO, "simt": L.CNB_ALGO_SIMT, "tc": L.CNB_ALGO_TC}[case.get("algo", "auto")]
    x = torch.randn(N, C, H, W).to(dtype)
    w = (torch.randn(K, C // g, R, S) / (C // g * R * S) ** 0.5)
    wq = w.to(dtype)  # the operand values the kernel sees
    b = torch.randn(K) if case.get("bias") else None
    res = {"case": case}
    xs, ws = (x.shape, w.shape)
    res["uses_tc"] = [F_.conv2d_uses_tc(xs, ws, st, pd, dl, g, dtype, d) if algo != L.CNB_ALGO_SIMT else False
                      for d in (L.CNB_FWD, L.CNB_DGRAD, L.CNB_WGRAD)]
    xd = x.to(dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    wd = w.to(dev).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    bd = None if b is None else b.to(dev).requires_grad_(True)
    wpack = F_.pack_weight(wd, dtype)
    y = F_.conv2d(xd, wd, bd, st, pd, dl, g, wpack=wpack, algo=algo)
    torch.cuda.synchronize()
    y_ref = O.conv2d(x, wq, b, st, pd, dl, g)
    res["fwd"] = diag("fwd", y, y_ref)
    dy = torch.randn(y_ref.shape).to(dtype)
    y.backward(dy.to(dev).contiguous(memory_format=torch.channels_last))
    torch.cuda.synchronize()
    dx_ref, dw_ref, db_ref = O.conv2d_grads(x, wq, b, dy, st, pd, dl, g)
    res["dgrad"] = diag("dgrad", xd.grad, dx_ref)
    res["wgrad"] = diag("wgrad", wd.grad, dw_ref)
    if b is not None:
        res["bgrad"] = diag("bgrad", bd.grad, db_ref)
    print("PROBE_RESULT " + json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
