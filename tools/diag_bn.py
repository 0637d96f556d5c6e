#!/usr/bin/env python
"""Diagnostic: capture the tensors around one BatchNorm2d of a converted zoo model during a train
step and compare the library's backward with torch's own fp64 BatchNorm+ReLU on the same tensors.
usage: python tools/diag_bn.py DenseNet 121 64 8 f32 features.9
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F


def main():
    family, kind, hw, batch, dt, name = sys.argv[1:7]
    hw, batch = int(hw), int(batch)
    if kind == "-":
        kind = ""
    dtype = torch.float32 if dt == "f32" else torch.bfloat16
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, hw, hw), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[family](s)
    cnb.convert(model, dtype=dtype)
    s.device.move(model)
    g = torch.Generator().manual_seed(21)
    x = torch.randn(batch, 3, hw, hw, generator=g)
    y = torch.randint(0, 10, (batch,), generator=g)
    mod = dict(model.named_modules())[name]
    cap = {}

    def fwd_hook(m, args, out):
        cap["x"] = args[0].detach().clone()
        cap["y"] = out.detach().clone()
        out.register_hook(lambda gr: cap.__setitem__("dy", gr.detach().clone()))
        if args[0].requires_grad:
            args[0].register_hook(lambda gr: cap.__setitem__("dx", gr.detach().clone()))

    mod.register_forward_hook(fwd_hook)
    model.train()
    out = model(x.cuda())
    loss = model.cost_function(out.float(), y.cuda())
    loss.backward()
    torch.cuda.synchronize()
    X, Y, DY = cap["x"], cap["y"], cap["dy"]
    print("x", tuple(X.shape), X.dtype, X.stride(), "dy", tuple(DY.shape), DY.dtype, DY.stride(),
          "relu", mod.fuse_relu)
    xr = X.double().requires_grad_(True)
    w = mod.weight.detach().double().requires_grad_(True)
    b = mod.bias.detach().double().requires_grad_(True)
    yr = F.batch_norm(xr, None, None, w, b, True, 0.1, mod.eps)
    if mod.fuse_relu:
        yr = F.relu(yr)
    yr.backward(DY.double())

    def rel(a, r):
        return ((a.double() - r.double()).norm() / r.double().norm().clamp_min(1e-300)).item()

    print("y     rel", rel(Y, yr))
    print("dbeta rel", rel(mod.bias.grad, b.grad), " dgamma rel", rel(mod.weight.grad, w.grad))
    if "dx" in cap:
        print("dx    rel", rel(cap["dx"], xr.grad), "(dx hook sees the SUM over all consumers of x)")
    from oracle import zoo as Z
    sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    # note: state after the step differs only in BN running stats / counters (no optimizer step ran)
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    st = {} if dtype == torch.float32 else {"storage": torch.bfloat16}
    _, _, g64, _ = Z.loss_and_grads(family, sd64, x.double(), y, kind, **st)
    ob = g64[name + ".bias"].double().cuda()
    print("dbeta vs ORACLE rel", rel(mod.bias.grad, ob), " torch-recomputed vs ORACLE", rel(b.grad, ob))
    eb = (mod.bias.grad.double() - ob).abs()
    print("channels with |err| > 1e-4*max:", int((eb > 1e-4 * ob.abs().max()).sum().item()), "of", eb.numel())
    worst = torch.topk(eb, 8).indices.tolist()
    ymask_ours = (Y > 0)
    ymask_ref = (yr > 0)
    print("mask mismatches:", int((ymask_ours != ymask_ref).sum().item()), "of", Y.numel())
    for c in worst:
        xs = X[:, c].flatten().double()
        print(f"ch {c}: dbeta ours {mod.bias.grad[c].item():.6e} ref {b.grad[c].item():.6e} oracle {ob[c].item():.6e} "
              f"x min {xs.min().item():.4e} max {xs.max().item():.4e} std {xs.std().item():.4e} "
              f"n_distinct {xs.unique().numel()}  y>0 ours {int(ymask_ours[:, c].sum())} ref {int(ymask_ref[:, c].sum())}")
        print("    y ours", [f"{v:.3e}" for v in Y[:, c].flatten().tolist()[:8]])
        print("    y ref ", [f"{v:.3e}" for v in yr[:, c].flatten().tolist()[:8]])


if __name__ == "__main__":
    main()
