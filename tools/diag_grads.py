#!/usr/bin/env python
"""Per-parameter gradient error of a converted zoo model against the fp64 oracle (diagnostic for
tests/test_gpu_models.py failures): prints ours / fp32-floor for every parameter in module order.
usage: python tools/diag_grads.py DenseNet 121 64 8 f32
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from oracle import ops as O
from oracle import zoo as Z


def main():
    family, kind, hw, batch, dt = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
    if kind == "-":
        kind = ""
    dtype = torch.float32 if dt == "f32" else torch.bfloat16
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, hw, hw), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[family](s)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    cnb.convert(model, dtype=dtype)
    s.device.move(model)
    g = torch.Generator().manual_seed(21)
    x = torch.randn(batch, 3, hw, hw, generator=g)
    y = torch.randint(0, 10, (batch,), generator=g)
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    st = {} if dtype == torch.float32 else {"storage": torch.bfloat16}
    lg64, loss64, g64, _ = Z.loss_and_grads(family, sd64, x.double(), y, kind, **st)
    _, _, g32, _ = Z.loss_and_grads(family, sd, x, y, kind, **st)
    model.train()
    out = model(x.cuda())
    loss = model.cost_function(out.float(), y.cuda())
    loss.backward()
    torch.cuda.synchronize()
    print(f"logits rel-L2 {O.rel_l2(out, lg64):.3e}  loss {loss.item():.6f} vs {loss64.item():.6f}")
    for k, p in model.named_parameters():
        ref = g64[k]
        ours = O.rel_l2(p.grad, ref)
        floor = O.rel_l2(g32[k], ref)
        flag = " <<<" if ours > max(1e-3, 8 * floor) else ""
        print(f"{k:60s} {tuple(p.shape)!s:22s} |g| {ref.norm().item():.3e} ours {ours:.3e} floor {floor:.3e}{flag}")


if __name__ == "__main__":
    main()
