#!/usr/bin/env python
"""GPU check of the batched filter pack (cnb_pack_filters, CNB_BATCHED_PACK=1) before it becomes the
default: every buffer it writes must equal the per-layer cnb_pack_filter output bit for bit, and a
few training steps with it must follow the per-layer trajectory exactly.
usage (GPU box): CNB_BATCHED_PACK=1 python tools/check_batched_pack.py [ResNet 18]
then:            CNB_BATCHED_PACK=1 python bench.py --steps 10 --warmup 4 --no-cpu-baseline
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def main():
    family = sys.argv[1] if len(sys.argv) > 1 else "ResNet"
    kind = sys.argv[2] if len(sys.argv) > 2 else "18"
    import convnets_b200 as cnb
    from convnets_b200 import functional as F_
    from convnets_b200 import layers as L
    from convnets_b200 import zoo as M
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, 64, 64), num_classes=10, dropout_rate=0.0)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[family](s)
    cnb.convert(model)
    s.device.move(model)
    group = next(m.__dict__["_cnb_group"] for m in model.modules() if "_cnb_group" in m.__dict__)
    group.refresh()
    torch.cuda.synchronize()
    bad = 0
    for m, (wb, wt) in zip(group.layers, group.bufs):
        rb, rt = F_.pack_filter(m.weight, transposed=wt is not None)
        ok = torch.equal(wb, rb) and (wt is None or torch.equal(wt, rt))
        bad += not ok
        if not ok:
            print("MISMATCH", tuple(m.weight.shape))
    print(f"{len(group.layers)} filters, {group.n_blocks} blocks, mismatches: {bad}")
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(20):
        group.refresh()
    t1.record()
    torch.cuda.synchronize()
    print(f"batched pack: {t0.elapsed_time(t1) / 20 * 1e3:.1f} us per call")
    t0.record()
    for _ in range(20):
        for m, (wb, wt) in zip(group.layers, group.bufs):
            F_.pack_filter(m.weight, transposed=wt is not None)
    t1.record()
    torch.cuda.synchronize()
    print(f"per-layer packs: {t0.elapsed_time(t1) / 20 * 1e3:.1f} us per model")
    # trajectory: 4 train steps with the group active (this process) must be finite and decrease
    g = torch.Generator().manual_seed(21)
    x = torch.randn(16, 3, 64, 64, generator=g).cuda()
    y = torch.randint(0, 10, (16,), generator=g).cuda()
    model.init_optimizer()
    model.train()
    losses = []
    for _ in range(4):
        model.optimizer.zero_grad(set_to_none=True)
        loss = model.cost_function(model(x).float(), y)
        loss.backward()
        model.optimizer.step()
        losses.append(loss.item())
    print("losses", [round(v, 4) for v in losses], "batched pack active:", L._BATCHED_PACK)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
