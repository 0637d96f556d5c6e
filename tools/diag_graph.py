#!/usr/bin/env python
"""Diagnostic: eager vs eager (determinism control) vs StepGraph trajectories at a small learning rate."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import convnets_b200 as cnb
from convnets_b200 import zoo as M

def make(family, kind, lr):
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, 32, 32), num_classes=10, dropout_rate=0.0)
    s.learning_rate = lr
    s.device.set_device("cuda:0")
    m = M.FAMILIES[family](s)
    cnb.convert(m, dtype=torch.bfloat16)
    s.device.move(m)
    m.init_optimizer()
    m.train()
    return m

def main():
    family, kind, lr = sys.argv[1], (sys.argv[2] if sys.argv[2] != "-" else ""), float(sys.argv[3])
    g = torch.Generator().manual_seed(3)
    batches = [(torch.randn(8, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (8,), generator=g).cuda()) for _ in range(7)]
    a, b, c = make(family, kind, lr), make(family, kind, lr), make(family, kind, lr)
    b.load_state_dict(a.state_dict()); c.load_state_dict(a.state_dict())
    sg = cnb.StepGraph(c, eager_steps=3)
    def eager(m, x, y):
        m.optimizer.zero_grad(); out = m(x); l = m.cost_function(out.float(), y); l.backward(); m.optimizer.step(); return l
    for i, (x, y) in enumerate(batches):
        la, lb = eager(a, x, y).item(), eager(b, x, y).item()
        lc = sg(x, y)[0].item()
        print(f"step {i}: eagerA {la:.6f} eagerB {lb:.6f} graph {lc:.6f}", flush=True)
    def dist(m1, m2):
        num = sum(((p.detach().double() - q.detach().double()) ** 2).sum().item() for p, q in zip(m1.parameters(), m2.parameters()))
        den = sum((p.detach().double() ** 2).sum().item() for p in m1.parameters())
        return (num / den) ** 0.5
    print("param rel dist eagerA-eagerB", dist(a, b), " eagerA-graph", dist(a, c))
    for (k, p), (_, q) in zip(a.named_buffers(), c.named_buffers()):
        if p.dtype.is_floating_point:
            d = (p.double() - q.double()).norm().item() / max(p.double().norm().item(), 1e-30)
            if d > 1e-3: print("buffer", k, d)

if __name__ == "__main__":
    main()
