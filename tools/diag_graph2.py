#!/usr/bin/env python
"""Diagnostic: first module whose forward output differs between an eager step and the graph replay."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import convnets_b200 as cnb
from convnets_b200 import zoo as M
from tools.diag_graph import make

def main():
    family, kind = sys.argv[1], (sys.argv[2] if sys.argv[2] != "-" else "")
    g = torch.Generator().manual_seed(3)
    batches = [(torch.randn(8, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (8,), generator=g).cuda()) for _ in range(5)]
    a, c = make(family, kind, 1e-4), make(family, kind, 1e-4)
    c.load_state_dict(a.state_dict())
    sg = cnb.StepGraph(c, eager_steps=3)
    capA, capC = {}, {}
    def hook(store):
        def mk(name):
            def h(mod, args, out):
                if torch.is_tensor(out):
                    store[name] = out if store is capC else out.detach().clone()
                    if store is capA and torch.is_tensor(args[0]):
                        store[name + "#in"] = args[0].detach().clone()
                    if store is capC and torch.is_tensor(args[0]):
                        store[name + "#in"] = args[0]
            return h
        return mk
    def eager(m, x, y):
        m.optimizer.zero_grad(); out = m(x); l = m.cost_function(out.float(), y); l.backward(); m.optimizer.step(); return l
    for i, (x, y) in enumerate(batches[:4]):
        if i == 3:
            for n, mod in a.named_modules():
                if not list(mod.children()): mod.register_forward_hook(hook(capA)(n))
            for n, mod in c.named_modules():
                if not list(mod.children()): mod.register_forward_hook(hook(capC)(n))
        if i == 3:
            num = sum(((p.detach().double() - q.detach().double()) ** 2).sum().item() for p, q in zip(a.parameters(), c.parameters()))
            print("param sq dist before step 3:", num)
        la = eager(a, x, y).item()
        lc = sg(x, y)[0].item()
        print(f"step {i}: eager {la:.6f} graph {lc:.6f}")
    torch.cuda.synchronize()
    # packed operand copies after the step (eager cache is from this step's forward; graph's too)
    ma, mc = dict(a.named_modules()), dict(c.named_modules())
    for n in ("features.2.res_block.0.0", "features.2.res_block.1.0"):
        pa, pc = ma[n].__dict__.get("_cnb_pack"), mc[n].__dict__.get("_cnb_pack")
        print(n, "pack keys", pa[0], pc[0])
        print("   wb diff", (pa[1][0].float() - pc[1][0].float()).abs().max().item(),
              " weight diff now", (ma[n].weight - mc[n].weight).abs().max().item(),
              " wb vs own weight (graph)", (pc[1][0].float() - mc[n].weight.detach().float()).abs().max().item(),
              " (eager)", (pa[1][0].float() - ma[n].weight.detach().float()).abs().max().item())
    for n in capA:
        if n not in capC: continue
        pa, pc = capA[n].double(), capC[n].double()
        d = (pa - pc).norm().item() / max(pa.norm().item(), 1e-30)
        flag = " <<<" if d > 1e-6 else ""
        print(f"{n:50s} {tuple(pa.shape)!s:20s} rel diff {d:.3e}{flag}")

if __name__ == "__main__":
    main()
