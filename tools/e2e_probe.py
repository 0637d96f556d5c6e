#!/usr/bin/env python
"""Where the time of one end-to-end step goes on the host: the train_model loop of bench.py's e2e leg
with wall-clock stamps around (next batch from the PrefetchLoader, graph call, loss.item(), score).
usage: python tools/e2e_probe.py resnext50 [--steps 8]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("model")
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--hw", type=int, default=224)
    a = ap.parse_args()
    import convnets_b200 as cnb
    from convnets_b200 import zoo as M
    from convnets_b200.loader import PrefetchLoader
    cls, kind, B = bench.WORKLOADS[a.model]
    torch.manual_seed(21)
    s = M.Settings(kind=kind, input_size=(3, a.hw, a.hw), num_classes=10)
    s.device.set_device("cuda:0")
    model = M.FAMILIES[cls](s)
    cnb.convert(model)
    s.device.move(model)
    model.init_optimizer()
    model.train()
    g = torch.Generator().manual_seed(3)
    host = [(torch.randint(0, 256, (B, a.hw, a.hw, 3), generator=g, dtype=torch.uint8).pin_memory(),
             torch.randint(0, 10, (B,), generator=g).pin_memory()) for _ in range(2)]
    sg = cnb.enable_step_graph(model, eager_steps=3)
    dev = torch.device("cuda:0")
    norm = dict(normalize=(bench.CINIC_MEAN, bench.CINIC_STD), nhwc=True)
    for rnd in range(2):
        loader = PrefetchLoader([host[i % 2] for i in range(a.steps)], dev, **norm)
        torch.cuda.synchronize()
        t_prev = time.perf_counter()
        rows = []
        it = iter(loader)
        while True:
            t0 = time.perf_counter()
            try:
                x, y = next(it)
            except StopIteration:
                break
            t1 = time.perf_counter()
            loss, out = sg(x, y)
            t2 = time.perf_counter()
            lv = loss.item()
            t3 = time.perf_counter()
            sc = model.score(out, y)
            t4 = time.perf_counter()
            rows.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3))
        torch.cuda.synchronize()
        print(f"round {rnd}: graph captured = {sg.graph is not None}")
        for i, r in enumerate(rows):
            print(f"  step {i}: next(loader) {r[0]*1e3:7.2f} ms  graph call {r[1]*1e3:7.2f} ms  "
                  f"loss.item {r[2]*1e3:7.2f} ms  score {r[3]*1e3:7.2f} ms")
    st = torch.cuda.memory_stats()
    print("cudaMalloc retries", st.get("num_alloc_retries"), "segments", st.get("segment.all.allocated"),
          "reserved GB", torch.cuda.memory_reserved() / 2**30)


if __name__ == "__main__":
    main()
