#!/usr/bin/env python
"""The reference's own published benchmark, reproduced on B200: `MultiClassBaseModel.test()`
(reference src/basemodel.py:601-722) on the nine tuned models of data/results.xlsx — 32x32x3 input,
batch 256, inference latency per image (`speed`) and throughput (`fps`).  The reference stores
cuda.Event milliseconds as seconds, so its `speed` is ms/image and its
`fps` images/ms; the same units are printed here, next to the published row (unnamed GPU, torch 1.6,
fp16 AMP).  One JSON line per model.   usage: python tools/infer_bench.py [--batches 40] [--eager]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

# data/results.xlsx sheet1: model -> (family, kind, published ms/img, published img/ms)
ROWS = [
    ("VGGNet-11", "VGGNet", "11", 0.054, 18.369), ("InceptionNet-v1", "InceptionNetV1", "", 0.051, 19.620),
    ("ResNet-26", "ResNet", "26", 0.053, 18.835), ("SqueezeNet-v1.1", "SqueezeNet", "1.1", 0.030, 33.421),
    ("DenseNet-121", "DenseNet", "121", 0.116, 8.643), ("MobileNet-v1", "MobileNetV1", "", 0.031, 32.154),
    ("ShuffleNet-v1-g4", "ShuffleNetV1", "g4", 0.051, 19.726), ("SEResNet-26", "SEResNet", "26", 0.061, 16.536),
    ("SKResNet-26", "SKResNet", "26", 0.071, 14.111),
]


class Loader(list):
    pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", type=int, default=40)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--hw", type=int, default=32)
    ap.add_argument("--eager", action="store_true", help="no CUDA graph for the forward")
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    import convnets_b200 as cnb
    from convnets_b200 import lib as L, zoo as M
    L.load()
    g = torch.Generator().manual_seed(21)
    data = Loader([(torch.randn(a.batch, 3, a.hw, a.hw, generator=g).pin_memory(),
                    torch.randint(0, 10, (a.batch,), generator=g).pin_memory()) for _ in range(a.batches)])
    data.dataset = range(a.batch * a.batches)
    for name, fam, kind, pub_speed, pub_fps in ROWS:
        if a.only and a.only not in name:
            continue
        torch.manual_seed(21)
        s = M.Settings(kind=kind, input_size=(3, a.hw, a.hw), num_classes=10, batch_size=a.batch)
        s.device.set_device("cuda:0")
        model = cnb.convert(M.FAMILIES[fam](s))
        s.device.move(model)
        if not a.eager:
            cnb.enable_eval_graph(model)
        n0 = L.launch_count()
        acc, times, fps = model.test(data)
        _, mean, std, _ = model.inference_time(times, a.batch * a.batches)
        print(json.dumps({"model": name, "input": f"{a.hw}x{a.hw}x3", "batch": a.batch, "dtype": "bf16",
                          "ms_per_image": mean, "ms_per_image_std": std, "images_per_ms": fps,
                          "published_ms_per_image": pub_speed, "published_images_per_ms": pub_fps,
                          "published_hardware": "not stated (torch 1.6 / CUDA 10.1 era, fp16 AMP)",
                          "speedup_vs_published": fps / pub_fps,
                          "launch": "eager" if a.eager else "one CUDA graph replay per batch",
                          "eager_launches_seen_by_host": L.launch_count() - n0}), flush=True)
        del model


if __name__ == "__main__":
    main()
