#!/usr/bin/env python
"""Generate tests/golden/* by running the UNMODIFIED reference (milan-pavlovic-ai/ConvNets).

TEST INFRASTRUCTURE ONLY.  Runs in the build container only (it imports /root/reference through
oracle/refimport.py); the fixtures it writes are committed, the GPU box never runs this.

    python oracle/make_golden.py            # all cases
    python oracle/make_golden.py ResNet     # cases of one family

The reference has no tests or golden vectors of its own.  What pins the oracle:
  1. param_counts.json — the `complexity` column of the reference's data/results.xlsx (nine tuned
     models at 32x32) next to the counts of the models constructed here from the reference classes.
  2. <Family>-<kind>-<hw>.json — for each case the reference module tree is built by the
     reference's own constructor, its state_dict is overwritten with oracle.fill's deterministic
     weights (re-creatable from keys and shapes alone), and on one seeded batch we record
       * state_dict keys and shapes (construction / naming parity),
       * eval-mode logits and train-mode logits + CrossEntropyLoss(sum) in float64,
       * per-parameter gradient norms and the first values of every gradient (float64),
       * BatchNorm running statistics after the training forward,
       * the result of ONE step of the reference's own train_model() in float32 (loss per image,
         accuracy, per-parameter norm of the Adam update).
"""
from __future__ import annotations

import copy
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

from oracle import fill, refimport  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")

# data/results.xlsx (reference), sheet1, column `complexity`: model -> (family, kind, params)
PUBLISHED_PARAMS = {
    "VGGNet-11": ("VGGNet", 11, 28146762),
    "InceptionNet-v1": ("InceptionNetV1", "", 5991082),
    "ResNet-26": ("ResNet", 26, 13966666),
    "SqueezeNet-v1.1": ("SqueezeNet", "1.1", 730580),
    "DenseNet-121": ("DenseNet", "121", 6964106),
    "MobileNet-v1": ("MobileNetV1", "", 3217226),
    "ShuffleNet-v1-g4": ("ShuffleNetV1", "g4", 890234),
    "SEResNet-26": ("SEResNet", 26, 15359306),
    "SKResNet-26": ("SKResNet", 26, 8283978),
}

# (family, kind, hw, batch): small enough for the CPU suite; every family of the zoo
CASES = [
    ("VGGNet", 11, 32, 4),
    ("VGGNet", 16, 32, 4),
    ("InceptionNetV1", "", 32, 4),
    ("ResNet", 18, 32, 4),
    ("ResNet", 50, 64, 4),
    ("ResNeXt", 50, 64, 4),
    ("SqueezeNet", "1.1", 64, 4),
    ("DenseNet", "121", 64, 4),
    ("MobileNetV1", "", 64, 4),
    ("ShuffleNetV1", "g3", 64, 4),
    ("SEResNet", 50, 64, 4),
    ("SENet", 26, 64, 4),
    ("SKResNet", 26, 64, 4),
    ("SKNet", 50, 64, 4),
]

NHEAD = 6  # leading values of each gradient kept in the fixture


def case_name(family, kind, hw):
    return f"{family}-{kind if kind != '' else 'v1'}-{hw}"


def _lst(t):
    return [float(v) for v in t.detach().double().flatten().tolist()]


def run_case(family, kind, hw, batch):
    model = refimport.make_model(family, kind, hw=hw, batch_size=batch, dropout_rate=0.0,
                                 mixed_precision=False)
    spec = fill.spec_of(model.state_dict())
    sd = fill.fill_state_dict(spec)
    model.load_state_dict(sd, strict=True)
    x, y = fill.batch(batch, hw)
    out = {"family": family, "kind": kind, "hw": hw, "batch": batch, "spec": spec,
           "n_params": sum(p.numel() for p in model.parameters()),
           "torch": torch.__version__}

    # ---- float64 execution of the reference module graph
    m64 = copy.deepcopy(model).double()
    m64.eval()
    with torch.no_grad():
        out["logits_eval"] = _lst(m64(x.double()))
    m64.train()
    logits = m64(x.double())
    loss = m64.cost_function(logits, y)           # CrossEntropyLoss(reduction='sum')
    loss.backward()
    out["logits_train"] = _lst(logits)
    out["loss"] = float(loss)
    grads = {}
    for k, p in m64.named_parameters():
        g = p.grad if p.grad is not None else torch.zeros_like(p)
        grads[k] = {"norm": float(g.norm()), "head": _lst(g.flatten()[:NHEAD])}
    out["grads"] = grads
    out["buffers"] = {k: {"norm": float(v.double().norm()), "head": _lst(v.flatten()[:NHEAD])}
                      for k, v in m64.state_dict().items()
                      if k.endswith(("running_mean", "running_var", "num_batches_tracked"))}

    # ---- one step of the reference's own train_model (float32, Adam), basemodel.py:231-302
    model.train()
    model.init_optimizer()
    before = {k: v.detach().clone() for k, v in model.named_parameters()}
    dl = torch.utils.data.DataLoader(torch.utils.data.TensorDataset(x, y), batch_size=batch)
    loss_per_img, acc = model.train_model(dl)
    out["train_model"] = {
        "loss_per_image": float(loss_per_img), "accuracy": float(acc),
        "lr": float(model.setting.learning_rate), "weight_decay": float(model.setting.weight_decay),
        "update_norm": {k: float((p.detach() - before[k]).double().norm())
                        for k, p in model.named_parameters()},
    }
    return out


def param_counts():
    res = {}
    for name, (family, kind, published) in PUBLISHED_PARAMS.items():
        m = refimport.make_model(family, kind, hw=32)
        res[name] = {"family": family, "kind": kind, "published": published,
                     "constructed": sum(p.numel() for p in m.parameters())}
        assert res[name]["constructed"] == published, (name, res[name])
    return res


def main():
    only = set(sys.argv[1:])
    os.makedirs(GOLDEN, exist_ok=True)
    if not only:
        with open(os.path.join(GOLDEN, "param_counts.json"), "w") as f:
            json.dump(param_counts(), f, indent=1)
        print("param_counts.json ok")
    for family, kind, hw, batch in CASES:
        if only and family not in only:
            continue
        res = run_case(family, kind, hw, batch)
        path = os.path.join(GOLDEN, case_name(family, kind, hw) + ".json")
        with open(path, "w") as f:
            json.dump(res, f)
        print(f"{os.path.basename(path)}: {res['n_params']} params, loss {res['loss']:.6f}, "
              f"{os.path.getsize(path) / 1024:.0f} KiB", flush=True)


if __name__ == "__main__":
    main()
