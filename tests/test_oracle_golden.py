"""CPU: pin the oracle (oracle/zoo.py, oracle/ops.py) to the real reference through the committed
golden fixtures (tests/golden/*.json, written by oracle/make_golden.py from /root/reference).

The reference has no tests of its own; these fixtures are outputs of the reference
itself on deterministic weights and inputs (oracle/fill.py), so a pass here means the functional
restatement reproduces the reference's module graph: logits (eval and train mode), the summed
cross-entropy loss, every parameter gradient, the BatchNorm running statistics, and one Adam step
of the reference's train_model.
"""
import glob
import json
import os

import numpy as np
import pytest
import torch

from oracle import fill
from oracle import ops as O
from oracle import zoo as Z

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIXTURES = sorted(p for p in glob.glob(os.path.join(GOLDEN, "*.json"))
                  if not p.endswith("param_counts.json"))


def _load(path):
    with open(path) as f:
        return json.load(f)


def test_all_families_have_a_fixture_and_an_oracle():
    fams = {_load(p)["family"] for p in FIXTURES}
    assert fams == set(Z.MODELS), (sorted(fams), sorted(Z.MODELS))


def test_published_parameter_counts():
    """data/results.xlsx `complexity` column == parameters of the state_dict specs we rebuild."""
    pc = _load(os.path.join(GOLDEN, "param_counts.json"))
    assert len(pc) == 9
    for name, row in pc.items():
        assert row["published"] == row["constructed"], name
    assert pc["VGGNet-11"]["published"] == 28146762
    assert pc["ResNet-26"]["published"] == 13966666


@pytest.mark.parametrize("path", FIXTURES, ids=lambda p: os.path.basename(p)[:-5])
def test_oracle_reproduces_reference(path):
    fx = _load(path)
    family, kind = fx["family"], fx["kind"]
    if family not in Z.MODELS:
        pytest.fail(f"oracle/zoo.py has no restatement of {family}")
    sd = fill.fill_state_dict(fx["spec"])
    n_params = sum(v.numel() for k, v in sd.items() if Z.is_param(k))
    assert n_params == fx["n_params"]
    x, y = fill.batch(fx["batch"], fx["hw"])
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}

    # eval-mode forward (running statistics)
    lg_eval, _ = Z.forward(family, sd64, x.double(), kind, training=False)
    ref = torch.tensor(fx["logits_eval"], dtype=torch.float64).reshape(lg_eval.shape)
    assert O.rel_l2(lg_eval, ref) < 1e-9

    # training forward + backward in float64
    logits, loss, grads, newbuf = Z.loss_and_grads(family, sd64, x.double(), y, kind)
    ref = torch.tensor(fx["logits_train"], dtype=torch.float64).reshape(logits.shape)
    assert O.rel_l2(logits, ref) < 1e-9
    assert abs(float(loss) - fx["loss"]) < 1e-8 * max(1.0, abs(fx["loss"]))
    gscale = max(v["norm"] for v in fx["grads"].values())
    assert set(grads) == set(fx["grads"])
    for k, g in grads.items():
        g = torch.zeros_like(sd64[k]) if g is None else g
        want = fx["grads"][k]
        assert abs(float(g.norm()) - want["norm"]) <= 1e-7 * gscale, k
        head = g.flatten()[:len(want["head"])]
        assert np.allclose(head.numpy(), np.array(want["head"]), rtol=1e-6, atol=1e-7 * gscale), k
    for k, want in fx["buffers"].items():
        v = newbuf[k].double()
        assert abs(float(v.norm()) - want["norm"]) <= 1e-9 * max(1.0, want["norm"]), k
        assert np.allclose(v.flatten()[:len(want["head"])].numpy(), np.array(want["head"]),
                           rtol=1e-9, atol=1e-12), k

    # one Adam step of train_model, float32 like the reference's own run
    tm = fx["train_model"]
    new_sd, lg32, loss32, _ = Z.train_step(family, sd, x, y, kind, {}, lr=tm["lr"],
                                           weight_decay=tm["weight_decay"])
    assert abs(float(loss32) / fx["batch"] - tm["loss_per_image"]) <= 1e-4 * max(1.0, tm["loss_per_image"])
    acc = float((lg32.argmax(1) == y).sum()) / fx["batch"]
    assert acc == tm["accuracy"]
    for k, want in tm["update_norm"].items():
        got = float((new_sd[k] - sd[k]).double().norm())
        # Adam's first step moves every element by ~lr regardless of gradient scale; elements
        # whose gradient is numerically zero may flip sign between two fp32 executions
        assert abs(got - want) <= 2e-2 * max(want, 1e-12) + 1e-9, (k, got, want)
