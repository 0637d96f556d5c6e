"""CPU: the zoo mirrors (convnets_b200/zoo) reproduce the reference's module trees — identical
state_dict keys, shapes and order — and, run as plain torch.nn modules on CPU (un-converted: the
CUDA library has no CPU path), the reference's own logits on the golden weights.  This checks the
host-side construction logic (factories, shape tracker, block wiring) against fixtures generated
from the real reference (tests/golden, oracle/make_golden.py)."""
import glob
import json
import os

import pytest
import torch

from oracle import fill
from oracle import ops as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIXTURES = sorted(p for p in glob.glob(os.path.join(GOLDEN, "*.json"))
                  if not p.endswith("param_counts.json"))


@pytest.mark.parametrize("path", FIXTURES, ids=lambda p: os.path.basename(p)[:-5])
def test_mirror_matches_reference(path):
    from convnets_b200 import zoo as M
    fx = json.load(open(path))
    s = M.Settings(kind=fx["kind"], input_size=(3, fx["hw"], fx["hw"]), num_classes=10,
                   dropout_rate=0.0, batch_size=fx["batch"])
    s.device.set_device("cpu")
    model = M.FAMILIES[fx["family"]](s)
    spec = [(k, list(v.shape)) for k, v in model.state_dict().items()]
    assert spec == [(k, list(sh)) for k, sh in fx["spec"]]
    model.load_state_dict(fill.fill_state_dict(fx["spec"]), strict=True)
    x, y = fill.batch(fx["batch"], fx["hw"])
    model.double().eval()
    with torch.no_grad():
        lg = model(x.double())
    ref = torch.tensor(fx["logits_eval"], dtype=torch.float64).reshape(lg.shape)
    assert O.rel_l2(lg, ref) < 1e-9
    model.train()
    lg = model(x.double())
    ref = torch.tensor(fx["logits_train"], dtype=torch.float64).reshape(lg.shape)
    assert O.rel_l2(lg, ref) < 1e-9
    loss = model.cost_function(lg, y)
    assert abs(float(loss) - fx["loss"]) < 1e-8 * max(1.0, abs(fx["loss"]))


def test_published_parameter_counts_from_mirrors():
    from convnets_b200 import zoo as M
    pc = json.load(open(os.path.join(GOLDEN, "param_counts.json")))
    for name, row in pc.items():
        s = M.Settings(kind=row["kind"], input_size=(3, 32, 32), num_classes=10)
        s.device.set_device("cpu")
        m = M.FAMILIES[row["family"]](s)
        assert sum(p.numel() for p in m.parameters()) == row["published"], name
