"""Adam: the oracle restatement against torch.optim.Adam on CPU (the reference's optimiser,
src/basemodel.py:58-66), and — on the GPU — cnb_adam_step / optim.FusedAdam against the oracle,
with the bf16 operand copies it writes checked bit for bit against cnb_pack_filter."""
import numpy as np
import pytest
import torch

from oracle import ops as O


def _ulps(a, b):
    a = np.asarray(a, dtype=np.float32).view(np.int32).astype(np.int64)
    b = np.asarray(b, dtype=np.float32).view(np.int32).astype(np.int64)
    return int(np.abs(a - b).max())


@pytest.mark.parametrize("wd", [0.0, 1e-4])
def test_oracle_adam_matches_torch_cpu(wd):
    g = torch.Generator().manual_seed(4)
    p0 = torch.randn(37, 19, generator=g)
    grads = [torch.randn(37, 19, generator=g) * (0.1 + i) for i in range(6)]
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([p], lr=3e-3, weight_decay=wd)
    pn, m, v = p0.numpy().copy(), np.zeros_like(p0.numpy()), np.zeros_like(p0.numpy())
    for t, gr in enumerate(grads, start=1):
        p.grad = gr.clone()
        opt.step()
        pn, m, v = O.adam_step(pn, gr.numpy(), m, v, t, lr=3e-3, weight_decay=wd)
        st = opt.state[p]
        # same operations, same order, same fused multiply-adds: the moments agree bit for bit,
        # the parameter to one unit in the last place (the vectorised sqrt / divide)
        assert _ulps(st["exp_avg"].numpy(), m) == 0
        assert _ulps(st["exp_avg_sq"].numpy(), v) == 0
        assert _ulps(p.detach().numpy(), pn) <= 1
        pn = p.detach().numpy().copy()


@pytest.mark.gpu
@pytest.mark.parametrize("wd", [0.0, 1e-4])
def test_fused_adam_matches_oracle_and_packs(wd):
    import convnets_b200  # noqa: F401
    from convnets_b200 import functional as F
    from convnets_b200 import layers as L
    from convnets_b200.optim import FusedAdam
    torch.manual_seed(2)
    convs = [L.Conv2d(64, 128, 3, padding=1, bias=False), L.Conv2d(128, 72, 1, bias=True),
             L.Conv2d(8, 40, 5, padding=2, bias=False)]
    model = torch.nn.Sequential(*convs, torch.nn.BatchNorm2d(40), torch.nn.Linear(40, 10)).cuda()
    for m in convs:
        m.weight.data = m.weight.data.contiguous(memory_format=torch.channels_last)
        m.compute_dtype = torch.bfloat16
    grp = L.PackGroup(convs)
    for m in convs:
        m.__dict__["_cnb_group"] = grp
    params = list(model.parameters())
    opt = FusedAdam(params, lr=2e-3, weight_decay=wd, model=model)
    host = [(p.detach().cpu().numpy().copy(), np.zeros(p.shape, np.float32), np.zeros(p.shape, np.float32))
            for p in params]
    g = torch.Generator().manual_seed(8)
    for t in range(1, 5):
        grads = [torch.randn(p.shape, generator=g) * 0.3 for p in params]
        for p, gr in zip(params, grads):
            fmt = torch.channels_last if p.dim() == 4 else torch.contiguous_format
            p.grad = gr.cuda().contiguous(memory_format=fmt)
        opt.step()
        host = [O.adam_step(h[0], gr.numpy(), h[1], h[2], t, lr=2e-3, weight_decay=wd)
                for h, gr in zip(host, grads)]
        for p, h in zip(params, host):
            st = opt.state[p]
            assert _ulps(st["exp_avg"].cpu().numpy(), h[1]) <= 2
            assert _ulps(st["exp_avg_sq"].cpu().numpy(), h[2]) <= 2
            assert np.abs(p.detach().cpu().numpy() - h[0]).max() <= 2e-6 * max(1.0, float(np.abs(h[0]).max()))
    assert opt.step_count() == 4
    # the operand copies written by the step == a fresh pack of the updated master weights
    for m, (wb, wt) in zip(convs, grp.bufs):
        ref_b, ref_t = F.pack_filter(m.weight, transposed=True)
        assert torch.equal(wb, ref_b)
        assert torch.equal(wt.view(-1), ref_t.view(-1))
        # and the layers' caches are valid under the current weight epoch (no re-pack on the next forward)
        assert m._packed(torch.bfloat16)[0] is wb
    sd = opt.state_dict()
    assert float(sd["state"][0]["step"]) == 4.0


@pytest.mark.gpu
def test_fused_adam_follows_lr_changes_and_state_dict_roundtrip():
    from convnets_b200.optim import FusedAdam
    torch.manual_seed(5)
    p = torch.nn.Parameter(torch.randn(1000, device="cuda"))
    q = torch.nn.Parameter(p.detach().clone())
    a = FusedAdam([p], lr=1e-2)
    b = torch.optim.Adam([q], lr=1e-2)
    for t in range(6):
        if t == 3:
            for o in (a, b):
                o.param_groups[0]["lr"] = 1e-3     # what ReduceLROnPlateau does
            a2 = FusedAdam([p], lr=1e-2)
            a2.load_state_dict(a.state_dict())      # resume from a checkpoint
            a = a2
            a.param_groups[0]["lr"] = 1e-3
        gr = torch.randn(1000, device="cuda")
        p.grad, q.grad = gr.clone(), gr.clone()
        a.step()
        b.step()
    assert torch.allclose(p, q, rtol=0, atol=2e-6)
    assert a.step_count() == 6
