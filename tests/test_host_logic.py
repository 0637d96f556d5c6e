"""CPU: host-side logic added in round 2 that needs no device — the bit order of the ReLU mask helper,
the single-use marking of the block forwards, the DenseNet block parameter order, the bench cost model
for the fused entry points."""
import torch
import torch.nn as nn


def test_mask_apply_bit_order():
    from convnets_b200 import functional as F
    n, c, h, w = 2, 32, 3, 5
    g = torch.Generator().manual_seed(1)
    bits = torch.rand(n, h, w, c, generator=g) < 0.5                    # NHWC order, bit j of byte k = element 8k + j
    packed = (bits.view(-1, 8).to(torch.uint8) << torch.arange(8, dtype=torch.uint8)).sum(1).to(torch.uint8)
    t = torch.randn(n, c, h, w, generator=g).contiguous(memory_format=torch.channels_last)
    assert torch.equal(F.mask_apply(t, packed), t * bits.permute(0, 3, 1, 2).float())


def test_single_use_marking_only_towards_plain_conv_blocks():
    import convnets_b200 as cnb
    from convnets_b200 import layers as L
    import sys
    CV = sys.modules["convnets_b200.convert"]
    seq = cnb.convert(nn.Sequential(nn.Conv2d(8, 8, 1), nn.BatchNorm2d(8), nn.ReLU(inplace=True)))
    conv = seq[0]
    assert isinstance(conv, L.Conv2d) and CV._is_conv_block(conv) and CV._is_conv_block(seq)
    assert not CV._is_conv_block(nn.Sequential(nn.Linear(4, 4)))
    t = torch.zeros(1, 8, 2, 2)
    assert not hasattr(CV._sole(t, conv), "_cnb_sole")                  # no BatchNorm holder: nothing to mark
    t._cnb_bn_holder = {"x": t}
    assert not hasattr(CV._sole(t, nn.Identity()), "_cnb_sole")         # consumer is not a convolution
    assert CV._sole(t, seq)._cnb_sole is True


def test_dense_block_parameters_in_layer_order():
    import convnets_b200 as cnb
    from convnets_b200 import dense, zoo as M
    s = M.Settings(kind="121", input_size=(3, 32, 32), num_classes=10)
    m = cnb.convert(M.FAMILIES["DenseNet"](s))
    blk = next(x for x in m.modules() if type(x).__name__ == "DenseBlock")
    ps = dense._block_params(blk)
    names = {id(p): k for k, p in blk.named_parameters()}
    got = [names[id(p)] for p in ps]
    assert got == [k for k, _ in blk.named_parameters()]                # the order autograd returns gradients in
    assert not dense.eligible(blk, torch.zeros(2, 64, 8, 8))            # CPU tensor: the per-layer path


def test_bench_cost_model_for_fused_entry_points():
    import bench
    d = {"M": 1000, "C": 64, "dtype": 1, "flags": 3, "eps": 1e-5, "momentum": 0.1, "num_batches_tracked": 0}
    full = bench.call_cost("cnb_bn_bwd", [d, {"_nulls": [2]}])
    lazy = bench.call_cost("cnb_bn_bwd", [d, {"_nulls": [2, 9]}])
    sums = bench.call_cost("cnb_bn_bwd_from_partials", [d, 3, {"_nulls": [10]}])
    assert full[2] == lazy[2] == sums[2] == "cnb_bn_bwd"
    assert full[1] - lazy[1] == 1000 * 64 * 2 and lazy[1] == sums[1]    # the residual gradient counts only when written
    assert bench.call_cost("cnb_adam_step", [5, {"_nulls": []}])[1] == 0
    assert bench.call_cost("cnb_bn_stats_from_partials", [d, 3, {"_nulls": []}])[1] == 0
