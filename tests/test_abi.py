"""CPU: the C-ABI library loads and exports exactly what include/convnets_b200.h declares; the
ctypes table in convnets_b200/lib.py matches the header; descriptor validation (no compute) works
without a GPU; the product path refuses to run without CUDA instead of falling back."""
import ctypes as C
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "convnets_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"\b(cnb_[a-z0-9_]+)\s*\(", src)


@pytest.fixture(scope="module")
def lib():
    from convnets_b200 import build, lib as L
    build.build()
    return L


def test_header_matches_ctypes_table(lib):
    names = _declared()
    assert len(names) == len(set(names)) and len(names) >= 30
    assert set(names) == set(lib.SIGNATURES), set(names) ^ set(lib.SIGNATURES)


def test_shared_library_exports_every_declared_symbol(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib.LIB_PATH], capture_output=True, text=True,
                         check=True).stdout
    exported = set(re.findall(r" T (cnb_[a-z0-9_]+)", out))
    assert set(_declared()) <= exported, set(_declared()) - exported
    h = lib.load()
    for name in _declared():
        assert getattr(h._cdll, name) is not None
    assert "sm_100a" in lib.version()


def test_library_is_sm100a_only(lib):
    out = subprocess.run(["cuobjdump", "-lelf", lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_convolution_object_code_is_tcgen05_and_tma(lib):
    """The dense convolution kernels are Blackwell-native in the OBJECT code: tcgen05.mma (UTCHMMA) with
    TMEM read-back (LDTM), TMA loads in tiled and im2col mode (UTMALDG[.IM2COL]), TMA stores (UTMASTG),
    mbarrier waits (SYNCS) — and no legacy mma.sync (HMMA) contraction."""
    from convnets_b200 import build
    obj = os.path.join(build.BUILD, "conv_tc.o")
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
    ops = re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", sass)
    count = lambda prefix: sum(1 for o in ops if o.startswith(prefix))
    assert count("UTCHMMA") >= 100 and count("LDTM") >= 20
    assert count("UTMALDG") >= 100 and sum(1 for o in ops if o.startswith("UTMALDG") and "IM2COL" in o) >= 20
    assert count("UTMASTG") >= 20 and count("SYNCS") >= 100
    assert count("HMMA") == 0


def test_descriptor_validation_without_a_gpu(lib):
    h = lib.load()
    d = lib.ConvDesc(2, 56, 56, 64, 128, 3, 3, 2, 2, 1, 1, 1, 1, 1, lib.CNB_BF16, 0, lib.CNB_ALGO_AUTO)
    P, Q = C.c_int32(), C.c_int32()
    assert h.cnb_conv2d_out_dims(C.byref(d), C.byref(P), C.byref(Q)) == 0
    assert (P.value, Q.value) == (28, 28)
    assert h.cnb_conv2d_uses_tc(C.byref(d), lib.CNB_FWD) == 1
    d.groups = 32  # grouped (2 -> 4 channels per group): tensor cores on a block-diagonal filter
    assert h.cnb_conv2d_uses_tc(C.byref(d), lib.CNB_FWD) == 1
    # ShuffleNet-v1 g = 4 (68 channels per group): a tile's channel slice would start 136 bytes into a
    # pixel row, which TMA cannot address -> CUDA cores, in every direction that has the problem
    g4 = lib.ConvDesc(8, 4, 4, 272, 136, 1, 1, 1, 1, 0, 0, 1, 1, 4, lib.CNB_BF16, 0, lib.CNB_ALGO_AUTO)
    assert [h.cnb_conv2d_uses_tc(C.byref(g4), k) for k in (lib.CNB_FWD, lib.CNB_DGRAD, lib.CNB_WGRAD)] == [0, 0, 0]
    g3 = lib.ConvDesc(8, 4, 4, 240, 240, 1, 1, 1, 1, 0, 0, 1, 1, 3, lib.CNB_BF16, 0, lib.CNB_ALGO_AUTO)
    assert [h.cnb_conv2d_uses_tc(C.byref(g3), k) for k in (lib.CNB_FWD, lib.CNB_DGRAD, lib.CNB_WGRAD)] == [1, 1, 1]
    d.dtype = lib.CNB_F32  # fp32-IO mode -> CUDA cores
    assert h.cnb_conv2d_uses_tc(C.byref(d), lib.CNB_FWD) == 0
    d.dtype = lib.CNB_BF16
    d.groups = 3   # 64 % 3 != 0 -> invalid, with torch's wording
    assert h.cnb_conv2d_out_dims(C.byref(d), C.byref(P), C.byref(Q)) == -1
    assert "divisible by groups" in lib.last_error()
    d.groups, d.R, d.H = 1, 9, 4  # kernel larger than padded input
    assert h.cnb_conv2d_out_dims(C.byref(d), C.byref(P), C.byref(Q)) == -1
    d2 = lib.ConvDesc(2, 8, 8, 60, 240, 1, 1, 1, 1, 0, 0, 1, 1, 3, lib.CNB_BF16, 0, lib.CNB_ALGO_TC)
    assert h.cnb_conv2d_uses_tc(C.byref(d2), lib.CNB_FWD) == -2  # forced TC with C % 8 != 0
    d3 = lib.ConvDesc(2, 8, 8, 64, 64, 3, 3, 1, 1, 1, 1, 1, 1, 64, lib.CNB_BF16, 0, lib.CNB_ALGO_TC)
    assert h.cnb_conv2d_uses_tc(C.byref(d3), lib.CNB_FWD) == -2  # forced TC on a depthwise conv


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    from convnets_b200 import functional as F_
    x = torch.randn(1, 8, 4, 4)
    w = torch.randn(8, 8, 1, 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        F_.conv2d(x, w)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        F_.relu(x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        F_.max_pool2d(x, 2)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "convnets_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_packed_weight_cache_key_moves_with_every_optimizer_step():
    """torch's fused Adam does not bump `Parameter._version`; the epoch that convnets_b200.layers
    adds to the pack-cache key must advance on every optimizer step of ANY optimizer."""
    from convnets_b200 import layers as L
    p = torch.nn.Parameter(torch.randn(4, 4))
    for kw in ({}, {"fused": True}, {"foreach": True}):
        opt = torch.optim.Adam([p], lr=1e-3, **kw)
        for _ in range(2):
            before = (L._WEIGHT_EPOCH[0], p._version)
            opt.zero_grad()
            (p * p).sum().backward()
            opt.step()
            assert (L._WEIGHT_EPOCH[0], p._version) != before
            assert L._WEIGHT_EPOCH[0] == before[0] + 1
    e = L._WEIGHT_EPOCH[0]
    L.invalidate_packed_weights()
    assert L._WEIGHT_EPOCH[0] == e + 1


def test_batched_pack_table_layout(lib):
    """cnb_pack_filters takes a device array of cnb_pack_item: the ctypes mirror must have the C
    layout, the block ranges must tile [0, total) in order, and convert() must enrol exactly the
    dense layers the tensor-core path takes (the table itself needs no GPU to build)."""
    import convnets_b200 as cnb
    from convnets_b200 import layers as L
    assert C.sizeof(lib.PackItem) == 40            # 3 pointers + 4 int32, no padding
    h = lib.load()
    assert h.cnb_pack_filter_blocks(64, 9, 64) == 9 * 2 * 2
    assert h.cnb_pack_filter_blocks(40, 1, 24) == 1 * 2 * 1
    assert h.cnb_pack_filter_blocks(0, 1, 8) == 0
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 16, 3, padding=1), torch.nn.Conv2d(16, 40, 1),
                              torch.nn.Conv2d(40, 40, 3, padding=1, groups=5),
                              torch.nn.Conv2d(40, 72, 3, padding=1, bias=False))
    cnb.convert(net)
    group = net[1].__dict__["_cnb_group"]
    assert [m is net[1] or m is net[3] for m in group.layers] == [True, True]
    assert "_cnb_group" not in net[0].__dict__ and "_cnb_group" not in net[2].__dict__
    group._build()
    items = (lib.PackItem * 2).from_buffer_copy(bytes(group.table.numpy().tobytes()))
    first = 0
    for it, m, (wb, wt) in zip(items, group.layers, group.bufs):
        k, c, r, s = m.weight.shape
        assert (it.K, it.RS, it.C, it.first_block) == (k, r * s, c, first)
        assert it.w == m.weight.data_ptr() and it.w_bf16 == wb.data_ptr() and it.wt_bf16 == wt.data_ptr()
        assert wb.shape == m.weight.shape and wb.is_contiguous(memory_format=torch.channels_last)
        assert wt.shape == (c, r * s, k)
        first += h.cnb_pack_filter_blocks(k, r * s, c)
    assert group.n_blocks == first
    assert L._BATCHED_PACK is True                 # the default since tests/test_gpu_fused_paths.py
