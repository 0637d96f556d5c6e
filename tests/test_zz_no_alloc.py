"""cnb_bn_fwd_train_static — the allocation-free BatchNorm training forward — against cnb_bn_fwd_train.

No counterpart in the reference: the entry point exists for callers that cannot call cudaMalloc (CUDA
graph capture, serving loops); its scratch is a static __device__ array inside the library.  The file
name sorts last on purpose: a failure here must not hide another
test.  The test calls through the C ABI only, twice in a row on different shapes (the second call
finds the first call's partial sums and coefficients in the shared array), and checks that the
entry point leaves torch's allocator untouched (the library itself never calls torch; that it
makes no cudaMalloc is by construction — see the entry point in csrc/bn.cu).
"""
import ctypes as C

import pytest
import torch

from convnets_b200 import lib as L

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2      # the bf16 bar of tests/test_gpu_ops.py; the two entries run the same kernels in
F32_TOL = 1e-5       # the same launch geometry, so the results are expected to be identical
CNB_ERR_WORKSPACE = -4   # include/convnets_b200.h


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _case(m, c, dtype, relu, add, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = (torch.randn(m, c, device="cuda", generator=g) * 1.7 + 0.3).to(dtype)
    res = torch.randn(m, c, device="cuda", generator=g).to(dtype) if add else None
    gamma = torch.rand(c, device="cuda", generator=g) + 0.5
    beta = torch.randn(c, device="cuda", generator=g) * 0.1
    return x, res, gamma, beta


def _outputs(m, c, dtype, with_mask):
    return dict(
        y=torch.empty(m, c, dtype=dtype, device="cuda"),
        mean=torch.empty(c, dtype=torch.float32, device="cuda"),
        rstd=torch.empty(c, dtype=torch.float32, device="cuda"),
        rmean=torch.zeros(c, dtype=torch.float32, device="cuda"),
        rvar=torch.ones(c, dtype=torch.float32, device="cuda"),
        nbt=torch.zeros((), dtype=torch.int64, device="cuda"),
        mask=torch.zeros(m * c // 8, dtype=torch.uint8, device="cuda") if with_mask else None,
    )


def _run_pair(lib, m, c, dtype, relu, add, seed):
    x, res, gamma, beta = _case(m, c, dtype, relu, add, seed)
    with_mask = relu and dtype == torch.bfloat16 and c % 8 == 0
    flags = (L.CNB_BN_RELU if relu else 0) | (L.CNB_BN_ADD if add else 0)
    dt = L.CNB_BF16 if dtype == torch.bfloat16 else L.CNB_F32
    a, b = _outputs(m, c, dtype, with_mask), _outputs(m, c, dtype, with_mask)

    # existing entry point: caller workspace
    da = L.BnDesc(m, c, dt, flags, 1e-5, 0.1, a["nbt"].data_ptr())
    ws = torch.empty(lib.cnb_bn_workspace_bytes(C.byref(da)), dtype=torch.uint8, device="cuda")
    L.check(lib.cnb_bn_fwd_train(C.byref(da), _p(x), _p(res), _p(gamma), _p(beta), _p(a["rmean"]),
                                 _p(a["rvar"]), _p(a["y"]), _p(a["mean"]), _p(a["rstd"]), None, 0,
                                 _p(a["mask"]), _p(ws), ws.numel(), _stream()), "cnb_bn_fwd_train")
    torch.cuda.synchronize()

    # new entry point: everything it touches was allocated above; it must allocate nothing
    db = L.BnDesc(m, c, dt, flags, 1e-5, 0.1, b["nbt"].data_ptr())
    torch_before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    L.check(lib.cnb_bn_fwd_train_static(C.byref(db), _p(x), _p(res), _p(gamma), _p(beta), _p(b["rmean"]),
                                        _p(b["rvar"]), _p(b["y"]), _p(b["mean"]), _p(b["rstd"]),
                                        _p(b["mask"]), _stream()), "cnb_bn_fwd_train_static")
    torch.cuda.synchronize()
    unchanged = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == torch_before
    return a, b, unchanged


def _compare(a, b, dtype):
    tol = BF16_TOL if dtype == torch.bfloat16 else F32_TOL
    torch.testing.assert_close(b["y"].float(), a["y"].float(), rtol=tol, atol=tol)
    for k in ("mean", "rstd", "rmean", "rvar"):
        torch.testing.assert_close(b[k], a[k], rtol=F32_TOL, atol=F32_TOL)
    assert int(b["nbt"]) == int(a["nbt"]) == 1
    if a["mask"] is not None:
        assert torch.equal(b["mask"], a["mask"])
    assert torch.isfinite(b["y"].float()).all()


def test_static_scratch_bn_forward_matches_workspace_entry_twice():
    lib = L.load()
    assert lib.cnb_bn_static_scratch_bytes() == 1 << 20
    # (rows, channels, dtype, relu, residual): a ResNet-50 layer3 slab, then a narrower fp32 tensor
    # with a ragged row count — the second call's partial-sum layout differs from what the first left
    cases = [
        (8 * 14 * 14, 1024, torch.bfloat16, True, True, 11),
        (3 * 7 * 7 + 5, 96, torch.float32, False, False, 12),
    ]
    # one call first: the first use on a device looks the array's address up (cudaGetSymbolAddress)
    _run_pair(lib, 64, 64, torch.bfloat16, True, False, 10)
    for m, c, dtype, relu, add, seed in cases:
        a, b, allocator_unchanged = _run_pair(lib, m, c, dtype, relu, add, seed)
        _compare(a, b, dtype)
        assert allocator_unchanged
    # the same inputs again: the array now holds this very call's leftovers
    m, c, dtype, relu, add, seed = cases[0]
    a, b, _ = _run_pair(lib, m, c, dtype, relu, add, seed)
    _compare(a, b, dtype)


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    lib = L.load()
    c = 16384 + 8
    m = 4
    x = torch.zeros(m, c, dtype=torch.bfloat16, device="cuda")
    y = torch.full_like(x, 7.0)
    mean = torch.empty(c, dtype=torch.float32, device="cuda")
    rstd = torch.empty(c, dtype=torch.float32, device="cuda")
    d = L.BnDesc(m, c, L.CNB_BF16, 0, 1e-5, 0.1, None)
    status = lib.cnb_bn_fwd_train_static(C.byref(d), _p(x), None, None, None, None, None, _p(y), _p(mean),
                                         _p(rstd), None, _stream())
    torch.cuda.synchronize()
    assert status == CNB_ERR_WORKSPACE
    assert "16384" in L.last_error()
    assert bool((y == 7.0).all())    # nothing was launched
