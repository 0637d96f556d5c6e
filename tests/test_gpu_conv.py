"""GPU parity of the convolution entry points (cnb_conv2d_fwd / dgrad / wgrad) against the
oracle, per layer signature.  Every case runs in its own subprocess (tests/tc_probe.py) with a
timeout, so a faulting tensor-core kernel fails one case instead of wedging the session.

Tolerances (BASELINE.json north_star): rel-L2 <= 1e-2 in bf16 mode, <= 1e-4 in fp32-IO mode.
Oracle inputs are pre-rounded to the storage dtype, so only accumulation/output rounding differ.
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE = os.path.join(ROOT, "tests", "tc_probe.py")


def c(N, C, H, W, K, R, S=None, **kw):
    d = dict(N=N, C=C, H=H, W=W, K=K, R=R, S=R if S is None else S)
    d.update(kw)
    return d


# tensor-core (tcgen05) cases: ResNet-50 / DenseNet / Inception style signatures at small batch
TC_CASES = [
    c(2, 64, 16, 16, 64, 1),                          # pure GEMM, one n-tile of 64
    c(3, 256, 7, 7, 128, 1),                          # M = 147 (ragged last m-tile), K-loop of 4
    c(2, 128, 14, 14, 512, 1),                        # BLOCK_N = 256, two n-tiles
    c(2, 64, 14, 14, 64, 3, pad=1),                   # 3x3 im2col, padding by zero-fill
    c(2, 128, 15, 15, 128, 3, pad=1, stride=2),       # strided 3x3 (parity-class dgrad)
    c(2, 256, 14, 14, 512, 1, stride=2),              # strided 1x1 projection
    c(2, 32, 12, 12, 64, 3, pad=2, dil=2),            # dilated, C < 64 (channel zero-fill)
    c(2, 16, 9, 9, 32, 5, pad=2),                     # 5x5 (Inception), C = 16
    c(2, 8, 20, 20, 64, 7, pad=3, stride=2),          # 7x7 stride 2 stem-like, C = 8
    c(1, 96, 10, 10, 208, 3, pad=1),                  # C, K not multiples of 64
    c(2, 64, 8, 8, 64, 3, pad=1, bias=True),          # bias epilogue + bias grad
    c(4, 512, 7, 7, 512, 3, pad=1),                   # deep K loop (72 k-blocks), NT=256 wgrad
    c(8, 64, 56, 56, 64, 3, pad=1),                   # many m-tiles, split-K wgrad
    c(8, 256, 28, 28, 128, 1),                        # wgrad NT=256 x 2 cout tiles? (K=128: 1 tile)
    c(4, 512, 2, 2, 2048, 1),                         # M = 16: one ragged m-tile (ResNet-50 @64x64 tail)
    c(4, 2048, 2, 2, 512, 1),                         # M = 16, 32 k-blocks
    c(4, 512, 2, 2, 512, 3, pad=1),                   # 3x3 on a 2x2 map: most taps are padding
    c(4, 512, 4, 4, 512, 3, pad=1, stride=2),         # strided 3x3, 4x4 -> 2x2
    c(4, 1024, 4, 4, 2048, 1, stride=2),              # strided projection, 4x4 -> 2x2
    # grouped layers on the tensor-core path (block-diagonal filter expansion per 64-channel slice)
    c(2, 128, 10, 10, 128, 3, pad=1, groups=32),                 # ResNeXt stage 1: 4 ch / group
    c(2, 256, 9, 9, 256, 3, pad=1, stride=2, groups=32),         # 8 ch / group, strided (parity dgrad)
    c(2, 512, 7, 7, 512, 3, pad=2, dil=2, groups=32),            # SK path 2: dilated, 16 ch / group
    c(3, 1024, 4, 4, 1024, 3, pad=1, groups=32),                 # 32 ch / group, 16 n-tiles
    c(2, 480, 7, 7, 120, 1, groups=3),                           # ShuffleNet g3 compress (Cg 160, Kg 40)
    c(2, 120, 7, 7, 480, 1, groups=3),                           # ShuffleNet g3 expand (Cg 40, Kg 160)
    c(2, 960, 4, 4, 240, 1, groups=3),                           # wide groups: Cg 320, Kg 80
    c(2, 240, 4, 4, 960, 1, groups=3),                           # wide groups: Cg 80, Kg 320
    c(2, 64, 12, 12, 128, 3, pad=1, groups=2, bias=True),        # Cg 32 -> Kg 64, bias
]

# the 22 tensor-core signatures of ResNet-50 @224 at their real H x W (batch 2;
# the 3 -> 64 7x7 stem is covered by tests/test_gpu_fused_paths.py::test_rgb_stem_vs_oracle at 224)
APPENDIX_A = [
    c(2, 64, 56, 56, 64, 1), c(2, 64, 56, 56, 64, 3, pad=1), c(2, 64, 56, 56, 256, 1),
    c(2, 256, 56, 56, 64, 1), c(2, 256, 56, 56, 128, 1), c(2, 128, 56, 56, 128, 3, pad=1, stride=2),
    c(2, 128, 28, 28, 512, 1), c(2, 256, 56, 56, 512, 1, stride=2), c(2, 512, 28, 28, 128, 1),
    c(2, 128, 28, 28, 128, 3, pad=1), c(2, 512, 28, 28, 256, 1), c(2, 256, 28, 28, 256, 3, pad=1, stride=2),
    c(2, 256, 14, 14, 1024, 1), c(2, 512, 28, 28, 1024, 1, stride=2), c(2, 1024, 14, 14, 256, 1),
    c(2, 256, 14, 14, 256, 3, pad=1), c(2, 1024, 14, 14, 512, 1), c(2, 512, 14, 14, 512, 3, pad=1, stride=2),
    c(2, 512, 7, 7, 2048, 1), c(2, 1024, 14, 14, 2048, 1, stride=2), c(2, 2048, 7, 7, 512, 1),
    c(2, 512, 7, 7, 512, 3, pad=1),
]

# CUDA-core cases: fp32-IO mode, grouped, tiny channel counts, stems
SIMT_CASES = [
    c(2, 3, 17, 17, 24, 3, pad=1, stride=2, dtype="f32"),              # RGB stem
    c(2, 3, 32, 32, 64, 7, pad=3, stride=2, dtype="bf16"),             # bf16 stem, C = 3
    c(2, 64, 9, 9, 64, 3, pad=1, dtype="f32"),                         # fp32-IO dense
    c(2, 128, 10, 10, 128, 3, pad=1, groups=32, dtype="bf16"),         # ResNeXt 4 ch/group
    c(2, 128, 10, 10, 128, 3, pad=2, dil=2, groups=32, stride=2, dtype="f32"),  # SK path 2
    c(2, 240, 7, 7, 60, 1, groups=3, dtype="bf16"),                    # ShuffleNet g3 compress
    c(2, 60, 7, 7, 240, 1, groups=3, dtype="f32", bias=True),          # ShuffleNet g3 expand
    c(3, 32, 1, 1, 64, 1, dtype="f32", bias=True),                     # SK attention conv [B,d,1,1]
    c(2, 64, 14, 14, 64, 3, pad=1, dtype="bf16", algo="simt"),         # forced CUDA-core bf16
    c(2, 32, 13, 13, 32, 3, pad=1, groups=32, dtype="bf16"),           # depthwise s1
    c(2, 64, 14, 14, 64, 3, pad=1, stride=2, groups=64, dtype="f32"),  # depthwise s2
    c(2, 60, 7, 7, 60, 3, pad=1, stride=2, groups=60, dtype="bf16"),   # depthwise, C % 8 != 0
    c(2, 64, 15, 15, 64, 3, pad=1, stride=2, groups=64, dtype="bf16"), # depthwise s2, odd extent
    c(2, 32, 7, 10, 32, 3, pad=1, groups=32, dtype="bf16", bias=True), # depthwise strips: ragged row, bias
    c(2, 40, 9, 6, 40, 3, pad=1, stride=2, groups=40, dtype="f32"),    # depthwise s2 fp32, W < one strip pair
    c(1, 1024, 7, 7, 1024, 3, pad=1, groups=1024, dtype="bf16"),       # MobileNet tail
    c(3, 120, 14, 14, 120, 3, pad=1, stride=2, groups=120, dtype="bf16"),  # ShuffleNet: 15 channel vectors
    c(4, 24, 9, 9, 60, 1, dtype="bf16"),                               # ShuffleNet first compress (narrow pointwise)
    c(3, 60, 11, 11, 216, 1, groups=3, dtype="bf16"),                  # ShuffleNet first expand: Cg 20, Kg 72
    c(5, 240, 13, 13, 60, 1, groups=3, dtype="bf16", bias=True),       # narrow pointwise, ragged 128-pixel tile
    c(2, 96, 6, 6, 96, 1, dtype="f32"),                                # pointwise path at its 96 x 96 limit
    c(16, 1024, 1, 1, 64, 1, dtype="f32"),                             # SK descriptor conv on [B,C,1,1]
    c(16, 64, 1, 1, 1024, 1, dtype="f32", bias=True),                  # SK attention conv (vector GEMM)
]


def run_probe(case, timeout=180):
    env = dict(os.environ)
    p = subprocess.run([sys.executable, PROBE, json.dumps(case)], capture_output=True, text=True,
                       timeout=timeout, env=env)
    line = [l for l in p.stdout.splitlines() if l.startswith("PROBE_RESULT ")]
    if p.returncode != 0 or not line:
        pytest.fail(f"probe failed rc={p.returncode}\nstdout:\n{p.stdout[-2000:]}\nstderr:\n{p.stderr[-3000:]}")
    return json.loads(line[-1][len("PROBE_RESULT "):])


def run_probe_many(cases, timeout=600):
    """All `cases` in ONE subprocess (one `import torch`), one PROBE_RESULT line per case."""
    p = subprocess.run([sys.executable, PROBE, json.dumps(cases)], capture_output=True, text=True,
                       timeout=timeout, env=dict(os.environ))
    lines = [l for l in p.stdout.splitlines() if l.startswith("PROBE_RESULT ")]
    if p.returncode != 0 or len(lines) != len(cases):
        pytest.fail(f"probe failed rc={p.returncode} ({len(lines)}/{len(cases)} results)\nstdout:\n"
                    f"{p.stdout[-2000:]}\nstderr:\n{p.stderr[-3000:]}")
    return [json.loads(l[len("PROBE_RESULT "):]) for l in lines]


def check(res, tol):
    msgs = []
    for k in ("fwd", "dgrad", "wgrad", "bgrad"):
        if k in res and not (res[k]["rel_l2"] <= tol):
            msgs.append(f"{k}: {json.dumps(res[k])}")
    assert not msgs, "parity failure for %s (uses_tc=%s):\n%s" % (
        json.dumps(res["case"]), res["uses_tc"], "\n".join(msgs))


@pytest.mark.gpu
@pytest.mark.parametrize("case", TC_CASES, ids=lambda d: "N{N}C{C}H{H}K{K}R{R}".format(**d) + "".join(
    f"{k}{v}" for k, v in d.items() if k in ("stride", "pad", "dil", "bias", "groups")))
def test_conv_tensor_core(case):
    case = dict(case, dtype="bf16", algo="tc")
    res = run_probe(case)
    assert res["uses_tc"] == [True, True, True]
    check(res, 1e-2)


@pytest.mark.gpu
def test_conv_resnet50_signatures_at_real_size():
    """every tensor-core signature of ResNet-50 @224 at its real H x W,
    forward / data-grad / weight-grad against the float64 oracle (rel-L2 <= 1e-2)."""
    results = run_probe_many([dict(cs, dtype="bf16", algo="tc") for cs in APPENDIX_A])
    worst = {}
    for res in results:
        assert res["uses_tc"] == [True, True, True]
        check(res, 1e-2)
        for k in ("fwd", "dgrad", "wgrad"):
            worst[k] = max(worst.get(k, 0.0), res[k]["rel_l2"])
    print("Appendix A worst rel-L2:", {k: f"{v:.2e}" for k, v in worst.items()})


@pytest.mark.gpu
@pytest.mark.parametrize("case", SIMT_CASES, ids=lambda d: "N{N}C{C}H{H}K{K}R{R}".format(**d) + "".join(
    f"{k}{v}" for k, v in d.items() if k in ("stride", "pad", "dil", "groups", "dtype", "bias", "algo")))
def test_conv_cuda_core(case):
    res = run_probe(case)
    check(res, 1e-4 if case["dtype"] == "f32" else 1e-2)
