"""Deterministic synthetic weights and inputs shared by oracle/make_golden.py and the tests.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Golden fixtures cannot carry whole state_dicts (ResNet-50 = 94 MB), so both the generating script
(which runs the real reference in the build container) and the tests (which run the oracle, or the
CUDA library on the GPU box) re-create identical weights from the state_dict KEYS and SHAPES alone:
every tensor is drawn from its own CPU generator seeded by crc32(key), with a scale chosen by the
tensor's role so that activations stay O(1) through deep nets (He-style fan-in scaling for convs,
BatchNorm near identity with non-trivial running statistics).  Nothing here depends on module
construction order or on torch's global RNG.
"""
from __future__ import annotations

import zlib
from typing import Dict, Sequence, Tuple

import torch


def _gen(key: str, salt: int = 0) -> torch.Generator:
    return torch.Generator().manual_seed((zlib.crc32(key.encode()) + 7919 * salt) & 0x7FFFFFFF)


def fill_tensor(key: str, shape: Sequence[int], salt: int = 0) -> torch.Tensor:
    """fp32 (int64 for num_batches_tracked) tensor for one state_dict entry."""
    shape = tuple(int(s) for s in shape)
    g = _gen(key, salt)
    if key.endswith("num_batches_tracked"):
        return torch.zeros(shape, dtype=torch.int64)
    if key.endswith("running_mean"):
        return 0.1 * torch.randn(shape, generator=g)
    if key.endswith("running_var"):
        return 1.0 + 0.2 * torch.rand(shape, generator=g)
    if len(shape) == 4:  # conv weight [K, C/g, R, S]: He fan-in
        fan_in = shape[1] * shape[2] * shape[3]
        return torch.randn(shape, generator=g) * (2.0 / fan_in) ** 0.5
    if len(shape) == 2:  # Linear weight [out, in]
        return torch.randn(shape, generator=g) * (1.0 / shape[1]) ** 0.5
    if key.endswith(".weight"):  # BatchNorm gamma
        return 1.0 + 0.1 * torch.randn(shape, generator=g)
    return 0.1 * torch.randn(shape, generator=g)  # biases / BatchNorm beta


def fill_state_dict(spec: Sequence[Tuple[str, Sequence[int]]], salt: int = 0) -> Dict[str, torch.Tensor]:
    """spec = [(key, shape), ...] in state_dict order."""
    return {k: fill_tensor(k, s, salt) for k, s in spec}


def spec_of(state_dict) -> list:
    return [(k, list(v.shape)) for k, v in state_dict.items()]


def batch(n: int, hw: int, num_classes: int = 10, seed: int = 21):
    """Synthetic CINIC-10-shaped batch: seed 21 = Settings.DEF_SEED (reference src/settings.py:222)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 3, hw, hw, generator=g)
    y = torch.randint(0, num_classes, (n,), generator=g)
    return x, y
