"""Host-side mirror of the reference's configuration objects for the training hot path.

`Settings` mirrors reference src/settings.py:169-319 (same keyword names, same DEF_* defaults,
None -> default fill) minus the random-search distributions (src/settings.py:66-166, used only by
the tuner, out of scope).  `DeviceMngr` mirrors src/mngrdevice.py:5-43.  These exist so that the
model zoo in this package can be constructed on a box where /root/reference is absent (the GPU
box); a reference `Settings` instance works just as well — only attribute reads are used.
"""
from __future__ import annotations

import torch


class DeviceMngr:
    """reference src/mngrdevice.py:5-43 (cuda:0 when available, else cpu; recursive move)."""

    def __init__(self, device_str: str | None = None, verbose: bool = False):
        if device_str is None:
            device_str = "cuda:0" if torch.cuda.is_available() else "cpu"
        self.device = torch.device(device_str)
        self.verbose = verbose

    def get_default_device(self):
        return torch.device("cuda:0" if torch.cuda.is_available() else "cpu")

    def set_device(self, device_str):
        self.device = torch.device(device_str)
        if self.verbose:
            print("You are using device:", device_str)

    def move(self, data):
        if isinstance(data, (list, tuple)):
            return [self.move(x) for x in data]
        return data.to(self.device, non_blocking=True)

    def move_to(self, data, device_str):
        return data.to(torch.device(device_str), non_blocking=True)


_DEFAULTS = dict(
    batch_size=256, batch_norm=True, epochs=50, learning_rate=0.01, lr_factor=0.1, lr_patience=10,
    weight_decay=1e-4, dropout_rate=0.5, loss_optim=False, data_augment=True, data_norm=True,
    early_stop=True, es_patience=12, grad_clip_norm=False, gc_max_norm=1, grad_clip_value=False,
    gc_value=1, init_params=True, distrib=None, sanity_check=False, debug=False, device=None,
    num_workers=16, mixed_precision=True, test_sample_size=90, seed=21,
)


class Settings:
    """Keyword-compatible with reference src/settings.py:224-300; unspecified (None) fields take
    the reference's DEF_* values (src/settings.py:177-222)."""

    def __init__(self, kind, input_size, num_classes, **kwargs):
        unknown = set(kwargs) - set(_DEFAULTS)
        if unknown:
            raise TypeError(f"Settings got unexpected keyword(s): {sorted(unknown)}")
        self.kind = kind
        self.input_size = tuple(input_size)
        self.num_classes = num_classes
        for name, default in _DEFAULTS.items():
            value = kwargs.get(name)
            setattr(self, name, default if value is None else value)
        if self.device is None:
            self.device = DeviceMngr()

    def to_dict(self):
        return dict(self.__dict__)
