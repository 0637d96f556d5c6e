"""Multi-tensor Adam over the C ABI (`cnb_adam_step`).

The reference builds `optim.Adam(self.parameters(), lr, weight_decay)` (src/basemodel.py:58-66)
and steps it once per batch (:284).  `FusedAdam` keeps that class's interface — param_groups (so
ReduceLROnPlateau keeps working), state keys `step` / `exp_avg` / `exp_avg_sq`, state_dict layout —
and replaces the update with ONE library call per parameter group: every parameter of the model in
1 + ceil(n_params / 56) launches, fp32 arithmetic in the order of torch's single-tensor Adam (the
reference's CPU path; oracle: oracle/ops.py adam_step), bias corrections in double.

For the dense tensor-core convolutions of a converted model the same pass also writes the two bf16
operand copies (KRSC for forward / weight-grad, CRSK for data-grad) into the layer group's
persistent buffers (layers.PackGroup), so the separate pack pass over the weights — and torch's
Adam kernels — leave the step.  The step count and the hyper-parameters live on the device
(16 doubles per group): a CUDA-graph replay of the step advances them without host code;
`sync_hyperparams()` (called by StepGraph before every replay) pushes a changed learning rate.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import layers as L
from . import lib as _L


def _dense(t: torch.Tensor) -> str:
    """Physical order of a dense tensor: 'c' (row-major) and / or 'l' (channels_last)."""
    f = ""
    if t.is_contiguous():
        f += "c"
    if t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last):
        f += "l"
    return f


class FusedAdam(torch.optim.Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, model=None):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        for g in self.param_groups:
            if g.get("amsgrad") or g.get("maximize"):
                raise NotImplementedError("FusedAdam: amsgrad / maximize are not used by the reference")
        self._dev_state = {}     # group index -> float64[16] on the device
        self._pushed = {}        # group index -> hyper-parameters last written to the device
        self._packs = {}         # id(param) -> (layer, PackGroup, index in the group)
        self._tables = {}        # group index -> (key, ctypes array, n)
        self._written = {}       # group index -> ids of the layers whose copies the last table writes
        if model is not None:
            self.attach(model)

    # ---------------------------------------------------------------- operand copies
    def attach(self, model) -> None:
        """Let the step maintain the bf16 operand copies of `model`'s dense tensor-core filters."""
        self._packs.clear()
        seen = set()
        for m in model.modules():
            grp = m.__dict__.get("_cnb_group") if isinstance(m, L.Conv2d) else None
            if grp is None or id(grp) in seen:
                continue
            seen.add(id(grp))
            for i, layer in enumerate(grp.layers):
                self._packs[id(layer.weight)] = (layer, grp, i)
        self._tables.clear()
        self._written.clear()

    def _groups_in_use(self):
        out, seen = [], set()
        for _layer, grp, _i in self._packs.values():
            if id(grp) not in seen:
                seen.add(id(grp))
                out.append(grp)
        return out

    def restamp(self) -> None:
        """The operand copies are current (this step wrote them): mark the layers' caches valid
        under the weight epoch the post-step hook just advanced."""
        for grp in self._groups_in_use():
            if grp.bufs is None:
                continue
            written = set().union(*self._written.values()) if self._written else set()
            for m, pair in zip(grp.layers, grp.bufs):
                if id(m) not in written:
                    continue     # e.g. a frozen layer: its cache key carries the old epoch, it re-packs
                w = m.weight
                m.__dict__["_cnb_pack"] = ((L._WEIGHT_EPOCH[0], w._version, w.data_ptr(), torch.bfloat16), pair)

    # ---------------------------------------------------------------- device-resident scalars
    def _state_for(self, gi, device):
        st = self._dev_state.get(gi)
        if st is None or st.device != device:
            st = torch.zeros(16, dtype=torch.float64, device=device)
            self._dev_state[gi] = st
            self._pushed.pop(gi, None)
        return st

    def sync_hyperparams(self) -> None:
        """Write lr / betas / eps / weight_decay of every group to its device block if they
        changed (ReduceLROnPlateau edits param_groups on the host).  Not capturable: call it
        outside a stream capture (StepGraph does, before each replay)."""
        for gi, g in enumerate(self.param_groups):
            params = [p for p in g["params"] if p.requires_grad]
            if not params:
                continue
            st = self._state_for(gi, params[0].device)
            lr = g["lr"]
            hp = (float(lr), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]))
            if self._pushed.get(gi) != hp:
                st[1:6].copy_(torch.tensor(hp, dtype=torch.float64))
                self._pushed[gi] = hp

    # ---------------------------------------------------------------- the step
    def _init_state(self, p):
        s = self.state[p]
        if len(s) == 0:
            s["step"] = torch.zeros((), dtype=torch.float32, device=p.device)
            s["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            s["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return s

    def ensure_state(self) -> None:
        """Allocate the moments of every trainable parameter now.  StepGraph calls this before a
        capture: a `zeros` issued INSIDE the capture would become a memset node and every replay
        would clear the moments."""
        for g in self.param_groups:
            for p in g["params"]:
                if p.requires_grad:
                    self._init_state(p)

    def _table(self, gi, params):
        key = tuple((p.data_ptr(), p.grad.data_ptr()) for p in params)
        cached = self._tables.get(gi)
        if cached is not None and cached[0] == key:
            return cached[1], cached[2]
        items = (_L.AdamItem * len(params))()
        written = set()
        for i, p in enumerate(params):
            s = self._init_state(p)
            g, m, v = p.grad, s["exp_avg"], s["exp_avg_sq"]
            if p.dtype != torch.float32 or g.dtype != torch.float32 or g.is_sparse:
                raise TypeError("FusedAdam updates dense fp32 master parameters with fp32 gradients")
            common = set(_dense(p)) & set(_dense(g)) & set(_dense(m)) & set(_dense(v))
            if not common:
                raise RuntimeError(
                    f"FusedAdam: parameter {tuple(p.shape)} and its gradient / moments are not dense in "
                    f"the same physical order (param {p.stride()}, grad {g.stride()})")
            wb = wt = None
            k, rs, c = 1, 0, p.numel()
            pk = self._packs.get(id(p))
            if pk is not None and p.is_cuda and "l" in common:
                _layer, grp, idx = pk
                if grp.bufs is None or grp.ptrs != tuple(q.weight.data_ptr() for q in grp.layers):
                    grp._build()
                wb_t, wt_t = grp.bufs[idx]
                k, rs, c = p.shape[0], p.shape[2] * p.shape[3], p.shape[1]
                wb = wb_t.data_ptr()
                wt = None if wt_t is None else wt_t.data_ptr()
                written.add(id(_layer))
            items[i] = _L.AdamItem(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), wb, wt,
                                   int(k), int(rs), int(c), 0)
        self._tables[gi] = (key, items, len(params))
        self._written[gi] = written
        return items, len(params)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not torch.cuda.is_current_stream_capturing():
            self.sync_hyperparams()
        lib = _L.load()
        for gi, g in enumerate(self.param_groups):
            params = [p for p in g["params"] if p.grad is not None]
            if not params:
                continue
            if not all(p.is_cuda for p in params):
                raise RuntimeError("FusedAdam runs on CUDA parameters only (no CPU fallback)")
            st = self._state_for(gi, params[0].device)
            if gi not in self._pushed:
                raise RuntimeError("FusedAdam: hyper-parameters were never pushed to the device "
                                   "(call sync_hyperparams() before capturing the step)")
            items, n = self._table(gi, params)
            _L.check(lib.cnb_adam_step(items, n, C.c_void_p(st.data_ptr()),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cnb_adam_step")
        return loss

    # ---------------------------------------------------------------- torch.optim.Adam-compatible state
    def step_count(self, gi: int = 0) -> int:
        st = self._dev_state.get(gi)
        return 0 if st is None else int(st[0].item())

    def state_dict(self):
        for gi, g in enumerate(self.param_groups):
            st = self._dev_state.get(gi)
            if st is None:
                continue
            for p in g["params"]:
                s = self.state.get(p)
                if s:
                    s["step"].copy_(st[0])
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._tables.clear()
        self._pushed.clear()
        for gi, g in enumerate(self.param_groups):
            for p in g["params"]:
                s = self.state.get(p)
                if s and "step" in s:
                    st = self._state_for(gi, p.device)
                    st[0] = float(s["step"])
                    break
