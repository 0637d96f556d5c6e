"""Data-parallel gradient exchange for the training hot path.

One process per GPU; every rank holds a full replica, draws its own batch shard and keeps its own
BatchNorm statistics (the reference has no multi-device code, src/mngrdevice.py:14-20; BASELINE
fixes per-replica BN).  The only exchange step is ONE all-reduce of all gradients per step with
op = SUM: the loss is CrossEntropyLoss(reduction='sum') (src/basemodel.py:46), so the full-batch
gradient is exactly the sum of the shard gradients.

`GradSync` keeps `train_model` (src/basemodel.py:231-302) unmodified: the all-reduces are launched
from per-parameter post-accumulate-grad hooks while the rest of backward is still running, in
buckets filled in reverse parameter order (the order gradients become ready), and an
end-of-backward callback makes the compute stream wait for them — so gradient clipping and
`optimizer.step()` (src/basemodel.py:281-284) always see the reduced gradients.

Gradients live in the bucket: after the hook copies a fresh gradient into its bucket slot, the
parameter's `.grad` is re-pointed at that slot (same shape and strides as the parameter), so the
reduced values are what the optimizer reads and no copy-back is needed.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist


class _Bucket:
    __slots__ = ("flat", "params", "offsets", "pending", "work")

    def __init__(self):
        self.flat: Optional[torch.Tensor] = None
        self.params: List[torch.nn.Parameter] = []
        self.offsets: List[int] = []
        self.pending = 0
        self.work = None


class GradSync:
    def __init__(self, model: torch.nn.Module, bucket_bytes: int = 25 << 20, group=None,
                 broadcast: bool = True):
        if not dist.is_initialized():
            raise RuntimeError("GradSync needs torch.distributed to be initialised")
        self.group = group
        self.world = dist.get_world_size(group)
        params = [p for p in model.parameters() if p.requires_grad]
        if broadcast:  # identical replicas: rank 0's weights and buffers win
            for t in list(model.parameters()) + list(model.buffers()):
                dist.broadcast(t.data, src=0, group=group)
        self.buckets: List[_Bucket] = []
        self._slot = {}
        cur, cur_bytes = _Bucket(), 0
        for p in reversed(params):
            nbytes = p.numel() * p.element_size()
            if cur.params and (cur_bytes + nbytes > bucket_bytes or p.dtype != cur.params[0].dtype
                               or p.device != cur.params[0].device):
                self.buckets.append(cur)
                cur, cur_bytes = _Bucket(), 0
            cur.offsets.append(sum(q.numel() for q in cur.params))
            cur.params.append(p)
            cur_bytes += nbytes
        if cur.params:
            self.buckets.append(cur)
        for bi, b in enumerate(self.buckets):
            n = sum(q.numel() for q in b.params)
            b.flat = torch.zeros(n, dtype=b.params[0].dtype, device=b.params[0].device)
            for p, off in zip(b.params, b.offsets):
                self._slot[p] = (bi, off)
        self._callback_queued = False
        self._handles = [p.register_post_accumulate_grad_hook(self._hook) for p in params]
        self.launched = 0  # all-reduces launched so far (bench.py reports it)

    # ---------------------------------------------------------------------------------------
    def _view(self, b: _Bucket, p: torch.Tensor, off: int) -> torch.Tensor:
        return b.flat.as_strided(p.shape, p.stride(), off)

    def _hook(self, p: torch.Tensor) -> None:
        bi, off = self._slot[p]
        b = self.buckets[bi]
        slot = self._view(b, p, off)
        if p.grad.data_ptr() != slot.data_ptr():
            slot.copy_(p.grad)
            p.grad = slot
        if b.pending == 0:
            b.pending = len(b.params)
        b.pending -= 1
        if not self._callback_queued:
            torch.autograd.Variable._execution_engine.queue_callback(self._finish)
            self._callback_queued = True
        if b.pending == 0:
            b.work = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self.launched += 1

    def _finish(self) -> None:
        self._callback_queued = False
        for b in self.buckets:
            if b.pending != 0:
                # parameters that received no gradient this step (unused branch): reduce what is
                # there so that ranks stay in lock-step
                for p, off in zip(b.params, b.offsets):
                    if p.grad is None or p.grad.data_ptr() != self._view(b, p, off).data_ptr():
                        self._view(b, p, off).zero_()
                b.work = dist.all_reduce(b.flat, op=dist.ReduceOp.SUM, group=self.group,
                                         async_op=True)
                self.launched += 1
                b.pending = 0
        for b in self.buckets:
            if b.work is not None:
                b.work.wait()
                b.work = None

    def remove(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles = []


def sum_scalar(value: float, device=None, group=None) -> float:
    """Epoch loss / score are per-rank Python floats (src/basemodel.py:287-288); sum them."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return float(t.item())
