"""One CUDA graph for the whole training step.

A converted zoo model enqueues 450 (ResNet-50) to 1 500 (DenseNet-121) kernels per step through
ctypes and the autograd engine; at 15-25 us of host time per launch the small-kernel models are
host-bound and every model pays for the bubble after the per-step `loss.item()` of the reference
loop (src/basemodel.py:286-288).  `StepGraph` captures

    zero_grad -> forward -> CrossEntropyLoss(sum) -> backward -> [gradient all-reduce] -> clip -> Adam

once (after a few eager steps have settled the allocator, the workspaces and the library's lazily
created state) and replays it per batch: inputs are copied into static device buffers, loss and
logits are read from static outputs.  Everything the library needs is capture-safe by
construction: kernels launch on the current stream, tensor maps are kernel parameters encoded on
the host, scratch memory comes from torch's graph-private pool, and the packed operand copies of
the weights are rebuilt INSIDE the graph (their cache is invalidated before the capture) so that
every replay sees the weights the previous replay's Adam step produced.
"""
from __future__ import annotations

import gc

import torch

from . import functional as F_
from . import layers as L


def _invalidate_packs(model: torch.nn.Module) -> None:
    for m in model.modules():
        if isinstance(m, L.Conv2d):
            m.__dict__.pop("_cnb_pack", None)


class StepGraph:
    """Callable `(inputs, targets) -> (loss, logits)` running one training step of `model`
    (a zoo model with `.optimizer`, `.cost_function`, `.gradient_clipping`).  The first
    `eager_steps` calls run eagerly; the next call captures and replays."""

    def __init__(self, model, eager_steps: int = 3):
        self.model = model
        self.eager_steps = int(eager_steps)
        self.calls = 0
        self.graph = None
        self.key = None
        self.x = self.y = self.loss = self.logits = None
        self.launches_per_step = 0
        self.stream = None  # side stream of the warm-up steps and the capture (PyTorch's recipe)

    # one eager step (also the body that gets captured)
    def _step(self, x, y):
        m = self.model
        m.optimizer.zero_grad(set_to_none=True)
        out = m(x)
        loss = m.cost_function(out.float(), y)
        loss.backward()
        m.gradient_clipping()
        m.optimizer.step()
        return loss, out

    def _capture(self, x, y):
        from . import lib as _L
        m = self.model
        dev = x.device
        self.x = torch.empty_like(x)
        self.y = torch.empty_like(y)
        self.x.copy_(x)
        self.y.copy_(y)
        _invalidate_packs(m)
        opt = m.optimizer
        if hasattr(opt, "sync_hyperparams"):
            # optim.FusedAdam: hyper-parameters go to the device now (not capturable), and the bf16
            # operand copies it maintains are brought up to date eagerly — inside the graph the
            # forward then finds them valid and the Adam step rewrites them: no pack kernel per replay
            opt.sync_hyperparams()
            opt.ensure_state()
            for grp in opt._groups_in_use():
                grp.refresh()
        F_._workspaces.clear()      # scratch must come from the graph's private pool
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        m.optimizer.zero_grad(set_to_none=True)
        n0 = _L.launch_count()
        if self.stream is None:
            self.stream = torch.cuda.Stream(device=dev)
        # No garbage collection while the stream is capturing: a collected cycle that owns CUDA
        # resources (an older model, graph or event) would issue cudaFree / cudaGraphExecDestroy
        # in the middle of the capture and invalidate it (seen once in the full GPU test run).
        gc.collect()
        gc_was_enabled = gc.isenabled()
        gc.disable()
        try:
            with torch.cuda.graph(g, stream=self.stream):
                self.loss, self.logits = self._step(self.x, self.y)
        finally:
            if gc_was_enabled:
                gc.enable()
        self.launches_per_step = _L.launch_count() - n0
        F_._workspaces.clear()      # eager code must not reuse the graph's scratch buffers
        self.graph = g
        self.key = (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype)

    def _eager_on_side_stream(self, x, y):
        # the eager steps before the capture run on the stream the capture will use, so that
        # every gradient accumulator and lazily created buffer is first touched there
        cur = torch.cuda.current_stream(x.device)
        if self.stream is None:
            self.stream = torch.cuda.Stream(device=x.device)
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            out = self._step(x, y)
        cur.wait_stream(self.stream)
        return out

    def __call__(self, x, y):
        self.calls += 1
        if not self.model.training:
            raise RuntimeError("StepGraph runs training steps: call model.train() first")
        if self.graph is None:
            if self.calls <= self.eager_steps:
                return self._eager_on_side_stream(x, y)
            self._capture(x, y)
        elif self.key != (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype):
            # ragged last batch of an epoch: run it eagerly, keep the graph for the full batches
            return self._step(x, y)
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x, non_blocking=True)
        if y.data_ptr() != self.y.data_ptr():
            self.y.copy_(y, non_blocking=True)
        opt = self.model.optimizer
        if hasattr(opt, "sync_hyperparams"):
            opt.sync_hyperparams()      # a learning rate changed by the scheduler reaches the replay
        self.graph.replay()
        # The replayed Adam step changed every weight without running any Python: neither the
        # optimizer post-step hook nor the tensor version counter moved.  Advance the weight epoch so
        # that the next EAGER forward (valid_model, a ragged last batch) re-packs its bf16 operand
        # copies instead of reusing the ones this replay built from the pre-step weights.
        L._bump_weight_epoch(opt)
        return self.loss, self.logits


class EvalGraph:
    """Callable `inputs -> logits` replaying ONE CUDA graph of the inference forward (reference
    valid_model / test, src/basemodel.py:304-346,601-722) for full batches; other shapes run
    eagerly.  At 32x32 the zoo's forward is 100-400 kernels of a few microseconds each, so the
    eager loop is bound by the host's launch rate, not by the GPU.  The folded filters / packed
    operand copies are built by one eager forward BEFORE the capture and are read-only inside the
    graph: re-capture (`reset()`) after the weights change."""

    def __init__(self, model):
        self.model = model
        self.graph = None
        self.key = None
        self.x = self.out = None
        self.epoch = None

    def reset(self):
        self.graph = None

    @torch.no_grad()
    def __call__(self, x):
        m = self.model
        if m.training:
            raise RuntimeError("EvalGraph runs inference: call model.eval() first")
        key = (tuple(x.shape), x.dtype)
        if self.graph is not None and (self.key != key or self.epoch != L._WEIGHT_EPOCH[0]):
            if self.key != key:
                return m(x)                       # ragged last batch
            self.reset()                          # weights changed since the capture
        if self.graph is None:
            self.x = torch.empty_like(x)
            self.x.copy_(x)
            m(self.x)                             # builds every folded / packed operand eagerly
            F_._workspaces.clear()
            torch.cuda.synchronize(x.device)
            g = torch.cuda.CUDAGraph()
            gc.collect()
            was = gc.isenabled()
            gc.disable()
            try:
                with torch.cuda.graph(g):
                    self.out = m(self.x)
            finally:
                if was:
                    gc.enable()
            F_._workspaces.clear()
            self.graph, self.key, self.epoch = g, key, L._WEIGHT_EPOCH[0]
        if x.data_ptr() != self.x.data_ptr():
            self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.out


def enable_eval_graph(model) -> EvalGraph:
    """Make `model.test()` (and anything else that reads `model._cnb_eval_graph`) run full-batch
    inference forwards through an EvalGraph."""
    eg = EvalGraph(model)
    model.__dict__["_cnb_eval_graph"] = eg
    return eg


def enable_step_graph(model, eager_steps: int = 3) -> StepGraph:
    """Make `model.train_model` run its steps through a StepGraph (zoo models only)."""
    sg = StepGraph(model, eager_steps)
    model._cnb_step_graph = sg
    return sg
