"""Host-side mirror of the reference's base model: layer factories (the drop-in seam) and the
train / validation loop that calls the hot path.

Mirrors reference src/basemodel.py:
  * MultiClassBaseModel.__init__ (:33-56), init_optimizer (:58-83), init_params (:85-108),
    score (:111-128), gradient_clipping (:131-178), forward_step (:180-201),
    train_model (:231-302), valid_model (:304-346)
  * shape tracker save_conv_outshape (:725-781), save_adapt_outshape (:783-810),
    num_flat_features (:812-816)
  * layer factories conv2d (:966-979), conv2d_depthwise (:981-997), conv2d_block (:999-1015),
    maxpool2d (:1017-1024), avgpool2d (:1026-1033), adapt_avgpool2d (:1035-1042) and
    class Conv2dBlock (:1045-1067)
The factories build plain torch.nn layers exactly as the reference does (same constructor
arguments, same child names, hence the same state_dict keys); `convnets_b200.convert()` then
swaps them for the CUDA-library-backed layers.  Epoch control, checkpoints, plotting and the
tuner are out of scope.
"""
from __future__ import annotations

import math
import os

import torch
import torch.nn as nn
from torch.nn.utils import clip_grad_norm_, clip_grad_value_


def _hw(v):
    return (v, v) if isinstance(v, int) else tuple(v)


class MultiClassBaseModel(nn.Module):
    def __init__(self, setting):
        super().__init__()
        self.setting = setting
        self.model_name = type(self).__name__ + str(setting.kind)
        self.in_channels, self.height, self.width = setting.input_size
        self.cost_function = nn.CrossEntropyLoss(reduction="sum")  # basemodel.py:46
        self.optimizer = None
        self.lr_scheduler = None
        self.grad_scaler = None

    # ------------------------------------------------------------------ shape book-keeping
    def save_conv_outshape(self, layer):
        """floor(1 + (size + 2p - d(k-1) - 1) / s) per axis; channels follow a Conv2d."""
        k, s, p = _hw(layer.kernel_size), _hw(layer.stride), _hw(layer.padding)
        d = _hw(getattr(layer, "dilation", 1))
        self.height = int(math.floor(1 + (self.height + 2 * p[0] - d[0] * (k[0] - 1) - 1) / s[0]))
        self.width = int(math.floor(1 + (self.width + 2 * p[1] - d[1] * (k[1] - 1) - 1) / s[1]))
        if isinstance(layer, nn.Conv2d):
            self.in_channels = layer.out_channels
        return self.in_channels, self.height, self.width

    def save_adapt_outshape(self, layer):
        self.height, self.width = _hw(layer.output_size)
        return self.in_channels, self.height, self.width

    def num_flat_features(self):
        return self.in_channels * self.height * self.width

    # ------------------------------------------------------------------ layer factories
    def conv2d(self, in_channels=None, num_filters=None, set_output=True, **kwargs):
        if in_channels is None:
            in_channels = self.in_channels
        layer = nn.Conv2d(in_channels, num_filters, **kwargs, bias=not self.setting.batch_norm)
        if set_output:
            self.save_conv_outshape(layer)
        return layer

    def conv2d_depthwise(self, in_channels=None, num_filters=None, set_output=True, **kwargs):
        if in_channels is None:
            in_channels = self.in_channels
        if num_filters is None:
            num_filters = self.in_channels
        layer = nn.Conv2d(in_channels, num_filters, groups=in_channels, **kwargs,
                          bias=not self.setting.batch_norm)
        if set_output:
            self.save_conv_outshape(layer)
        return layer

    def conv2d_block(self, in_channels=None, num_filters=None, set_output=True, activation=True,
                     **kwargs):
        """Sequential(Conv2d, [BatchNorm2d], [ReLU(inplace)]) with integer child names."""
        layers = [self.conv2d(in_channels, num_filters, set_output, **kwargs)]
        if self.setting.batch_norm:
            layers.append(nn.BatchNorm2d(num_features=num_filters))
        if activation:
            layers.append(nn.ReLU(inplace=True))
        return nn.Sequential(*layers)

    def maxpool2d(self, set_output=True, **kwargs):
        layer = nn.MaxPool2d(**kwargs)
        if set_output:
            self.save_conv_outshape(layer)
        return layer

    def avgpool2d(self, set_output=True, **kwargs):
        layer = nn.AvgPool2d(**kwargs)
        if set_output:
            self.save_conv_outshape(layer)
        return layer

    def adapt_avgpool2d(self, set_output=True, **kwargs):
        layer = nn.AdaptiveAvgPool2d(**kwargs)
        if set_output:
            self.save_adapt_outshape(layer)
        return layer

    # ------------------------------------------------------------------ optimiser / init
    def init_optimizer(self):
        s = self.setting
        params = list(self.parameters())
        # same update rule as the reference's Adam (basemodel.py:58-66); on a CUDA device the
        # multi-tensor ("fused") implementation runs it in a handful of launches instead of ~10
        # per parameter
        fused = bool(params) and all(p.is_cuda for p in params)
        if fused and os.environ.get("CNB_FUSED_ADAM", "1") == "1":
            # the library's multi-tensor Adam: the whole update in 1 + ceil(n / 56) launches, writing
            # the bf16 operand copies of the tensor-core filters in the same pass (optim.FusedAdam)
            from ..optim import FusedAdam
            self.optimizer = FusedAdam(params, lr=s.learning_rate, weight_decay=s.weight_decay, model=self)
        else:
            # capturable: the step counters live on the device, so the update can sit inside the
            # CUDA graph of a whole training step (convnets_b200.graph.StepGraph)
            self.optimizer = torch.optim.Adam(params, lr=s.learning_rate, weight_decay=s.weight_decay,
                                              fused=fused, capturable=fused)
        self.lr_scheduler = torch.optim.lr_scheduler.ReduceLROnPlateau(
            self.optimizer, mode="min" if s.loss_optim else "max", factor=s.lr_factor,
            patience=s.lr_patience)
        self.grad_scaler = torch.amp.GradScaler(enabled=False)

    def init_params(self):
        """Kaiming-normal(fan_out, relu) convs, BN (1, 0), Linear N(0, 0.01), zero biases."""
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
            elif isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, mean=0, std=0.01)
            else:
                continue
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)

    # ------------------------------------------------------------------ the training hot loop
    @torch.no_grad()
    def score(self, outputs, targets):
        return (outputs.argmax(dim=1) == targets).sum().item()

    def gradient_clipping(self):
        s = self.setting
        if s.grad_clip_norm:
            clip_grad_norm_(self.parameters(), max_norm=s.gc_max_norm)
        if s.grad_clip_value:
            clip_grad_value_(self.parameters(), clip_value=s.gc_value)

    def forward_step(self, batch):
        inputs, targets = self.setting.device.move(batch)
        outputs = self(inputs)
        loss = self.cost_function(outputs.float(), targets)
        return loss, self.score(outputs, targets)

    @torch.enable_grad()
    def training_step(self, batch):
        return self.forward_step(batch)

    @torch.no_grad()
    def validation_step(self, batch):
        return self.forward_step(batch)

    @torch.enable_grad()
    def train_model(self, dataloader):
        """One pass over `dataloader`: zero_grad -> forward -> CE(sum) -> backward -> clip ->
        Adam step, accumulating loss.item() and the correct-prediction count; returns both
        divided by len(dataloader.dataset).  Low precision lives inside the converted layers
        (bf16 operands, fp32 accumulate, fp32 master weights), so the reference's fp16
        autocast + GradScaler branch (basemodel.py:253-270) is not taken."""
        self.train()
        epoch_loss, epoch_score = 0.0, 0
        graph = getattr(self, "_cnb_step_graph", None)
        if graph is not None:
            # same step, replayed from one CUDA graph (convnets_b200.enable_step_graph)
            for batch in dataloader:
                inputs, targets = self.setting.device.move(batch)
                loss, outputs = graph(inputs, targets)
                epoch_loss += loss.item()
                epoch_score += self.score(outputs, targets)
                if self.setting.sanity_check:
                    break
            n = len(dataloader.dataset)
            return epoch_loss / n, epoch_score / n
        for batch in dataloader:
            self.optimizer.zero_grad()
            loss, score = self.training_step(batch)
            loss.backward()
            self.gradient_clipping()
            self.optimizer.step()
            epoch_loss += loss.item()
            epoch_score += score
            if self.setting.sanity_check:
                break
        n = len(dataloader.dataset)
        return epoch_loss / n, epoch_score / n

    @torch.no_grad()
    def valid_model(self, dataloader):
        self.eval()
        epoch_loss, epoch_score = 0.0, 0
        for batch in dataloader:
            loss, score = self.validation_step(batch)
            epoch_loss += loss.item()
            epoch_score += score
            if self.setting.sanity_check:
                break
        n = len(dataloader.dataset)
        return epoch_loss / n, epoch_score / n


    # ------------------------------------------------------------------ the reference's benchmark
    def inference_time(self, total_times_ms, num_instances):
        """Reference src/basemodel.py:579-599.  `total_times_ms` are cuda.Event.elapsed_time values,
        i.e. MILLISECONDS — the reference prints and stores them as seconds (:595-597), which is why
        data/results.xlsx's `speed` is really ms/image and `fps` images/ms.
        Returns (entire_time, mean per image, std per image, images per time unit) in that same unit."""
        bs = self.setting.batch_size
        n = max(len(total_times_ms), 1)
        entire = float(sum(total_times_ms))
        per = [t / bs for t in total_times_ms]
        mean = sum(per) / n
        std = (sum((p - mean) ** 2 for p in per) / n) ** 0.5
        return entire, mean, std, (num_instances / entire if entire > 0 else float("inf"))

    @torch.no_grad()
    def test(self, dataloader, repetitions=50):
        """The timing half of reference test() (src/basemodel.py:601-722): 50 warm-up forwards on a
        random batch (:644-648), then per batch cuda.Event.record -> forward + argmax -> record ->
        synchronize -> elapsed_time (:659-668).  Returns (accuracy, total_times_ms, images per ms).
        The sklearn classification report / confusion matrix (:696-708) are out of scope.  With
        convnets_b200.enable_eval_graph(model) the forward of a full batch is one CUDA-graph replay."""
        self.eval()
        bs, (ch, h, w) = self.setting.batch_size, self.setting.input_size
        rand_input = self.setting.device.move(torch.randn(bs, ch, h, w, dtype=torch.float))
        fwd = getattr(self, "_cnb_eval_graph", None) or self
        for _ in range(repetitions):
            fwd(rand_input)
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        times, correct, seen = [], 0, 0
        for batch in dataloader:
            inputs, targets = self.setting.device.move(batch)
            start.record()
            outputs = fwd(inputs)
            predictions = outputs.argmax(dim=1)
            end.record()
            torch.cuda.synchronize()
            times.append(start.elapsed_time(end))
            correct += int((predictions == targets).sum().item())
            seen += int(targets.numel())
            if self.setting.sanity_check:
                break
        _, _, _, fps = self.inference_time(times, seen)
        return correct / max(seen, 1), times, fps


class Conv2dBlock(nn.Sequential):
    """Named-children variant of conv2d_block: `conv`, [`bn`], [`relu`] (basemodel.py:1045-1067)."""

    def __init__(self, network, in_channels=None, num_filters=None, set_output=True,
                 activation=True, **kwargs):
        super().__init__()
        self.add_module("conv", network.conv2d(in_channels, num_filters, set_output, **kwargs))
        if network.setting.batch_norm:
            self.add_module("bn", nn.BatchNorm2d(num_features=num_filters))
        if activation:
            self.add_module("relu", nn.ReLU(inplace=True))
