"""Device prefetch for the training loop's input batches.

The reference moves each batch to the device at the top of the step
(`forward_step` -> `DeviceMngr.move`, reference src/basemodel.py:185-186, src/mngrdevice.py:30-36),
on the compute stream, so the 154 MB fp32 copy of a 256 x 3 x 224 x 224 batch sits on the step's
critical path.  `PrefetchLoader` wraps any iterable of (inputs, targets) host batches (a
`torch.utils.data.DataLoader` with `pin_memory=True`, or a list of pinned tensors) and yields
device tensors instead: batch i+1 is copied on a side stream while step i computes, and the
compute stream only waits for the copy event.  `train_model` stays unmodified — `move` on a
tensor that already lives on the device is a no-op — and every batch still crosses PCIe once.

With `normalize=(mean, std)` the loader accepts uint8 image batches ([N, C, H, W], or [N, H, W, C]
with nhwc=True — what the dataset holds before transforms.ToTensor) and runs ToTensor + Normalize
(+ RandomCrop / RandomHorizontalFlip index arithmetic) on the device, on the copy stream
(`functional.normalize_u8`): the bus carries a quarter of the bytes (38.5 MB instead of 154 MB for
256 x 3 x 224 x 224) and the model still receives the fp32 NCHW tensor the reference's DataLoader
would have produced (src/mngrdata.py:139-165), bit for bit.  RandomAffine (:148) stays a host
transform (resampling, not index arithmetic): apply it to the uint8 image before batching.
"""
from __future__ import annotations

import torch


class PrefetchLoader:
    def __init__(self, loader, device, depth: int = 2, normalize=None, nhwc: bool = False,
                 crop=None, padding: int = 0, flip_p: float = 0.0, generator=None):
        self.loader = loader
        self.device = torch.device(device)
        self.depth = max(1, int(depth))
        self.stream = torch.cuda.Stream(device=self.device)
        self.normalize = normalize          # (mean, std) per channel, or None: batches pass through
        self.nhwc = bool(nhwc)
        self.crop, self.padding, self.flip_p = crop, int(padding), float(flip_p)
        self.generator = generator
        if hasattr(loader, "dataset"):
            self.dataset = loader.dataset  # train_model divides by len(dataloader.dataset)

    def __len__(self):
        return len(self.loader)

    def draw_augmentation(self, n, h, w):
        """Per image, in torchvision's order: RandomCrop.get_params (top, then left; uniform over the
        padded image), then RandomHorizontalFlip's `rand(1) < p`.  int32 [n, 3] on the host."""
        th = tw = int(self.crop) if self.crop else None
        if th is None:
            th, tw = h, w
        hp, wp = h + 2 * self.padding, w + 2 * self.padding
        aug = torch.zeros((n, 3), dtype=torch.int32)
        g = self.generator
        for i in range(n):
            if self.crop:
                aug[i, 0] = int(torch.randint(0, hp - th + 1, (1,), generator=g))
                aug[i, 1] = int(torch.randint(0, wp - tw + 1, (1,), generator=g))
            if self.flip_p > 0:
                aug[i, 2] = int(float(torch.rand(1, generator=g)) < self.flip_p)
        return aug

    def _issue(self, batch):
        with torch.cuda.stream(self.stream):
            moved = [t.to(self.device, non_blocking=True) for t in batch]
            if self.normalize is not None and moved[0].dtype == torch.uint8:
                from . import functional as F
                x = moved[0]
                n = x.shape[0]
                h, w = (x.shape[1], x.shape[2]) if self.nhwc else (x.shape[2], x.shape[3])
                aug = None
                if self.crop or self.flip_p > 0:
                    aug = self.draw_augmentation(n, h, w).pin_memory().to(self.device, non_blocking=True)
                moved[0] = F.normalize_u8(x, self.normalize[0], self.normalize[1], nhwc=self.nhwc,
                                          crop=self.crop, padding=self.padding, aug=aug)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return moved, ev

    def __iter__(self):
        it = iter(self.loader)
        queue = []
        cur = torch.cuda.current_stream(self.device)
        try:
            while len(queue) < self.depth:
                queue.append(self._issue(next(it)))
        except StopIteration:
            it = None
        while queue:
            moved, ev = queue.pop(0)
            cur.wait_event(ev)
            for t in moved:
                t.record_stream(cur)  # the caching allocator must not recycle it under the step
            if it is not None:
                try:
                    queue.append(self._issue(next(it)))
                except StopIteration:
                    it = None
            yield moved
