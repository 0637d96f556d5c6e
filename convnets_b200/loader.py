"""Device prefetch for the training loop's input batches.

The reference moves each batch to the device at the top of the step
(`forward_step` -> `DeviceMngr.move`, reference src/basemodel.py:185-186, src/mngrdevice.py:30-36),
on the compute stream, so the 154 MB fp32 copy of a 256 x 3 x 224 x 224 batch sits on the step's
critical path.  `PrefetchLoader` wraps any iterable of (inputs, targets) host batches (a
`torch.utils.data.DataLoader` with `pin_memory=True`, or a list of pinned tensors) and yields
device tensors instead: batch i+1 is copied on a side stream while step i computes, and the
compute stream only waits for the copy event.  `train_model` stays unmodified — `move` on a
tensor that already lives on the device is a no-op — and every batch still crosses PCIe once.
"""
from __future__ import annotations

import torch


class PrefetchLoader:
    def __init__(self, loader, device, depth: int = 2):
        self.loader = loader
        self.device = torch.device(device)
        self.depth = max(1, int(depth))
        self.stream = torch.cuda.Stream(device=self.device)
        if hasattr(loader, "dataset"):
            self.dataset = loader.dataset  # train_model divides by len(dataloader.dataset)

    def __len__(self):
        return len(self.loader)

    def _issue(self, batch):
        with torch.cuda.stream(self.stream):
            moved = [t.to(self.device, non_blocking=True) for t in batch]
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
