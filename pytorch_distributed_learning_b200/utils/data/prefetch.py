"""``CudaPrefetcher``: iterate a DataLoader one batch ahead of the GPU.

The reference's ResNet loop moves every batch with a blocking ``inputs.to(device)`` at the top of the step
(example_launch.py:83, example_mp.py:103), so the GPU idles for the PCIe copy; its ConvNet loop uses ``.cuda(non_blocking=True)`` on
the compute stream (launch_dist.py:82-83), which still serialises copy and compute.  Wrapping the loader --

    for inputs, targets in CudaPrefetcher(loader, device):   # loader built with pin_memory=True
        ...                                                  # batches arrive as device tensors

-- issues the copy of batch i+1 on a copy stream before batch i is handed out, so it overlaps step i.  (A graphed
``TrainStep.step_async`` does the same internally for host batches; this is the counterpart for eager loops.)  On a
CPU device it is a transparent pass-through.
"""
from __future__ import annotations

import torch


def _to_device(obj, device):
    if isinstance(obj, torch.Tensor):
        return obj.to(device, non_blocking=True)
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_device(o, device) for o in obj)
    if isinstance(obj, dict):
        return {k: _to_device(v, device) for k, v in obj.items()}
    return obj


def _record_stream(obj, stream):
    if isinstance(obj, torch.Tensor):
        if obj.is_cuda:
            obj.record_stream(stream)  # allocated on the copy stream, consumed on `stream`
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            _record_stream(o, stream)
    elif isinstance(obj, dict):
        for o in obj.values():
            _record_stream(o, stream)


class CudaPrefetcher:
    def __init__(self, loader, device):
        self.loader = loader
        self.device = torch.device(device)
        self._stream = None

    def __len__(self):
        return len(self.loader)

    @property
    def sampler(self):  # so that ``prefetcher.sampler.set_epoch(e)`` keeps working
        return self.loader.sampler

    def _issue(self, it):
        try:
            batch = next(it)
        except StopIteration:
            return None
        with torch.cuda.stream(self._stream):
            moved = _to_device(batch, self.device)
            ready = torch.cuda.Event()
            ready.record(self._stream)
        return moved, ready

    def __iter__(self):
        if self.device.type != "cuda":
            yield from self.loader
            return
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=self.device)
        it = iter(self.loader)
        nxt = self._issue(it)
        while nxt is not None:
            batch, ready = nxt
            current = torch.cuda.current_stream(self.device)
            current.wait_event(ready)
            _record_stream(batch, current)
            nxt = self._issue(it)  # batch i+1 starts crossing PCIe before the caller launches step i
            yield batch
