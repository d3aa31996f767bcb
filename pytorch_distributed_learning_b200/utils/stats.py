"""Per-step statistics: device-timed step, exposed communication, bus bandwidth."""
from __future__ import annotations

import torch


def busbw_gbs(nbytes, seconds, world):
    """Allreduce bus bandwidth (NCCL convention): S/t * 2(W-1)/W, in GB/s."""
    if world <= 1 or seconds <= 0:
        return 0.0
    return nbytes / seconds * 2.0 * (world - 1) / world / 1e9


class CudaStepTimer:
    """CUDA-event timer on the current stream; ``ms()`` synchronises."""

    def __init__(self):
        self.start = torch.cuda.Event(enable_timing=True)
        self.stop = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.start.record()
        return self

    def __exit__(self, *exc):
        self.stop.record()

    def ms(self):
        self.stop.synchronize()
        return self.start.elapsed_time(self.stop)


def max_over_ranks(value, group=None):
    """Max of a python float over the group (device-timed numbers are reported as the max over ranks)."""
    from ..parallel import distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(value)
    pg = dist._resolve(group)
    return max(pg.allgather_object(float(value)))


class RunningMetrics:
    """Running loss / top-1 accuracy of a training loop, accumulated on the device.

    The reference's ResNet loop computes ``loss.item()`` and ``eq(outputs.argmax(1), targets).sum().item()`` after every
    step (example_launch.py:91-93) -- two host syncs per iteration -- to print a line every 25 steps.  ``update`` is one
    small kernel and no sync; ``read`` (one D2H copy) is called when a line is actually printed::

        metrics = RunningMetrics(device)
        for step, (x, y) in enumerate(loader):
            out = net(x); loss = criterion(out, y); ...
            metrics.update(out, y, loss)
            if (step + 1) % 25 == 0:
                mean_loss, acc, seen = metrics.read()

    CPU tensors take the obvious PyTorch path, so the class works in the CPU tests as well.
    """

    def __init__(self, device=None):
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.acc = torch.zeros(3, dtype=torch.float64, device=self.device)  # loss sum, correct, samples
        self.steps = 0

    def reset(self):
        self.acc.zero_()
        self.steps = 0

    @torch.no_grad()
    def update(self, outputs, targets, loss=None):
        self.steps += 1
        outputs = outputs.detach()
        if outputs.is_cuda and outputs.dtype == torch.float32 and outputs.dim() == 2 and targets.dtype == torch.int64:
            from .. import _lib
            from ..ops.functional import _ok, _p, _stream

            outputs, targets = outputs.contiguous(), targets.contiguous()
            lp = None
            if loss is not None:
                loss = loss.detach()
                if not (loss.is_cuda and loss.dtype == torch.float32 and loss.numel() == 1):
                    loss = torch.as_tensor(float(loss), dtype=torch.float32, device=outputs.device)
                lp = _p(loss)
            _ok(_lib.cuda().pdl_step_metrics(_p(outputs), _p(targets), lp, outputs.shape[0], outputs.shape[1], _p(self.acc), _stream(outputs)),
                "step metrics")
            return
        if loss is not None:
            self.acc[0] += float(loss)
        self.acc[1] += float((outputs.argmax(dim=1) == targets).sum())
        self.acc[2] += outputs.shape[0]

    def read(self):
        """(mean loss per step, top-1 accuracy in [0, 1], samples seen); synchronises on the accumulator only."""
        loss_sum, correct, seen = self.acc.tolist()
        return loss_sum / max(self.steps, 1), (correct / seen if seen else 0.0), int(seen)
