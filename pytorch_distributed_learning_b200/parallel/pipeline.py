"""Host <-> device plumbing of a training step that must not stall the GPU.

The reference's loop (launch_dist.py:82-92) copies the batch to the GPU with ``non_blocking=True`` and reads
``loss.item()`` only every 100th step; between two steps the GPU still idles for the PCIe copy of the next batch,
and every ``item()`` drains the device.  Two small pieces remove both stalls without changing what is computed:

* ``InputStager`` -- the H2D copy of batch i+1 runs on a copy stream while step i computes (double-buffered
  device landing zones, ordered against their consumers with events);
* ``LossRing`` / ``PendingLoss`` -- the scalar loss of every step is copied to a ring of pinned host slots right
  behind the step; ``PendingLoss.item()`` waits for exactly that copy, not for the device.  Read it one step late
  and the host never blocks the GPU.
"""
from __future__ import annotations

import torch


def refresh_inputs(x_dst, y_dst, images, labels):
    """Copy a device-resident batch into a graph's static inputs on the current stream: one kernel for both tensors when
    they are plain contiguous CUDA tensors of the destination's dtype (two ``copy_`` calls otherwise)."""
    if (images.is_cuda and labels.is_cuda and images.dtype == x_dst.dtype and labels.dtype == y_dst.dtype
            and images.is_contiguous() and labels.is_contiguous() and images.numel() == x_dst.numel() and labels.numel() == y_dst.numel()
            and images.device == x_dst.device and labels.device == y_dst.device):
        import ctypes

        from .. import _lib

        L = _lib.cuda()
        stream = ctypes.c_void_p(torch.cuda.current_stream(x_dst.device).cuda_stream)
        rc = L.pdl_copy_pair(ctypes.c_void_p(x_dst.data_ptr()), ctypes.c_void_p(images.data_ptr()), images.numel() * images.element_size(),
                             ctypes.c_void_p(y_dst.data_ptr()), ctypes.c_void_p(labels.data_ptr()), labels.numel() * labels.element_size(), stream)
        if rc != 0:
            raise RuntimeError("input refresh kernel failed: %s" % L.pdl_last_error().decode())
        return
    x_dst.view(images.shape).copy_(images, non_blocking=True)
    y_dst.copy_(labels, non_blocking=True)


class PendingLoss:
    """Loss of one training step on its way to (or already in) pinned host memory."""

    __slots__ = ("_ring", "_slot", "_serial")

    def __init__(self, ring, slot, serial):
        self._ring, self._slot, self._serial = ring, slot, serial

    def ready(self):
        return self._ring.events[self._slot].query()

    def item(self):
        ring = self._ring
        if ring.serial - self._serial > ring.DEPTH:  # the publish numbered serial + DEPTH has taken this slot
            raise RuntimeError("PendingLoss expired: its host slot was reused (%d steps were issued since; the ring keeps %d)"
                               % (ring.serial - self._serial, ring.DEPTH))
        ring.events[self._slot].synchronize()
        return float(ring.host[self._slot])

    __float__ = item

    def __repr__(self):
        return "PendingLoss(step=%d, ready=%s)" % (self._serial, self.ready())


class LossRing:
    DEPTH = 64

    def __init__(self, device):
        self.device = device
        self.host = torch.zeros(self.DEPTH, dtype=torch.float32).pin_memory()
        self.events = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.serial = 0

    def publish(self, loss, stream, after=None):
        """Queue ``host[slot] <- loss`` on ``stream`` (after event ``after``, if given); returns (PendingLoss, event)."""
        slot = self.serial % self.DEPTH
        if after is not None:
            stream.wait_event(after)
        with torch.cuda.stream(stream):
            self.host[slot:slot + 1].copy_(loss.detach().reshape(1), non_blocking=True)
            self.events[slot].record(stream)
        pending = PendingLoss(self, slot, self.serial)
        self.serial += 1
        return pending, self.events[slot]


class InputStager:
    """Two device landing zones for (images, labels) filled on a copy stream.

    ``stage(images, labels)`` queues the H2D copies of a host batch on the copy stream and returns
    ``(x, y, ready, slot)``: device tensors plus the event the consumer stream must wait for.  After the consumer has
    queued its last read of ``x``/``y`` it calls ``consumed(slot, stream)``; the copy that next reuses the zone waits for
    that.  With the host running ahead of the GPU, batch i+1 crosses PCIe while step i computes.
    """

    def __init__(self, device, zones=None):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self.zones = zones  # [(x, y), (x, y)] supplied by the owner, or allocated on first use
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.next = 0

    def stage(self, images, labels):
        if self.zones is None:
            self.zones = [(torch.empty(images.shape, dtype=images.dtype, device=self.device),
                           torch.empty(labels.shape, dtype=labels.dtype, device=self.device)) for _ in range(2)]
        slot, self.next = self.next, self.next ^ 1
        x, y = self.zones[slot]
        cs = self.stream
        cs.wait_event(self.free[slot])  # never recorded yet == nothing to wait for
        with torch.cuda.stream(cs):
            x.view(images.shape).copy_(images, non_blocking=True)
            y.copy_(labels, non_blocking=True)
            self.ready[slot].record(cs)
        return x, y, self.ready[slot], slot

    def consumed(self, slot, stream):
        self.free[slot].record(stream)
