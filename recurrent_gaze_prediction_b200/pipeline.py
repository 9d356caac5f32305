"""Double-buffered host -> device feeding of the model's input "placeholders".

The reference feeds every ``session.run`` from host NumPy arrays (``gaze_rnn.py:520-531``: 0.38 GB per step at B=28,
1.9 GB at B=256).  Over PCIe that copy is several times longer than the B200 step itself, so the feeder uploads the
NEXT step's batch on a copy stream into a staging slot while the compute stream runs the CURRENT step; a device-to-
device move (0.3 ms/GB) brings a ready slot into the buffers the captured CUDA graph reads.  Steady-state step time is
max(H2D, compute) instead of their sum.  Plumbing only: torch streams / events / copies, no arithmetic.
"""
from __future__ import annotations

import torch


class PrefetchFeeder(object):
    def __init__(self, dst_tensors, device=None, slots=2):
        """dst_tensors: the device tensors the model reads (e.g. [model.c3d_input, model.gt_gazemap])."""
        self.dst = list(dst_tensors)
        self.device = device or self.dst[0].device
        self.slots = slots
        self.stage = [[torch.empty_like(t) for t in self.dst] for _ in range(slots)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.ready = [torch.cuda.Event() for _ in range(slots)]
        self.free = [torch.cuda.Event() for _ in range(slots)]
        self.n_prefetched = 0
        self.n_acquired = 0

    def prefetch(self, host_tensors):
        """Start the H2D copy of one batch (pinned host tensors, same shapes as dst) into the next free slot."""
        if self.n_prefetched - self.n_acquired >= self.slots:
            raise RuntimeError("all %d staging slots are in flight; call acquire() first" % self.slots)
        k = self.n_prefetched % self.slots
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.free[k])
            for d, h in zip(self.stage[k], host_tensors):
                d.copy_(h, non_blocking=True)
            self.ready[k].record(self.copy_stream)
        self.n_prefetched += 1

    def acquire(self):
        """On the current (compute) stream: wait for the oldest prefetched batch and move it into dst."""
        if self.n_acquired >= self.n_prefetched:
            raise RuntimeError("acquire() without a prefetched batch")
        k = self.n_acquired % self.slots
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.ready[k])
        for d, s in zip(self.dst, self.stage[k]):
            d.copy_(s, non_blocking=True)
        self.free[k].record(cur)
        self.n_acquired += 1

    def drain(self):
        """Drop batches that were prefetched but never consumed (end of a run)."""
        self.copy_stream.synchronize()
        self.n_acquired = self.n_prefetched
