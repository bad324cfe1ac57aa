"""Host -> device batch prefetcher: overlaps the pinned-memory copy of batch i+1 with the training step of batch i.

The reference feeds the network synchronously: `prep_data` copies every batch with `.to(device)` right before
`model.forward` (training/train_ubresnet2018_wlarcv2.py:322-332), so the PCIe copy (134 MB for 32 crops of
image + int64 label + weight) sits on the critical path.  Here the copy of the next batch runs on a side stream into a
second set of device buffers while the current step computes; the consumer only waits on the copy's event.
"""
import torch


class DevicePrefetcher:
    """Wraps an iterator of tuples of pinned CPU tensors; yields tuples of device tensors.

    Double buffered: at most two batches are resident on the device.  Every batch is copied exactly once, on
    `self.stream`; `__next__` makes the current stream wait for that copy and immediately queues the following one."""

    def __init__(self, batches, device, depth=2):
        self.it = iter(batches)
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.depth = depth
        self.slots = []            # ring of (tensors, copy-done event, consumed event)
        self.pending = []
        for _ in range(depth - 1):
            self._issue()

    def _issue(self):
        try:
            host = next(self.it)
        except StopIteration:
            return
        for t in host:
            if not t.is_pinned():
                raise RuntimeError("DevicePrefetcher: host tensors must be pinned (tensor.pin_memory())")
        with torch.cuda.stream(self.stream):
            dev = tuple(t.to(self.device, non_blocking=True) for t in host)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.pending.append((dev, ev))

    def __iter__(self):
        return self

    def __next__(self):
        if not self.pending:
            self._issue()
        if not self.pending:
            raise StopIteration
        dev, ev = self.pending.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for t in dev:
            t.record_stream(cur)       # the caching allocator must not reuse the buffer before the step has read it
        self._issue()                  # next batch starts copying while this one is being used
        return dev


class AssembledPrefetcher:
    """Host wire-format batches -> ASSEMBLED device batches, one step ahead.

    Wraps an iterator of (image planes, ROI-label planes) pinned host tensors and a `fill.DeviceBatchFiller`: the copy of
    batch i+1 AND its device-side assembly (transpose, ROI -> class table, pixel weights: csrc/batch_fill.cu) run on the
    prefetch stream while step i computes, into one of two staging sets of (image, label, weight).  `__next__` makes the
    current stream wait for the set and returns it; a set is refilled only after everything the consumer enqueued on it
    (up to its next `__next__` call) has finished."""

    def __init__(self, batches, filler, rows, cols, batch, device):
        self.it = iter(batches)
        self.filler, self.rows, self.cols = filler, rows, cols
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        mk = lambda dt: torch.empty((batch, rows, cols), dtype=dt, device=self.device)
        self.sets = [(torch.empty((batch, 1, rows, cols), dtype=torch.float32, device=self.device), mk(torch.int64),
                      mk(torch.float32)) for _ in range(2)]
        self.free_ev = [None, None]        # main-stream event after which set k may be overwritten
        self.pending = []
        self.k = 0
        self._issue()

    def _issue(self):
        try:
            planes, rois = next(self.it)
        except StopIteration:
            return
        k = self.k
        self.k ^= 1
        with torch.cuda.stream(self.stream):
            if self.free_ev[k] is not None:
                self.stream.wait_event(self.free_ev[k])
            dp = planes.to(self.device, non_blocking=True)
            dr = rois.to(self.device, non_blocking=True)
            self.filler.fill(dp, dr, self.rows, self.cols, out=self.sets[k])
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.pending.append((k, ev))

    def __iter__(self):
        return self

    def __next__(self):
        if not self.pending:
            self._issue()
        if not self.pending:
            raise StopIteration
        k, ev = self.pending.pop(0)
        cur = torch.cuda.current_stream(self.device)
        # everything enqueued so far has consumed the OTHER set (handed out by the previous call): it may be refilled
        done = torch.cuda.Event()
        done.record(cur)
        self.free_ev[k ^ 1] = done
        cur.wait_event(ev)
        self._issue()                      # fills set k^1 for the following step, after `done`
        return self.sets[k]
