"""Double-buffered host -> device input pipeline.  The reference's training loop copies each batch with
`samples.cuda(non_blocking=True)` on the compute stream right before the step (main.py:224-226), so the PCIe copy
(154 MB per 256-image batch, ~2.8 ms) and the step run back to back.  Here batch i+1 travels on a copy stream into the
other of two device buffers while step i computes; the compute stream waits on the copy's event only.

    pf = DevicePrefetcher(device)
    pf.submit(host_batch_0)                  # list of pinned host tensors
    for i in range(steps):
        if i + 1 < steps: pf.submit(host_batch(i + 1))
        batch = pf.next()                    # device tensors of batch i; compute stream ordered after the copy
        loss = step(batch)
        pf.release()                         # batch i's buffer may be overwritten once the work enqueued so far is done
"""
import torch


class DevicePrefetcher:
    def __init__(self, device, depth=2):
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.depth = depth
        self.bufs = [None] * depth
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [None] * depth            # event after the last consumer of the slot
        self.head = 0                         # next slot to fill
        self.tail = 0                         # next slot to hand out
        self.pending = 0

    def submit(self, host):
        if self.pending >= self.depth:
            raise RuntimeError("DevicePrefetcher: all %d buffers are in flight" % self.depth)
        s = self.head
        if self.bufs[s] is None or any(b.shape != h.shape or b.dtype != h.dtype for b, h in zip(self.bufs[s], host)):
            self.bufs[s] = [torch.empty(h.shape, dtype=h.dtype, device=self.device) for h in host]
        with torch.cuda.stream(self.stream):
            if self.free[s] is not None:
                self.stream.wait_event(self.free[s])
            for b, h in zip(self.bufs[s], host):
                b.copy_(h, non_blocking=True)
            self.ready[s].record(self.stream)
        self.head = (s + 1) % self.depth
        self.pending += 1

    def next(self):
        if self.pending == 0:
            raise RuntimeError("DevicePrefetcher: nothing submitted")
        s = self.tail
        torch.cuda.current_stream(self.device).wait_event(self.ready[s])
        self._out = s
        self.tail = (s + 1) % self.depth
        self.pending -= 1
        return self.bufs[s]

    def release(self):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.free[self._out] = ev
