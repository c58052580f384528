"""Input pipeline of the training / evaluation loops: ``scripts/train.py:176-183`` and ``evaluate.py:205-217``
take ``(constants, prescribed, prognostic, target)`` batches from a ``DataLoader`` and call ``.to(device)`` on each tensor in
front of the model call -- four pageable-memory copies the GPU waits for, every step.

``DevicePrefetcher`` wraps the same iterable: a background thread copies each batch into pinned staging buffers (reused per
slot: no ``cudaHostAlloc`` in the loop) and a side stream brings it to the device ``depth`` batches ahead of the consumer, so
the copies run under the previous step's kernels.  The reference's conventions are kept: a tensor that holds NaNs stands for
"this model has no such input" (``train.py:180-181``) and is handed on as ``None``; everything else arrives as a device tensor
with the batch's dtype.  Iterating yields tuples in the loader's order.

    for constants, prescribed, prognostic, target in DevicePrefetcher(train_dataloader, device):
        output = model(constants=constants, prescribed=prescribed, prognostic=prognostic)

There is no host fallback: without a CUDA device the constructor raises (as the modules do).
"""
from __future__ import annotations

import queue
import threading
from typing import Iterable, Iterator, Sequence

import torch

from . import native


class DevicePrefetcher:
    def __init__(self, loader: Iterable, device="cuda", depth: int = 2, nan_means_absent: bool = True):
        self.device = torch.device(device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise native.NativeError("DevicePrefetcher copies to a CUDA device (no CPU fallback)")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        if depth < 1:
            raise ValueError("depth must be at least 1")
        self.loader, self.depth, self.nan_means_absent = loader, int(depth), bool(nan_means_absent)
        self._stream = torch.cuda.Stream(self.device)
        self._staging = [dict() for _ in range(self.depth + 1)]      # slot -> {position in the batch: pinned buffer}

    def __len__(self):
        return len(self.loader)

    def _pinned(self, slot: int, pos: int, t: torch.Tensor) -> torch.Tensor:
        buf = self._staging[slot].get(pos)
        if buf is None or buf.shape != t.shape or buf.dtype != t.dtype:
            buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            self._staging[slot][pos] = buf
        buf.copy_(t)
        return buf

    def _stage(self, slot: int, batch: Sequence) -> list:
        """Host side of one batch (worker thread): NaN check, copy into this slot's pinned buffers."""
        out = []
        for pos, t in enumerate(batch):
            if not torch.is_tensor(t):
                out.append(t)
            elif self.nan_means_absent and t.is_floating_point() and bool(torch.isnan(t).any()):
                out.append(None)
            elif t.device.type == "cuda":
                out.append(t)
            else:
                out.append(self._pinned(slot, pos, t))
        return out

    def __iter__(self) -> Iterator[tuple]:
        q: "queue.Queue" = queue.Queue(maxsize=self.depth)
        free = threading.Semaphore(self.depth + 1)       # staging slots the consumer has released
        stop = threading.Event()
        done = object()

        def worker():
            try:
                torch.cuda.set_device(self.device)
                for n, batch in enumerate(self.loader):
                    free.acquire()
                    if stop.is_set():
                        return
                    slot = n % (self.depth + 1)
                    staged = self._stage(slot, batch if isinstance(batch, (tuple, list)) else (batch,))
                    with torch.cuda.stream(self._stream):
                        dev = [t.to(self.device, non_blocking=True) if torch.is_tensor(t) and t.device.type == "cpu" else t for t in staged]
                        ev = torch.cuda.Event()
                        ev.record(self._stream)
                    q.put((dev, ev))
                q.put(done)
            except BaseException as exc:      # hand the loader's exception to the consumer
                q.put(exc)

        th = threading.Thread(target=worker, name="dlwp-prefetch", daemon=True)
        th.start()
        try:
            while True:
                item = q.get()
                if item is done:
                    break
                if isinstance(item, BaseException):
                    raise item
                dev, ev = item
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)
                for t in dev:
                    if torch.is_tensor(t):
                        t.record_stream(cur)      # the caching allocator must not hand the block out while the consumer's kernels run
                # the slot's pinned buffers are free once its copies have run: the event is complete when the consumer's
                # stream has passed the wait, and the worker only reuses a slot depth + 1 batches later
                ev.synchronize()
                free.release()
                yield tuple(dev)
        finally:
            stop.set()
            free.release()
            # drain so that a worker blocked on the queue can finish
            while th.is_alive():
                try:
                    q.get(timeout=0.05)
                except queue.Empty:
                    pass
