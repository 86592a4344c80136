"""Host-buffer front end of the attack path: a stream of pinned host batches goes through an attack with the
host->device copy of batch i+1 and the device->host copy of result i-1 overlapped with the attack on batch i
(three CUDA streams, two device input buffers, events instead of synchronisation).

The reference loads one image at a time and moves it with `.to(device)` (ResNet.py:62-63,
defense_experiments.py:649-662) and reads results back with `.cpu()`; this is the batched, asynchronous
equivalent for callers that hold their images in host memory.  There is no CPU fallback: the device work is
the native attack that `attack_fn` wraps."""
from __future__ import annotations

from typing import Callable, Iterable, Iterator, Optional

import torch


class HostBatchPipeline:
    """Runs `attack_fn(x_dev) -> x_adv_dev` over pinned host batches with copy/compute overlap.

    attack_fn    : callable taking a CUDA fp32 [N,3,H,W] tensor, returning a CUDA tensor of the same shape
                   (e.g. a functools.partial of advb200.pgd_linf_attack with the labels of the batch bound).
    device       : CUDA device the attack runs on.
    Every batch must have the shape of the first one (the device buffers are allocated once)."""

    def __init__(self, attack_fn: Callable[[torch.Tensor], torch.Tensor], device) -> None:
        self.fn = attack_fn
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostBatchPipeline needs a CUDA device (no CPU fallback)")
        self.h2d = torch.cuda.Stream(device=self.device)
        self.d2h = torch.cuda.Stream(device=self.device)
        self._in = [None, None]            # device input buffers
        self._in_free = [None, None]       # event: the attack that read buffer k has finished
        self._in_ready = [None, None]      # event: the H2D copy into buffer k has landed
        self._out = [None, None]           # device staging buffers the D2H copies read from
        self._out_free = [None, None]      # event: the D2H copy out of staging buffer k has finished

    def _stage(self, k: int, x_host: torch.Tensor) -> None:
        if not x_host.is_pinned():
            raise ValueError("host batches must live in pinned memory (tensor.pin_memory())")
        if self._in[k] is None:
            self._in[k] = torch.empty(x_host.shape, dtype=torch.float32, device=self.device)
        with torch.cuda.stream(self.h2d):
            if self._in_free[k] is not None:
                self.h2d.wait_event(self._in_free[k])
            self._in[k].copy_(x_host, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.h2d)
            self._in_ready[k] = ev

    def run(self, batches: Iterable[torch.Tensor], outs: Optional[Iterable[torch.Tensor]] = None) -> Iterator[torch.Tensor]:
        """Yields one pinned host tensor of adversarial images per input batch, in order.  `outs` optionally
        supplies the pinned result buffers (otherwise they are allocated; two buffers used in turn are enough —
        batch i-2 is always handed out before batch i is written).  A yielded tensor is complete."""
        compute = torch.cuda.current_stream(self.device)
        it = iter(batches)
        out_it = iter(outs) if outs is not None else None
        nxt = next(it, None)
        if nxt is None:
            return
        self._stage(0, nxt)
        i = 0
        # Results are handed out TWO batches late: the host then never waits on a copy that finished less than a
        # whole attack ago, and it enqueues the launches of batch i while the device is still working on batch
        # i-1.  (Handing them out one batch late kept the host only one attack ahead of the device; an occasional
        # host-side stall of a few tens of ms then starved the device: 1 run in 8 lost ~10 %.)
        pending = []                        # [(host result, its D2H event)], oldest first, at most two
        while nxt is not None:
            k = i & 1
            cur = nxt
            nxt = next(it, None)
            if nxt is not None:
                self._stage(k ^ 1, nxt)     # next batch's H2D overlaps this batch's attack
            if len(pending) == 2:           # retire batch i-2 first: its host buffer may be the one batch i reuses
                host_done, ev_done = pending.pop(0)
                ev_done.synchronize()
                yield host_done
            compute.wait_event(self._in_ready[k])
            x_adv = self.fn(self._in[k])
            # The result is copied (device to device, ~50 us) into a staging buffer this pipeline owns and the D2H
            # reads from there.  Handing the attack's own output tensor to the copy stream (record_stream) makes
            # the caching allocator wait on a cross-stream event before it can reuse the block; when that event is
            # late it falls back to cudaMalloc, which synchronises the device.
            if self._out[k] is None:
                self._out[k] = torch.empty_like(x_adv)
            if self._out_free[k] is not None:
                compute.wait_event(self._out_free[k])
            self._out[k].copy_(x_adv, non_blocking=True)
            del x_adv
            done = torch.cuda.Event()
            done.record(compute)
            self._in_free[k] = done
            host = next(out_it) if out_it is not None else torch.empty(cur.shape, dtype=torch.float32).pin_memory()
            with torch.cuda.stream(self.d2h):
                self.d2h.wait_event(done)
                host.copy_(self._out[k], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.d2h)
            self._out_free[k] = ev
            pending.append((host, ev))
            i += 1
        for host_done, ev_done in pending:
            ev_done.synchronize()
            yield host_done

    def synchronize(self) -> None:
        self.h2d.synchronize()
        self.d2h.synchronize()
        torch.cuda.current_stream(self.device).synchronize()
