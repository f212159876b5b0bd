"""A second CUDA stream for the weight-gradient GEMMs of a hand-written backward pass.

In the backward pass of a layer, ``dW = dZ^T X`` and ``dX = dZ W`` are independent; only ``dX`` is on the critical path to
the next layer - the weight gradients are needed by the optimizer alone.  Issuing every weight-gradient GEMM on a side
stream (fork after ``dZ`` exists, one join before the optimizer) turns the step's chain of ~24 dependent launches into a
shorter main chain with a parallel branch; under stream capture the fork / join events become edges of the CUDA graph.
"""

from __future__ import annotations

from typing import Callable, Optional

import torch


class WgradBranch:
    def __init__(self, device: torch.device, enabled: bool = True) -> None:
        self.device = device
        self.enabled = enabled and device.type == "cuda"
        self._side: Optional[torch.cuda.Stream] = None
        self._dirty = False

    def run(self, fn: Callable[[], None]) -> None:
        """Launch ``fn``'s kernels after everything issued so far on the current stream, but off its critical path."""
        if not self.enabled:
            fn()
            return
        main = torch.cuda.current_stream(self.device)
        if self._side is None:
            self._side = torch.cuda.Stream(self.device)
        ev = torch.cuda.Event()
        ev.record(main)
        self._side.wait_event(ev)
        with torch.cuda.stream(self._side):
            fn()
        self._dirty = True

    def join(self) -> None:
        """The current stream waits for everything launched through :meth:`run` (call before the optimizer step)."""
        if self._dirty:
            ev = torch.cuda.Event()
            ev.record(self._side)
            torch.cuda.current_stream(self.device).wait_event(ev)
            self._dirty = False
