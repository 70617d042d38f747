"""soil.timer / soil.duration (reference: source/soillib/util/timer.hpp:15-69, python/source/util.cpp:47-73).

The reference timer is wall-clock only, so around asynchronous launches it measures
launch time.  This one keeps that behaviour by default and, with sync=True,
brackets the region with a device synchronise so the count is the GPU time.
"""
from __future__ import annotations

import enum
import time


class duration(enum.Enum):
    s = 1.0
    ms = 1e3
    us = 1e6
    ns = 1e9


s, ms, us, ns = duration.s, duration.ms, duration.us, duration.ns


class timer:
    def __init__(self, unit: duration = duration.ms, sync: bool = False):
        self.unit, self.sync, self.count, self._t0 = unit, sync, 0, 0.0

    def _sync(self):
        if self.sync:
            import torch
            if torch.cuda.is_available():
                torch.cuda.synchronize()

    def start(self):
        self._sync()
        self._t0 = time.perf_counter()

    def stop(self):
        self._sync()
        self.count = int((time.perf_counter() - self._t0) * self.unit.value)
        return self.count

    def __enter__(self):
        self.start()
        return self

    def __exit__(self, *exc):
        self.stop()
        return False
