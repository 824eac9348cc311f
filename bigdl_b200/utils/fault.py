"""Fault injection for recovery tests (reference: the ``ExceptionTest`` layer of S/optim/DistriOptimizerSpec.scala, which
throws on chosen forward calls so the driver's retry-from-snapshot loop, DistriOptimizer.scala:881-963, can be
exercised) plus a step watchdog.

``FaultInjector`` is an identity layer that raises on the configured forward call numbers (counted per process, across
retries). ``StepWatchdog`` turns a training step that hangs — a peer that never arrives at a collective — into an
exception on the host after ``timeout`` seconds, which the same retry loop then handles."""
from __future__ import annotations

import threading
import time

from ..nn.module import Module


class InjectedFault(RuntimeError):
    pass


class FaultInjector(Module):
    """Identity; raises ``InjectedFault`` when the running count of training forwards hits one of ``fail_at``."""

    calls = 0                      # process-wide, survives model reloads like the reference's static counter

    def __init__(self, fail_at=(), message: str = "injected fault"):
        super().__init__()
        self.fail_at, self.message = {int(v) for v in fail_at}, message

    @classmethod
    def reset(cls):
        cls.calls = 0

    def updateOutput(self, input):
        if self.isTraining():
            FaultInjector.calls += 1
            if FaultInjector.calls in self.fail_at:
                raise InjectedFault(f"{self.message} (forward call {FaultInjector.calls})")
        return input

    def updateGradInput(self, input, gradOutput):
        return gradOutput


class StepWatchdog:
    """``with StepWatchdog(timeout) as w: ... w.kick()`` — a daemon thread raises the alarm callback (default: records
    the expiry so the training loop can raise) if ``kick()`` is not called for ``timeout`` seconds."""

    def __init__(self, timeout: float, on_expire=None):
        self.timeout, self.on_expire = float(timeout), on_expire
        self.expired = False
        self._last = time.monotonic()
        self._stop = threading.Event()
        self._thread = None

    def kick(self):
        self._last = time.monotonic()

    def check(self):
        if self.expired:
            raise TimeoutError(f"no training step completed within {self.timeout:.1f} s")

    def _run(self):
        while not self._stop.wait(min(0.05, self.timeout / 4)):
            if time.monotonic() - self._last > self.timeout:
                self.expired = True
                if self.on_expire is not None:
                    self.on_expire()
                return

    def __enter__(self):
        self._last = time.monotonic()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=1.0)
        return False
