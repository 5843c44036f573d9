"""Host-side watchdog for the peer collectives.

The device side never spins for ever: every cross-GPU wait is bounded and a timed-out barrier leaves ``1 + peer`` in the
communicator's status word (``csrc/comm/collectives.cu``).  Somebody still has to look at that word.  The trainers call
``check_health()`` once per epoch; this watchdog does it every ``interval_s`` seconds from a daemon thread, on a stream of its
own so that it never waits behind the training stream, and hands the first fault to ``on_fault`` (default: log it and
terminate the process with exit code 70, so that the launcher tears the job down instead of leaving the healthy ranks
training against garbage).  The reference has no failure handling beyond ``mp.spawn`` propagating a child's exception
(model_parallel.py:162).
"""
from __future__ import annotations

import os
import sys
import threading
from typing import Callable, Optional


class CommWatchdog:
    """Polls ``probe()`` (raises on a fault) every ``interval_s`` seconds until ``stop()``."""

    def __init__(self, probe: Callable[[], None], interval_s: float = 5.0, on_fault: Optional[Callable[[BaseException], None]] = None,
                 name: str = "dmp-comm-watchdog"):
        self.probe, self.interval_s = probe, float(interval_s)
        self.on_fault = on_fault if on_fault is not None else self._default_on_fault
        self.fault: Optional[BaseException] = None
        self.polls = 0
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name=name, daemon=True)

    @staticmethod
    def _default_on_fault(exc: BaseException) -> None:
        print("[dmp watchdog] %s -- terminating this rank" % exc, file=sys.stderr, flush=True)
        os._exit(70)

    def _run(self):
        while not self._stop.wait(self.interval_s):
            try:
                self.probe()
                self.polls += 1
            except BaseException as e:  # noqa: BLE001  (a fault is whatever the probe raises)
                self.fault = e
                try:
                    self.on_fault(e)
                finally:
                    return

    def start(self) -> "CommWatchdog":
        self._thread.start()
        return self

    def stop(self, timeout: float = 2.0) -> None:
        self._stop.set()
        if self._thread.is_alive():
            self._thread.join(timeout)

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()
        return False
