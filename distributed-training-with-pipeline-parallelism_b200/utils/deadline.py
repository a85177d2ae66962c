"""A wall-clock deadline around a block that may never return (a collective whose peer stalled, a kernel that spins on
a flag).  The reference's launcher bounds a whole worker with ``join(timeout)`` (nb:323-331); this
bounds ONE optional section of a worker so that what was measured before it can still be reported.

    with Deadline(120.0, on_timeout) as dl:
        optional_check()              # C-level waits release the GIL, so the watchdog thread keeps running
    if not dl.completed: ...          # on_timeout() won the race and owns the process from here

Exactly one of the two sides wins: either the block finishes first (``completed`` is True, the handler never runs) or
the handler runs (typically: print what is already known, ``os._exit(0)``)."""
from __future__ import annotations

import threading
from typing import Callable


class Deadline:
    def __init__(self, seconds: float, on_timeout: Callable[[], None]):
        self.seconds = float(seconds)
        self.on_timeout = on_timeout
        self.completed = False
        self._done = threading.Event()
        self._fired = False
        self._lock = threading.Lock()
        self._thread = threading.Thread(target=self._watch, name="ppb200-deadline", daemon=True)

    def _watch(self) -> None:
        if self._done.wait(self.seconds):
            return
        with self._lock:
            if self._done.is_set():
                return
            self._fired = True
        self.on_timeout()

    def __enter__(self) -> "Deadline":
        self._thread.start()
        return self

    def __exit__(self, exc_type, exc, tb) -> bool:
        with self._lock:
            self.completed = not self._fired
            self._done.set()
        return False
