"""ngsolve.ngstd stand-in: Timer only (minres.py:6, bramble_pasciak_cg.py:5). Test infrastructure."""
import time as _time


class Timer:
    def __init__(self, name=""):
        self.name = name
        self.time = 0.0
        self._t0 = None

    def Start(self):
        self._t0 = _time.perf_counter()

    def Stop(self):
        if self._t0 is not None:
            self.time += _time.perf_counter() - self._t0
            self._t0 = None
