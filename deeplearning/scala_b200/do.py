"""A synchronous model of RAII.scala's `Do` (com.thoughtworks.raii.asynchronous.Do): a computation that yields a value
together with the continuation that releases it.  The Scala host keeps the real asynchronous monad; this Python host
mirror runs the same acquire / release protocol eagerly on the calling thread, which is all the tensor plugins rely on
(`flatMap` keeps upstream resources alive until the downstream value is released; `shared` is reference counted and
releases on the last release -- the hook CumulativeTensorLayers uses to flush accumulated deltas, CTL:39-52).
"""
from __future__ import annotations

from typing import Callable, Sequence, Tuple


class Do:
    __slots__ = ("_acquire",)

    def __init__(self, acquire: Callable[[], Tuple[object, Callable[[], None]]]):
        self._acquire = acquire

    def acquire(self):
        """-> (value, release)"""
        return self._acquire()

    @staticmethod
    def now(value) -> "Do":
        return Do(lambda: (value, _noop))

    @staticmethod
    def delay(thunk) -> "Do":
        return Do(lambda: (thunk(), _noop))

    def map(self, f) -> "Do":
        def acquire():
            value, release = self._acquire()
            return f(value), release
        return Do(acquire)

    def flat_map(self, f) -> "Do":
        def acquire():
            value, release_up = self._acquire()
            try:
                value2, release_down = f(value).acquire()
            except BaseException:
                release_up()
                raise

            def release():
                release_down()
                release_up()
            return value2, release
        return Do(acquire)

    def shared(self) -> "Do":
        count, cell = [0], [None, None]

        def acquire():
            if count[0] == 0:
                cell[0], cell[1] = self._acquire()
            count[0] += 1
            released = [False]

            def release():
                if released[0]:
                    return
                released[0] = True
                count[0] -= 1
                if count[0] == 0:
                    rel, cell[0], cell[1] = cell[1], None, None
                    rel()
            return cell[0], release
        return Do(acquire)

    def run(self):
        """`Do.garbageCollected(...).run`: acquire, keep the value, release."""
        value, release = self._acquire()
        release()
        return value


def _noop():
    return None


def sequence(dos: Sequence[Do]) -> Do:
    """`traverse` / parallelApply (TL:36-38) -- evaluated left to right here."""
    dos = list(dos)

    def acquire():
        values, releases = [], []
        try:
            for d in dos:
                v, r = d.acquire()
                values.append(v)
                releases.append(r)
        except BaseException:
            for r in reversed(releases):
                r()
            raise

        def release():
            for r in reversed(releases):
                r()
        return values, release
    return Do(acquire)
