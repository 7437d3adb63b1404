"""Small helpers layered on actors and tasks: ``ActorPool`` and ``Queue`` (``ray.util.actor_pool`` /
``ray.util.queue``), which data-loading and serving code written next to RayDP leans on."""
from __future__ import annotations

import collections
import time
from typing import Any, Callable, Iterable, Iterator, List, Optional


class ActorPool:
    """Round-robin work over a fixed set of actors: ``pool.map(lambda actor, v: actor.f.remote(v), values)``."""

    def __init__(self, actors: Iterable[Any]):
        self._idle: List[Any] = list(actors)
        self._future_to_actor: dict = {}
        self._index_to_future: dict = {}
        self._next_task, self._next_return = 0, 0
        self._pending: collections.deque = collections.deque()

    def submit(self, fn: Callable[[Any, Any], Any], value: Any) -> None:
        if self._idle:
            actor = self._idle.pop()
            ref = fn(actor, value)
            self._future_to_actor[ref.hex()] = (self._next_task, actor, ref)
            self._index_to_future[self._next_task] = ref
            self._next_task += 1
        else:
            self._pending.append((fn, value))

    def has_next(self) -> bool:
        return bool(self._future_to_actor) or bool(self._pending)

    def has_free(self) -> bool:
        return bool(self._idle) and not self._pending

    def _return_actor(self, actor: Any) -> None:
        self._idle.append(actor)
        if self._pending:
            self.submit(*self._pending.popleft())

    def get_next(self, timeout: Optional[float] = None, ignore_if_timedout: bool = False) -> Any:
        from raydp_b200 import runtime
        if not self.has_next():
            raise StopIteration("No more results to get")
        while self._next_return not in self._index_to_future:       # queued behind busy actors
            self._wait_any(timeout)
        ref = self._index_to_future.pop(self._next_return)
        ready, _ = runtime.wait([ref], num_returns=1, timeout=timeout)
        if not ready:
            self._index_to_future[self._next_return] = ref
            raise TimeoutError("Timed out waiting for result")
        self._next_return += 1
        _, actor, _ = self._future_to_actor.pop(ref.hex())
        self._return_actor(actor)
        return runtime.get(ref)

    def _wait_any(self, timeout: Optional[float]) -> None:
        from raydp_b200 import runtime
        refs = [entry[2] for entry in self._future_to_actor.values()]
        ready, _ = runtime.wait(refs, num_returns=1, timeout=timeout)
        if not ready:
            raise TimeoutError("Timed out waiting for result")

    def get_next_unordered(self, timeout: Optional[float] = None) -> Any:
        from raydp_b200 import runtime
        if not self.has_next():
            raise StopIteration("No more results to get")
        refs = [entry[2] for entry in self._future_to_actor.values()]
        ready, _ = runtime.wait(refs, num_returns=1, timeout=timeout)
        if not ready:
            raise TimeoutError("Timed out waiting for result")
        ref = ready[0]
        index, actor, _ = self._future_to_actor.pop(ref.hex())
        self._index_to_future.pop(index, None)
        if index == self._next_return:
            self._next_return += 1
            while self._next_return < self._next_task and self._next_return not in self._index_to_future:
                self._next_return += 1
        self._return_actor(actor)
        return runtime.get(ref)

    def map(self, fn: Callable[[Any, Any], Any], values: Iterable[Any]) -> Iterator[Any]:  # noqa: A003
        for v in values:
            self.submit(fn, v)
        while self.has_next():
            yield self.get_next()

    def map_unordered(self, fn: Callable[[Any, Any], Any], values: Iterable[Any]) -> Iterator[Any]:
        for v in values:
            self.submit(fn, v)
        while self.has_next():
            yield self.get_next_unordered()

    def push(self, actor: Any) -> None:
        self._return_actor(actor)

    def pop_idle(self) -> Optional[Any]:
        return self._idle.pop() if self.has_free() else None


class Empty(Exception):
    pass


class Full(Exception):
    pass


class _QueueActor:
    def __init__(self, maxsize: int):
        self.maxsize = maxsize
        self.items: collections.deque = collections.deque()

    def qsize(self) -> int:
        return len(self.items)

    def put_nowait(self, item) -> bool:
        if self.maxsize > 0 and len(self.items) >= self.maxsize:
            return False
        self.items.append(item)
        return True

    def get_nowait(self):
        if not self.items:
            return False, None
        return True, self.items.popleft()

    def get_batch(self, n: int):
        if len(self.items) < n:
            return False, None
        return True, [self.items.popleft() for _ in range(n)]


class Queue:
    """A first-in-first-out queue shared between tasks and actors (the state lives in one actor)."""

    def __init__(self, maxsize: int = 0, actor_options: Optional[dict] = None):
        from raydp_b200 import runtime
        self.maxsize = maxsize
        options = dict(actor_options or {})
        options.setdefault("num_cpus", 0)
        self.actor = runtime.remote(_QueueActor).options(**options).remote(maxsize)

    def __len__(self) -> int:
        return self.size()

    def size(self) -> int:
        from raydp_b200 import runtime
        return runtime.get(self.actor.qsize.remote())

    qsize = size

    def empty(self) -> bool:
        return self.size() == 0

    def full(self) -> bool:
        return self.maxsize > 0 and self.size() >= self.maxsize

    def put(self, item: Any, block: bool = True, timeout: Optional[float] = None) -> None:
        from raydp_b200 import runtime
        deadline = None if timeout is None else time.monotonic() + timeout
        while True:
            if runtime.get(self.actor.put_nowait.remote(item)):
                return
            if not block or (deadline is not None and time.monotonic() >= deadline):
                raise Full
            time.sleep(0.005)

    def put_nowait(self, item: Any) -> None:
        self.put(item, block=False)

    def get(self, block: bool = True, timeout: Optional[float] = None) -> Any:
        from raydp_b200 import runtime
        deadline = None if timeout is None else time.monotonic() + timeout
        while True:
            ok, item = runtime.get(self.actor.get_nowait.remote())
            if ok:
                return item
            if not block or (deadline is not None and time.monotonic() >= deadline):
                raise Empty
            time.sleep(0.005)

    def get_nowait(self) -> Any:
        return self.get(block=False)

    def get_nowait_batch(self, num_items: int) -> List[Any]:
        from raydp_b200 import runtime
        ok, items = runtime.get(self.actor.get_batch.remote(num_items))
        if not ok:
            raise Empty(f"Cannot get {num_items} items from queue of size {self.size()}.")
        return items

    def put_nowait_batch(self, items: Iterable[Any]) -> None:
        for item in items:
            self.put_nowait(item)

    def shutdown(self, force: bool = False, grace_period_s: int = 5) -> None:
        from raydp_b200 import runtime
        if self.actor is not None:
            runtime.kill(self.actor)
        self.actor = None
