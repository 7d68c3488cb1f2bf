"""A minimal stand-in for dask.distributed.Client (dask is not installed here): scatter / submit / map / gather with eager
execution, futures resolved when passed as arguments.  It records what was submitted so tests can count tasks."""
from typing import Any, Callable, List


class FakeFuture:
    def __init__(self, value: Any):
        self.value = value


def _resolve(x: Any) -> Any:
    if isinstance(x, FakeFuture):
        return x.value
    if isinstance(x, list):
        return [_resolve(v) for v in x]
    if isinstance(x, tuple):
        return tuple(_resolve(v) for v in x)
    if isinstance(x, dict):
        return {k: _resolve(v) for k, v in x.items()}
    return x


class FakeClient:
    def __init__(self):
        self.submitted: List[str] = []
        self.scattered: List[Any] = []

    def scatter(self, obj: Any, broadcast: bool = False) -> FakeFuture:
        self.scattered.append(obj)
        return FakeFuture(obj)

    def submit(self, fn: Callable, *args, **kwargs) -> FakeFuture:
        self.submitted.append(getattr(fn, "__name__", str(fn)))
        return FakeFuture(fn(*[_resolve(a) for a in args], **{k: _resolve(v) for k, v in kwargs.items()}))

    def map(self, fn: Callable, futures) -> List[FakeFuture]:
        return [self.submit(fn, f) for f in futures]

    def gather(self, x: Any) -> Any:
        return _resolve(x)

    def count(self, name: str) -> int:
        return sum(1 for s in self.submitted if s == name)
