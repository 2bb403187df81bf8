"""B200 data-parallel runtime: symmetric heap over NVLink peer memory, fused collective kernels, sync and async
engines, device work queue.  Importing this package does not load the native library; the engines do."""
from .engine import DataParallelEngine

__all__ = ["DataParallelEngine", "AsyncEngine", "AsyncShardState", "DeviceWorkQueue", "SymmetricHeap"]


def __getattr__(name):
    if name in ("AsyncEngine", "AsyncShardState"):
        from . import async_engine
        return getattr(async_engine, name)
    if name == "DeviceWorkQueue":
        from .workqueue import DeviceWorkQueue
        return DeviceWorkQueue
    if name == "SymmetricHeap":
        from .symm import SymmetricHeap
        return SymmetricHeap
    raise AttributeError(name)
