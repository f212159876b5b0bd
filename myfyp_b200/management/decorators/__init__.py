"""The reference builds its logger by stacking decorators (``management/logger/decorators/``: ``LoggerDecorator``,
``AsyncLogger``, ``FileLogger``, ``WebP2PFLogger``, ``RayP2PFLogger``).  Here ONE logger class owns the sinks
(``enable_async`` / ``enable_file_logging`` / ``connect_web``); these wrappers keep the decorator spelling for scripts that
compose a logger that way.  A wrapper switches the sink on and forwards everything else to the wrapped logger."""

from __future__ import annotations

from typing import Any, Optional


class LoggerDecorator:
    """Delegation base (reference: decorators/logger_decorator.py:30-243)."""

    def __init__(self, p2pflogger: Any) -> None:
        self._p2pflogger = p2pflogger

    def __getattr__(self, name: str) -> Any:  # everything not overridden goes to the wrapped logger
        return getattr(self._p2pflogger, name)


class AsyncLogger(LoggerDecorator):
    """Records are handed to a background thread through a queue (reference: async_logger.py:29-58)."""

    def __init__(self, p2pflogger: Any) -> None:
        super().__init__(p2pflogger)
        p2pflogger.enable_async()


class FileLogger(LoggerDecorator):
    """Rotating file sink under ``Settings.LOG_DIR`` (reference: file_logger.py:30-53)."""

    def __init__(self, p2pflogger: Any, log_dir: Optional[str] = None) -> None:
        super().__init__(p2pflogger)
        p2pflogger.enable_file_logging(log_dir)


class WebP2PFLogger(LoggerDecorator):
    """Pushes logs / metrics / system metrics to the web platform (reference: web_logger.py:36-196)."""

    def __init__(self, p2pflogger: Any, url: Optional[str] = None, key: Optional[str] = None) -> None:
        super().__init__(p2pflogger)
        if url is not None and key is not None:
            p2pflogger.connect_web(url, key)

    def connect_web(self, url: str, key: str) -> None:
        self._p2pflogger.connect_web(url, key)


def _merge_logs(dst: dict, src: dict) -> dict:
    """Union of two metric-store dumps (nested dicts whose leaves are lists of (step | round, value) pairs)."""
    for k, v in src.items():
        if isinstance(v, dict):
            _merge_logs(dst.setdefault(k, {}), v)
        else:
            have = dst.setdefault(k, [])
            have.extend(x for x in v if x not in have)
    return dst


class RayP2PFLogger(LoggerDecorator):
    """One logger view across processes.  The reference makes its logger a detached Ray actor so that learners running in other
    processes log into the same stores (decorators/ray_logger.py:32-250).  Here the processes are the ``torchrun`` ranks (one per
    GPU), each with its own in-process logger; this wrapper gives any rank the federation-wide view by merging the ranks' stores
    over ``torch.distributed`` (an object all-gather on the host - observability plumbing, not a hot path):

    ``get_local_logs()`` / ``get_global_logs()`` return the union over all ranks when a process group exists (every rank must
    call, like any collective), and the plain local stores otherwise; ``all_ranks=False`` never communicates."""

    def _gather(self, mine: dict) -> dict:
        try:
            import torch.distributed as dist
        except ImportError:  # pragma: no cover
            return mine
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return mine
        parts = [None] * dist.get_world_size()
        dist.all_gather_object(parts, mine)
        out: dict = {}
        for p in parts:
            _merge_logs(out, p or {})
        return out

    def get_local_logs(self, all_ranks: bool = True):
        mine = self._p2pflogger.get_local_logs()
        return self._gather(mine) if all_ranks else mine

    def get_global_logs(self, all_ranks: bool = True):
        mine = self._p2pflogger.get_global_logs()
        return self._gather(mine) if all_ranks else mine

    def get_nodes(self, all_ranks: bool = True):
        mine = {k: {"rank": None} for k in getattr(self._p2pflogger, "_nodes", {})}
        return self._gather(mine) if all_ranks else mine


DistributedP2PFLogger = RayP2PFLogger  # what it is, next to the reference's name
