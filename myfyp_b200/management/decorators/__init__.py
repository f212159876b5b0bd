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


class RayP2PFLogger(LoggerDecorator):
    """The reference wraps its logger in a Ray actor so that simulated nodes in other processes share it.  Nodes of a process
    share the logger object here (``FitPool`` runs learners in-process), so this is a plain pass-through kept for the name."""
