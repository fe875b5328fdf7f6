"""Process-wide logger with per-task component labels, plus ``log_exec``.

Parity: reference ``nanofed/utils/logger.py`` — ``Logger`` singleton wrapping
``logging.getLogger("nanofed")`` (``:59-77``), ``context(component, sub)``
(``:79-88``), ``configure(LogConfig)`` (``:90-115``), colour formatter
(``:138-167``) and the ``log_exec`` timing decorator (``:189-226``).

Difference by design: the reference keeps ONE mutable ``current_context`` on the
singleton, so concurrent coroutines mislabel each other's lines
(``tutorial.ipynb:684``).  Here the label lives in a ``contextvars.ContextVar``,
which asyncio copies per task, so each coroutine sees its own label.
``current_context`` is still readable/assignable for code that pokes at it.
"""

from __future__ import annotations

import asyncio
import contextvars
import logging
import sys
import time
from contextlib import AbstractContextManager, contextmanager
from dataclasses import dataclass
from enum import Enum, auto
from functools import wraps
from pathlib import Path
from typing import Any, Callable, ClassVar, Iterator, Literal, ParamSpec, TypeVar

from .dates import get_current_time

P = ParamSpec("P")
R = TypeVar("R")

_LOGGER_NAME = "nanofed"
_DEFAULT_LABEL = "nanofed"


class LogLevel(Enum):
    DEBUG = auto()
    INFO = auto()
    WARNING = auto()
    ERROR = auto()


_STD_LEVEL = {
    LogLevel.DEBUG: logging.DEBUG,
    LogLevel.INFO: logging.INFO,
    LogLevel.WARNING: logging.WARNING,
    LogLevel.ERROR: logging.ERROR,
}


@dataclass(slots=True, frozen=True)
class LogConfig:
    """Logger settings (field-for-field the reference's ``LogConfig``)."""

    level: LogLevel
    color: bool
    format: str
    output: Literal["console", "file", "both"]
    log_dir: Path | None = None


@dataclass(slots=True)
class LogContext:
    """``component[.subcomponent]`` label printed in each record."""

    _component: str
    _subcomponent: str | None = None

    def __str__(self) -> str:
        return (
            f"{self._component}.{self._subcomponent}"
            if self._subcomponent
            else self._component
        )


_label_var: contextvars.ContextVar[LogContext | None] = contextvars.ContextVar(
    "nanofed_b200_log_label", default=None
)


class Formatter(logging.Formatter):
    """``time - label - LEVEL - message`` with optional ANSI level colours."""

    COLORS: ClassVar[dict[str, str]] = {
        "DEBUG": "\033[36m",
        "INFO": "\033[32m",
        "WARNING": "\033[33m",
        "ERROR": "\033[31m",
        "RESET": "\033[0m",
    }

    def __init__(self, use_color: bool = True) -> None:
        super().__init__("%(asctime)s - %(context)s - %(levelname)s - %(message)s")
        self.use_color = use_color

    def format(self, record: logging.LogRecord) -> str:
        if not hasattr(record, "context"):
            record.context = _DEFAULT_LABEL
        if self.use_color and record.levelname in self.COLORS:
            # Work on a copy: the same record may go to a plain file handler too.
            record = logging.makeLogRecord(record.__dict__)
            record.levelname = (
                f"{self.COLORS[record.levelname]}{record.levelname}{self.COLORS['RESET']}"
            )
        return super().format(record)


class Logger:
    """Singleton façade over the stdlib logger named ``"nanofed"``."""

    _instance: ClassVar["Logger | None"] = None
    _initialized: bool = False

    def __new__(cls) -> "Logger":
        if cls._instance is None:
            cls._instance = super().__new__(cls)
        return cls._instance

    def __init__(self) -> None:
        if self._initialized:
            return
        self._logger = logging.getLogger(_LOGGER_NAME)
        self._logger.setLevel(logging.DEBUG)
        if not self._logger.handlers:
            handler = logging.StreamHandler(sys.stdout)
            handler.setFormatter(Formatter(use_color=True))
            self._logger.addHandler(handler)
        self._initialized = True

    # -- label handling -------------------------------------------------
    @property
    def current_context(self) -> LogContext | None:
        return _label_var.get()

    @current_context.setter
    def current_context(self, value: LogContext | None) -> None:
        _label_var.set(value)

    @contextmanager
    def context(
        self, component: str, subcomponent: str | None = None
    ) -> Iterator["Logger"]:
        token = _label_var.set(LogContext(component, subcomponent))
        try:
            yield self
        finally:
            try:
                _label_var.reset(token)
            except ValueError:
                # Exited in a different Context than entered (async generator
                # finalisation); nothing to restore in this one.
                _label_var.set(None)

    # -- configuration --------------------------------------------------
    def configure(self, config: LogConfig) -> None:
        """Replace handlers according to ``config``."""
        for handler in list(self._logger.handlers):
            self._logger.removeHandler(handler)
        self._logger.setLevel(_STD_LEVEL[config.level])
        if config.output in ("console", "both"):
            console = logging.StreamHandler(sys.stdout)
            console.setFormatter(Formatter(config.color))
            self._logger.addHandler(console)
        if config.output in ("file", "both") and config.log_dir:
            target = config.log_dir / f"nanofed_{get_current_time():%Y%m%d_%H%M%S}.log"
            target.parent.mkdir(parents=True, exist_ok=True)
            to_file = logging.FileHandler(target)
            to_file.setFormatter(Formatter(use_color=False))
            self._logger.addHandler(to_file)

    # -- emitters ---------------------------------------------------------
    def _log(self, level: int, msg: str) -> None:
        label = _label_var.get()
        self._logger.log(
            level, msg, extra={"context": str(label) if label else _DEFAULT_LABEL}
        )

    def debug(self, msg: str) -> None:
        self._log(logging.DEBUG, msg)

    def info(self, msg: str) -> None:
        self._log(logging.INFO, msg)

    def warning(self, msg: str) -> None:
        self._log(logging.WARNING, msg)

    def error(self, msg: str) -> None:
        self._log(logging.ERROR, msg)


class LoggerContextManager(AbstractContextManager):
    """Class-based twin of ``Logger.context`` (reference ``logger.py:170-186``)."""

    __slots__ = ("_logger", "_context", "_token")

    def __init__(self, logger: Logger, context: LogContext) -> None:
        self._logger = logger
        self._context = context
        self._token: contextvars.Token | None = None

    def __enter__(self) -> Logger:
        self._token = _label_var.set(self._context)
        return self._logger

    def __exit__(self, *exc: Any) -> None:
        if self._token is not None:
            _label_var.reset(self._token)
            self._token = None


def log_exec(func: Callable[P, R]) -> Callable[P, R]:
    """Log ``Starting f`` / ``Completed f in X.XXs`` at DEBUG; errors at ERROR.

    Works for plain functions and coroutines.  Wall-clock (monotonic) host
    time; device-side phases are timed with CUDA events in
    ``nanofed_b200.utils.timing`` instead.
    """
    name = func.__name__

    def _done(t0: float) -> None:
        Logger().debug(f"Completed {name} in {time.perf_counter() - t0:.2f}s")

    if asyncio.iscoroutinefunction(func):

        @wraps(func)
        async def async_wrapper(*args: P.args, **kwargs: P.kwargs) -> R:
            t0 = time.perf_counter()
            Logger().debug(f"Starting {name}")
            try:
                out = await func(*args, **kwargs)  # type: ignore[misc]
            except Exception as exc:
                Logger().error(f"Error in {name}: {exc}")
                raise
            _done(t0)
            return out

        return async_wrapper  # type: ignore[return-value]

    @wraps(func)
    def sync_wrapper(*args: P.args, **kwargs: P.kwargs) -> R:
        t0 = time.perf_counter()
        Logger().debug(f"Starting {name}")
        try:
            out = func(*args, **kwargs)
        except Exception as exc:
            Logger().error(f"Error in {name}: {exc}")
            raise
        _done(t0)
        return out

    return sync_wrapper
