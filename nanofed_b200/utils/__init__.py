"""Layer L0 utilities (reference ``nanofed/utils/__init__.py:1-4``) plus device timers."""

from .dates import get_current_time
from .logger import LogConfig, Logger, LogLevel, log_exec
from .timing import PhaseTimer

__all__ = ["Logger", "log_exec", "LogLevel", "LogConfig", "get_current_time", "PhaseTimer"]
