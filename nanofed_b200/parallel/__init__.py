"""Layer B/C: cohort scheduling and in-box peer transport."""

from __future__ import annotations

__all__ = ["FedSimulation", "CohortConfig", "RoundOutcome"]


def __getattr__(name: str):
    if name in __all__:
        from . import cohort

        return getattr(cohort, name)
    raise AttributeError(name)
