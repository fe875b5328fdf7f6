"""``coordinate``: drain a coordinator's round generator (reference ``orchestration/utils.py:5-25``)."""

from __future__ import annotations

from nanofed_b200.utils import Logger

from .coordinator import Coordinator


async def coordinate(coordinator: Coordinator) -> None:
    logger = Logger()
    with logger.context("coordinator.run"):
        try:
            async for _ in coordinator.start_training():
                pass
        except Exception as exc:
            logger.error(f"Error while running coordinator: {exc}")
            raise
        finally:
            logger.info("Coordinator run completed.")
