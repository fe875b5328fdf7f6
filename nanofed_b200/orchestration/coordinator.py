"""Server-side round state machine.

Parity: reference ``orchestration/coordinator.py:26-405`` — ``CoordinatorConfig``
fields; constructor creating ``<base>/metrics``, ``<base>/data``,
``<base>/models/{configs,models}``, calling ``model_manager.set_dirs`` (which seeds
version ``_001``) and ``server.set_coordinator``; ``train_round()`` =
wait → rebuild ``ModelUpdate``s → weights → ``aggregator.aggregate`` →
``model_manager.save_model(config={round_id, client_metrics, client_weights,
start_time, status, num_clients}, metrics=…)`` → ``metrics_round_<id>.json``
(``indent=4``, same schema) ; ``start_training()`` is an async generator yielding one
``RoundMetrics`` per round and finally ``server.stop_training()``;
``training_progress`` / ``_get_global_metrics``.  On-disk layout: SURVEY §5.4.

Fixes:

* ``privacy_spent`` is read with ``.get`` — at the reference's HEAD every HTTP
  round dies with ``KeyError('privacy_spent')`` (``coordinator.py:319``);
* the server's round counter is advanced after each round, so later rounds are
  not all "round 0" on the wire and stale updates are rejected (``server.py:260``
  shows that was the intent);
* updates are taken from the server under its lock (``take_updates``) and the
  wait is event-driven, not a 1 Hz poll over a dict the handlers mutate;
* ``required_clients = max(1, int(min_clients · min_completion_rate))``;
* ``_clients`` is populated, so ``active_clients`` is meaningful;
* optional ``fault_tolerance=FaultTolerantCoordinator(...)`` checkpoints every
  COMPLETED round and restores the global model after a recoverable failure, and
  ``validator=`` screens updates before aggregation (both are unwired libraries
  in the reference).

In-box rounds do not go through this wait loop at all: a ``CohortEngine``
(``parallel.cohort``) passed as ``engine=`` runs the whole round — broadcast,
local epochs, fused aggregate — on the GPUs and this class only does the
bookkeeping and checkpointing around it.
"""

from __future__ import annotations

import asyncio
import json
from dataclasses import dataclass
from datetime import datetime
from pathlib import Path
from typing import TYPE_CHECKING, Any, AsyncGenerator, Callable, Sequence

import torch

from nanofed_b200.core import ModelManagerProtocol, ModelUpdate
from nanofed_b200.server import BaseAggregator
from nanofed_b200.server.fault_tolerance import FaultTolerantCoordinator, RoundState
from nanofed_b200.utils import Logger, get_current_time, log_exec

from .types import ClientInfo, RoundMetrics, RoundStatus, TrainingProgress

if TYPE_CHECKING:
    from nanofed_b200.communication import HTTPServer


@dataclass(slots=True, frozen=True)
class CoordinatorConfig:
    num_rounds: int
    min_clients: int
    min_completion_rate: float
    round_timeout: int
    base_dir: Path


def _as_tensor_state(raw: dict[str, Any]) -> dict[str, torch.Tensor]:
    from nanofed_b200.communication.http.codec import tensor_from_jsonable

    return {k: tensor_from_jsonable(v) for k, v in raw.items()}


def _as_time(value: Any) -> datetime:
    if isinstance(value, datetime):
        return value
    try:
        return datetime.fromisoformat(value)
    except (TypeError, ValueError):
        return get_current_time()


class Coordinator:
    def __init__(
        self,
        model_manager: ModelManagerProtocol,
        aggregator: BaseAggregator,
        server: "HTTPServer",
        config: CoordinatorConfig,
        *,
        engine: Any = None,
        fault_tolerance: FaultTolerantCoordinator | None = None,
        validator: Any = None,
    ) -> None:
        self._model_manager = model_manager
        self._aggregator = aggregator
        self._server = server
        self._config = config
        self._engine = engine
        self._ft = fault_tolerance
        self._validator = validator
        self._logger = Logger()

        self._current_round = 0
        self._clients: dict[str, ClientInfo] = {}
        self._round_metrics: list[RoundMetrics] = []
        self._status = RoundStatus.INITIALIZED
        self._round_lock = asyncio.Lock()

        base = Path(config.base_dir)
        self._metrics_dir = base / "metrics"
        self._data_dir = base / "data"
        self._models_dir = base / "models"
        self._model_configs_dir = self._models_dir / "configs"
        self._model_weights_dir = self._models_dir / "models"
        self._setup_directories()

        self._model_manager.set_dirs(self._model_weights_dir, self._model_configs_dir)
        if self._server is not None:
            self._server.set_coordinator(self)
        if self._engine is not None and hasattr(self._engine, "attach"):
            self._engine.attach(self)

    # -- accessors ---------------------------------------------------------
    @property
    def server(self) -> "HTTPServer":
        return self._server

    @property
    def data_dir(self) -> Path:
        return self._data_dir

    @property
    def model_manager(self) -> ModelManagerProtocol:
        return self._model_manager

    @property
    def aggregator(self) -> BaseAggregator:
        return self._aggregator

    @property
    def config(self) -> CoordinatorConfig:
        return self._config

    @property
    def current_round(self) -> int:
        return self._current_round

    @property
    def round_metrics(self) -> list[RoundMetrics]:
        return list(self._round_metrics)

    def _setup_directories(self) -> None:
        with self._logger.context("coordinator.setup"):
            for directory in (
                self._metrics_dir,
                self._data_dir,
                self._model_configs_dir,
                self._model_weights_dir,
            ):
                try:
                    directory.mkdir(parents=True, exist_ok=True)
                    self._logger.debug(f"Created directory: {directory}")
                except Exception as exc:
                    self._logger.error(f"Failed to create directory {directory}: {exc}")
                    raise

    @property
    def training_progress(self) -> TrainingProgress:
        return {
            "current_round": self._current_round,
            "total_rounds": self._config.num_rounds,
            "active_clients": len(self._clients),
            "global_metrics": self._get_global_metrics(),
            "status": self._status.name,
        }

    def _get_global_metrics(self) -> dict[str, float]:
        series: dict[str, list[float]] = {}
        for rm in self._round_metrics:
            for name, value in rm.agg_metrics.items():
                series.setdefault(name, []).append(value)
        return {name: sum(vals) / len(vals) for name, vals in series.items()}

    # -- waiting ----------------------------------------------------------------
    @property
    def required_clients(self) -> int:
        return max(1, int(self._config.min_clients * self._config.min_completion_rate))

    async def _wait_for_clients(self, timeout: int) -> bool:
        with self._logger.context("coordinator"):
            need = self.required_clients
            waiter = getattr(self._server, "wait_for_updates", None)
            if waiter is not None:
                ok = await waiter(need, float(timeout))
            else:  # duck-typed servers that only expose ``_updates``
                loop = asyncio.get_running_loop()
                end = loop.time() + timeout
                ok = False
                while loop.time() < end:
                    if len(self._server._updates) >= need:
                        ok = True
                        break
                    await asyncio.sleep(0.05)
            got = len(self._server._updates)
            if ok:
                self._logger.info(
                    f"Sufficient clients completed training: {got}/{self._config.min_clients}"
                )
            else:
                self._logger.error(
                    f"Timeout waiting for clients. Got {got}/{self._config.min_clients} "
                    f"(needed {need})"
                )
            return ok

    # -- persistence ----------------------------------------------------------------
    def _save_metrics(self, metrics: RoundMetrics, client_metrics: list[dict]) -> None:
        with self._logger.context("coordinator.metrics", f"round_{metrics.round_id}"):
            target = self._metrics_dir / f"metrics_round_{metrics.round_id}.json"
            record = {
                "round_id": metrics.round_id,
                "start_time": metrics.start_time.isoformat() if metrics.start_time else None,
                "end_time": metrics.end_time.isoformat() if metrics.end_time else None,
                "num_clients": metrics.num_clients,
                "agg_metrics": metrics.agg_metrics,
                "status": metrics.status.name,
                "client_metrics": client_metrics,
            }
            try:
                with target.open("w") as fh:
                    json.dump(record, fh, indent=4)
                self._logger.info(f"Saved metrics for round {metrics.round_id} to {target}")
            except Exception as exc:
                self._logger.error(
                    f"Failed to save metrics for round {metrics.round_id}: {exc}"
                )

    # -- one round -----------------------------------------------------------------
    async def _collect_updates(self) -> Sequence[ModelUpdate]:
        taker = getattr(self._server, "take_updates", None)
        raw = await taker() if taker is not None else dict(self._server._updates)
        updates: list[ModelUpdate] = []
        for item in raw.values():
            updates.append(
                ModelUpdate(
                    client_id=item["client_id"],
                    round_number=item["round_number"],
                    model_state=_as_tensor_state(item["model_state"]),
                    metrics=item["metrics"],
                    timestamp=_as_time(item["timestamp"]),
                    privacy_spent=item.get("privacy_spent"),
                )
            )
        return updates

    def _screen(self, updates: Sequence[ModelUpdate]) -> Sequence[ModelUpdate]:
        if self._validator is None:
            return updates
        from nanofed_b200.server.validation import ValidationConfig, ValidationResult

        shapes = {k: v.shape for k, v in self._model_manager.model.state_dict().items()}
        cfg = getattr(self._validator, "_config", ValidationConfig())
        kept = []
        for u in updates:
            verdicts = (
                self._validator.validate_shape(u, shapes),
                self._validator.validate_range(u, cfg),
                self._validator.validate_statistics(u, updates),
            )
            if all(v == ValidationResult.VALID for v in verdicts):
                kept.append(u)
            else:
                self._logger.warning(
                    f"Rejected update from {u['client_id']}: {[v.name for v in verdicts]}"
                )
        return kept

    @log_exec
    async def train_round(self) -> RoundMetrics:
        with self._logger.context("coordinator", f"round_{self._current_round}"):
            async with self._round_lock:
                try:
                    self._status = RoundStatus.IN_PROGRESS
                    start_time = get_current_time()

                    if self._engine is not None:
                        return await self._engine_round(start_time)

                    if not await self._wait_for_clients(self._config.round_timeout):
                        self._logger.warning(
                            f"Skipping round {self._current_round} due to timeout."
                        )
                        self._status = RoundStatus.FAILED
                        raise TimeoutError(
                            f"Round {self._current_round} timed out waiting for clients"
                        )

                    self._status = RoundStatus.AGGREGATING
                    client_updates = self._screen(await self._collect_updates())
                    if not client_updates:
                        raise RuntimeError("All client updates were rejected by validation")

                    weights = self._aggregator._compute_weights(client_updates)
                    client_weights = {
                        u["client_id"]: w for u, w in zip(client_updates, weights)
                    }
                    client_metrics = [
                        {
                            "client_id": u.get("client_id"),
                            "metrics": u.get("metrics", {}),
                            "weight": client_weights[str(u.get("client_id", ""))],
                        }
                        for u in client_updates
                    ]
                    now = get_current_time()
                    for u in client_updates:
                        self._clients[u["client_id"]] = ClientInfo(
                            client_id=u["client_id"], status="completed",
                            last_update=now, metrics=dict(u.get("metrics", {})),
                        )

                    result = self._aggregator.aggregate(
                        self._model_manager.model, client_updates
                    )
                    return await self._finish_round(
                        start_time, result.metrics, client_metrics, client_weights,
                        len(client_updates),
                    )
                except Exception as exc:
                    self._status = RoundStatus.FAILED
                    self._logger.error(f"Error in round {self._current_round}: {exc}")
                    raise

    async def _engine_round(self, start_time: datetime) -> RoundMetrics:
        """In-box round: the cohort engine trains and aggregates on the GPUs."""
        loop = asyncio.get_running_loop()
        outcome = await loop.run_in_executor(None, getattr(self._engine, "run_round_for_coordinator", getattr(self._engine, "run_round", None)), self._current_round)
        self._status = RoundStatus.AGGREGATING
        for cm in outcome.client_metrics:
            self._clients[cm["client_id"]] = ClientInfo(
                client_id=cm["client_id"], status="completed",
                last_update=get_current_time(), metrics=dict(cm["metrics"]),
            )
        weights = {cm["client_id"]: cm["weight"] for cm in outcome.client_metrics}
        return await self._finish_round(
            start_time, outcome.agg_metrics, outcome.client_metrics, weights,
            outcome.num_clients,
        )

    async def _finish_round(
        self,
        start_time: datetime,
        agg_metrics: dict[str, float],
        client_metrics: list[dict],
        client_weights: dict[str, float],
        num_clients: int,
    ) -> RoundMetrics:
        config = {
            "round_id": self._current_round,
            "client_metrics": client_metrics,
            "client_weights": client_weights,
            "start_time": start_time.isoformat(),
            "status": self._status.name,
            "num_clients": num_clients,
        }
        version = self._model_manager.save_model(config=config, metrics=agg_metrics)

        self._current_round += 1
        self._status = RoundStatus.COMPLETED
        metrics = RoundMetrics(
            round_id=self._current_round - 1,
            start_time=start_time,
            end_time=get_current_time(),
            num_clients=num_clients,
            agg_metrics=agg_metrics,
            status=self._status,
        )
        self._round_metrics.append(metrics)
        self._save_metrics(metrics, client_metrics)

        if self._ft is not None:
            state = {
                k: v.detach().cpu()
                for k, v in self._model_manager.model.state_dict().items()
            }
            self._ft.checkpoint_round(
                metrics.round_id, {}, getattr(version, "version_id", ""), state,
                RoundState.COMPLETED,
            )

        advance = getattr(self._server, "advance_round", None)
        if advance is not None:
            await advance(self._current_round)
        elif self._server is not None:
            self._server._updates.clear()
        return metrics

    # -- whole run -----------------------------------------------------------------
    async def _recover(self, failure: Exception) -> bool:
        if self._ft is None:
            return False
        restored = self._ft.handle_failure(failure, self._current_round)
        if restored is None:
            return False
        meta, state = restored
        self._model_manager.model.load_state_dict(state)
        # the in-box engine keeps the global model in its arena and syncs arena -> model after
        # every round: restore it there too, or the next round overwrites the recovered weights
        load = getattr(self._engine, "load_global_state", None)
        if load is not None:
            load(self._model_manager.model.state_dict())
        invalidate = getattr(self._server, "invalidate_model_cache", None)
        if invalidate is not None:
            invalidate()
        self._logger.info(f"Restored global model from round {meta.round_id}")
        return True

    async def start_training(
        self,
        progress_callback: Callable[[TrainingProgress], None] | None = None,
        max_recoveries: int = 1,
    ) -> AsyncGenerator[RoundMetrics, None]:
        with self._logger.context("coordinator"):
            try:
                recoveries = 0
                while self._current_round < self._config.num_rounds:
                    try:
                        metrics = await self.train_round()
                    except Exception as exc:
                        if recoveries < max_recoveries and await self._recover(exc):
                            recoveries += 1
                            continue
                        raise
                    if progress_callback:
                        progress_callback(self.training_progress)
                    yield metrics
                if self._server is not None:
                    await self._server.stop_training()
            except Exception as exc:
                self._logger.error(f"Training failed: {exc}")
                raise
            finally:
                self._logger.info("Training completed")
