"""Round checkpoints and recovery policy.

Parity: reference ``server/fault_tolerance.py:14-212`` — ``RoundState``,
``CheckpointMetadata`` (+ ``to_dict`` / ``from_dict`` turning JSON lists back into
tensors), ``FileStateStore`` writing
``<base_dir>/checkpoints/round_<id>/{metadata.json,state.pt}`` and loading with
``weights_only=True``, ``SimpleRecoveryStrategy`` (recoverable =
``TimeoutError | ConnectionError | RuntimeError``; recovery point = highest
COMPLETED round) and ``FaultTolerantCoordinator`` (``test_fault_tolerance.py:59-211``).

Differences: tensors inside ``client_updates`` are converted to nested lists on
save (the reference hands raw tensors to ``json.dump`` and fails unless the
caller pre-converted them, ``fault_tolerance.py:34-42``); ``list_checkpoints``
orders by numeric round id (the reference sorts paths lexicographically, so
``round_10`` precedes ``round_2``); and — unlike the reference, where nothing
calls this module — ``Coordinator(fault_tolerance=…)`` wires it into the round
loop (checkpoint after every COMPLETED round, restore on a recoverable failure).
"""

from __future__ import annotations

import json
import re
from dataclasses import dataclass
from datetime import datetime
from enum import Enum, auto
from pathlib import Path
from typing import Any, Protocol

import torch

from nanofed_b200.core import ModelUpdate
from nanofed_b200.utils import Logger, get_current_time


class RoundState(Enum):
    INITIALIZED = auto()
    IN_PROGRESS = auto()
    FAILED = auto()
    COMPLETED = auto()


def _jsonable(value: Any) -> Any:
    if isinstance(value, torch.Tensor):
        return value.detach().cpu().tolist()
    if isinstance(value, dict):
        return {k: _jsonable(v) for k, v in value.items()}
    if isinstance(value, (list, tuple)):
        return [_jsonable(v) for v in value]
    if isinstance(value, datetime):
        return value.isoformat()
    if isinstance(value, Enum):
        return value.name
    if hasattr(value, "__dataclass_fields__"):
        return {k: _jsonable(getattr(value, k)) for k in value.__dataclass_fields__}
    return value


@dataclass(slots=True, frozen=True)
class CheckpointMetadata:
    round_id: int
    timestamp: datetime
    num_clients: int
    client_updates: dict[str, ModelUpdate]
    global_model_version: str
    state: RoundState

    def to_dict(self) -> dict[str, Any]:
        return {
            "round_id": self.round_id,
            "timestamp": self.timestamp.isoformat(),
            "num_clients": self.num_clients,
            "client_updates": _jsonable(self.client_updates),
            "global_model_version": self.global_model_version,
            "state": self.state.name,
        }

    @staticmethod
    def from_dict(data: dict[str, Any]) -> "CheckpointMetadata":
        updates = data["client_updates"]
        for update in updates.values():
            tensors = update.get("model_state") or {}
            for name, value in tensors.items():
                if not isinstance(value, torch.Tensor):
                    tensors[name] = torch.tensor(value)
        return CheckpointMetadata(
            round_id=data["round_id"],
            timestamp=datetime.fromisoformat(data["timestamp"]),
            num_clients=data["num_clients"],
            client_updates=updates,
            global_model_version=data["global_model_version"],
            state=RoundState[data["state"]],
        )


class StateStore(Protocol):
    def save_checkpoint(self, metadata: CheckpointMetadata, state: dict[str, Any]) -> None: ...

    def load_checkpoint(
        self, round_id: int
    ) -> tuple[CheckpointMetadata, dict[str, Any]] | None: ...

    def list_checkpoints(self) -> list[CheckpointMetadata]: ...


class RecoveryStrategy(Protocol):
    def should_recover(self, failure: Exception) -> bool: ...

    def get_recovery_point(
        self, checkpoints: list[CheckpointMetadata]
    ) -> CheckpointMetadata | None: ...


_ROUND_DIR = re.compile(r"round_(-?\d+)$")


class FileStateStore(StateStore):
    """One directory per round under ``<base_dir>/checkpoints``."""

    def __init__(self, base_dir: Path) -> None:
        self._base_dir = Path(base_dir) / "checkpoints"
        self._base_dir.mkdir(parents=True, exist_ok=True)
        self._logger = Logger()

    def _dir(self, round_id: int) -> Path:
        return self._base_dir / f"round_{round_id}"

    def save_checkpoint(self, metadata: CheckpointMetadata, state: dict[str, Any]) -> None:
        target = self._dir(metadata.round_id)
        target.mkdir(exist_ok=True)
        with open(target / "metadata.json", "w") as fh:
            json.dump(metadata.to_dict(), fh)
        torch.save(state, target / "state.pt")
        self._logger.info(f"Saved checkpoint for round {metadata.round_id}")

    def _read_metadata(self, path: Path) -> CheckpointMetadata:
        with open(path) as fh:
            return CheckpointMetadata.from_dict(json.load(fh))

    def load_checkpoint(
        self, round_id: int
    ) -> tuple[CheckpointMetadata, dict[str, Any]] | None:
        source = self._dir(round_id)
        if not source.exists():
            return None
        metadata = self._read_metadata(source / "metadata.json")
        state = torch.load(source / "state.pt", weights_only=True)
        self._logger.info(f"Loaded checkpoint for round {round_id}")
        return metadata, state

    def list_checkpoints(self) -> list[CheckpointMetadata]:
        found: list[tuple[int, Path]] = []
        for path in self._base_dir.glob("round_*"):
            m = _ROUND_DIR.search(path.name)
            if m and (path / "metadata.json").exists():
                found.append((int(m.group(1)), path))
        return [self._read_metadata(p / "metadata.json") for _, p in sorted(found)]


class SimpleRecoveryStrategy(RecoveryStrategy):
    """Recover from transient failures using the newest COMPLETED checkpoint."""

    RECOVERABLE = (TimeoutError, ConnectionError, RuntimeError)

    def should_recover(self, failure: Exception) -> bool:
        return isinstance(failure, self.RECOVERABLE)

    def get_recovery_point(
        self, checkpoints: list[CheckpointMetadata]
    ) -> CheckpointMetadata | None:
        done = [c for c in checkpoints if c.state == RoundState.COMPLETED]
        return max(done, key=lambda c: c.round_id) if done else None


class FaultTolerantCoordinator:
    """Checkpoint/restore façade used by ``Coordinator`` and usable on its own."""

    def __init__(
        self,
        base_dir: Path,
        state_store: StateStore | None = None,
        recovery_strategy: RecoveryStrategy | None = None,
    ) -> None:
        self._state_store = state_store or FileStateStore(base_dir)
        self._recovery = recovery_strategy or SimpleRecoveryStrategy()
        self._logger = Logger()

    def checkpoint_round(
        self,
        round_id: int,
        client_updates: dict[str, ModelUpdate],
        model_version: str,
        state: dict[str, Any],
        round_state: RoundState,
    ) -> None:
        self._state_store.save_checkpoint(
            CheckpointMetadata(
                round_id=round_id,
                timestamp=get_current_time(),
                num_clients=len(client_updates),
                client_updates=client_updates,
                global_model_version=model_version,
                state=round_state,
            ),
            state,
        )

    def restore_round(
        self, round_id: int
    ) -> tuple[CheckpointMetadata, dict[str, Any]] | None:
        return self._state_store.load_checkpoint(round_id)

    def handle_failure(
        self, failure: Exception, current_round: int
    ) -> tuple[CheckpointMetadata, dict[str, Any]] | None:
        if not self._recovery.should_recover(failure):
            self._logger.error(f"Unrecoverable failure in round {current_round}: {failure}")
            return None
        point = self._recovery.get_recovery_point(self._state_store.list_checkpoints())
        if point is None:
            self._logger.error("No valid recovery point found")
            return None
        self._logger.info(f"Recovering from round {point.round_id}")
        return self.restore_round(point.round_id)
