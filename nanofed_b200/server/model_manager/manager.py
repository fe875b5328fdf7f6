"""Versioned on-disk store for the global model.

Parity (format is a compatibility requirement): reference
``server/model_manager/manager.py:13-210`` —

* ids ``model_v_%Y%m%d_%H%M%S_{counter:03d}`` (UTC, per-process counter);
* ``<models_dir>/<id>.pt`` = ``torch.save(model.state_dict())`` — a plain
  ``dict[str, Tensor]`` on CPU, loadable with ``weights_only=True`` by the
  reference's own ``ModelManager.load_model``;
* ``<configs_dir>/<id>.json`` = ``{"version_id", "timestamp" (ISO-8601 UTC),
  "config"}`` with ``indent=2``;
* ``load_model(None)`` picks the lexicographically last config file;
* ``set_dirs`` seeds a first version with ``{"name": "default", "version": "1.0"}``
  when the store is empty.

Differences: ``metrics`` passed to ``save_model`` are persisted under an extra
top-level ``"metrics"`` key (the reference silently drops them, ``manager.py:101``;
its loader ignores unknown keys); tensors are moved to CPU before saving so a
device-resident master copy still yields a portable checkpoint;
``save_flat(layout, flat)`` checkpoints straight from the arena's flat master
vector (one D2H copy of N floats, no per-tensor traffic).
"""

from __future__ import annotations

import json
from dataclasses import asdict
from datetime import datetime
from pathlib import Path
from typing import Any

import torch

from nanofed_b200.core import ModelManagerError, ModelProtocol, ModelVersion
from nanofed_b200.utils import Logger, get_current_time, log_exec


def make_json_serializable(data: Any) -> Any:
    """Recursively coerce ``data`` to JSON types; unknown objects become ``str(obj)``."""
    if isinstance(data, dict):
        return {str(k) if not isinstance(k, (str, int, float, bool, type(None))) else k:
                make_json_serializable(v) for k, v in data.items()}
    if isinstance(data, (list, tuple)):
        return [make_json_serializable(v) for v in data]
    if hasattr(data, "__dataclass_fields__") and not isinstance(data, type):
        return make_json_serializable(asdict(data))
    if isinstance(data, (int, float, str, bool, type(None))):
        return data
    return str(data)


class ModelManager:
    """Saves, lists and restores versions of one model object."""

    def __init__(self, model: ModelProtocol) -> None:
        self._model = model
        self._logger = Logger()
        self._current_version: ModelVersion | None = None
        self._version_counter = 0
        self._models_dir: Path | None = None
        self._configs_dir: Path | None = None

    # -- wiring -------------------------------------------------------------
    def set_dirs(self, models_dir: Path, configs_dir: Path) -> None:
        self._models_dir = Path(models_dir)
        self._configs_dir = Path(configs_dir)
        if not self.list_versions():
            self._logger.info("No model versions found. Saving initial model.")
            self.save_model(config={"name": "default", "version": "1.0"})

    @property
    def current_version(self) -> ModelVersion | None:
        return self._current_version

    @property
    def model(self) -> ModelProtocol:
        return self._model

    def _dirs(self) -> tuple[Path, Path]:
        if not self._models_dir or not self._configs_dir:
            raise ModelManagerError("Directories not set. Call set_dirs first.")
        return self._models_dir, self._configs_dir

    def _generate_version_id(self) -> str:
        self._version_counter += 1
        return f"model_v_{get_current_time():%Y%m%d_%H%M%S}_{self._version_counter:03d}"

    # -- save -----------------------------------------------------------------
    def _write(
        self,
        state: dict[str, torch.Tensor],
        config: dict[str, Any],
        metrics: dict[str, float] | None,
    ) -> ModelVersion:
        models_dir, configs_dir = self._dirs()
        with self._logger.context("model_manager", "save"):
            version_id = self._generate_version_id()
            weights_path = models_dir / f"{version_id}.pt"
            torch.save(state, weights_path)

            record: dict[str, Any] = {
                "version_id": version_id,
                "timestamp": get_current_time().isoformat(),
                "config": make_json_serializable(config),
            }
            if metrics is not None:
                record["metrics"] = make_json_serializable(metrics)
            try:
                with open(configs_dir / f"{version_id}.json", "w") as fh:
                    json.dump(record, fh, indent=2)
            except TypeError as exc:
                raise ModelManagerError(f"Failed to serialize config data: {exc}") from exc

            version = ModelVersion(
                version_id=version_id,
                timestamp=get_current_time(),
                config=config,
                path=weights_path,
            )
            self._current_version = version
            self._logger.info(f"Saved model version: {version_id}")
            return version

    @log_exec
    def save_model(
        self, config: dict[str, Any], metrics: dict[str, float] | None = None
    ) -> ModelVersion:
        """Persist ``model.state_dict()`` plus its JSON sidecar."""
        self._dirs()
        state = self._model.state_dict()
        if any(isinstance(t, torch.Tensor) and t.device.type != "cpu" for t in state.values()):
            state = {k: (t.detach().cpu() if isinstance(t, torch.Tensor) else t)
                     for k, t in state.items()}
        return self._write(state, config, metrics)

    def save_flat(
        self,
        layout: "Any",
        flat: torch.Tensor,
        config: dict[str, Any],
        metrics: dict[str, float] | None = None,
        sync_model: bool = True,
    ) -> ModelVersion:
        """Checkpoint from a flat master vector (see ``runtime.flat.FlatLayout``)."""
        self._dirs()
        host = flat.detach().to("cpu", torch.float32)
        state = layout.unflatten(host)
        if sync_model:
            self._model.load_state_dict(state)
        return self._write(state, config, metrics)

    # -- load -----------------------------------------------------------------
    @log_exec
    def load_model(self, version_id: str | None = None) -> ModelVersion:
        models_dir, configs_dir = self._dirs()
        with self._logger.context("model_manager", "load"):
            if version_id is None:
                candidates = sorted(configs_dir.glob("*.json"))
                if not candidates:
                    raise ModelManagerError("No model versions found")
                config_path = candidates[-1]
            else:
                config_path = configs_dir / f"{version_id}.json"
                if not config_path.exists():
                    raise ModelManagerError(f"Version {version_id} not found")

            with open(config_path) as fh:
                record = json.load(fh)

            weights_path = models_dir / f"{record['version_id']}.pt"
            if not weights_path.exists():
                raise ModelManagerError(f"Model file not found for version {version_id}")
            try:
                self._model.load_state_dict(torch.load(weights_path, weights_only=True))
            except Exception as exc:
                raise ModelManagerError(f"Failed to load model: {exc}") from exc

            version = ModelVersion(
                version_id=record["version_id"],
                timestamp=datetime.fromisoformat(record["timestamp"]),
                config=record["config"],
                path=weights_path,
            )
            self._current_version = version
            self._logger.info(f"Loaded model version: {version.version_id}")
            return version

    def list_versions(self) -> list[ModelVersion]:
        models_dir, configs_dir = self._dirs()
        found: list[ModelVersion] = []
        for config_path in sorted(configs_dir.glob("*.json")):
            with open(config_path) as fh:
                record = json.load(fh)
            found.append(
                ModelVersion(
                    version_id=record["version_id"],
                    timestamp=datetime.fromisoformat(record["timestamp"]),
                    config=record["config"],
                    path=models_dir / f"{record['version_id']}.pt",
                )
            )
        return found
