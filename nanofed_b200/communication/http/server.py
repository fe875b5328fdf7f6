"""aiohttp server: the cross-host transport endpoint.

Parity: reference ``communication/http/server.py:29-340`` and SURVEY Appendix B —
routes ``GET /model``, ``POST /update``, ``GET /status``, ``GET /test``; the same
JSON envelopes, status codes and messages ("Server not initialized with
coordinator" → 500, ``status:"terminated"`` + ``round_number:-1`` once training is
done, "Missing keys: …" / "Invalid round number" → 400, ``update_id =
update_<cid>_<round>``); attributes ``_app``, ``_updates``, ``_current_round``,
``_lock``, ``_is_training_done`` that the reference's tests and coordinator touch.

Fixes: an optional ``privacy_spent`` in the request is stored (the
reference drops it and then crashes reading it, ``coordinator.py:319``);
``message`` is read under its real name (``"mesage"`` still accepted);
``advance_round()`` lets the coordinator move ``_current_round`` forward so stale
updates are rejected; updates are handed over through ``take_updates()`` under
the lock instead of the coordinator reaching into ``_updates`` unlocked.

Speed-ups that keep the wire format: the encoded ``/model`` body is cached per
model version (the reference re-runs ``tolist()`` + ``json.dumps`` of ~27 MB for
every client), encoding/decoding of large bodies runs in a worker thread so the
event loop keeps serving, and both endpoints also speak the binary frame of
``codec.py`` when the client asks (``Accept`` / ``Content-Type:
application/octet-stream``).  In-box clients do not use this server at all: they
exchange models through the peer-memory arena (``parallel.peer``).
"""

from __future__ import annotations

import asyncio
import copy
import json
from dataclasses import dataclass
from typing import TYPE_CHECKING, Any

import torch
from aiohttp import web

from nanofed_b200.core import ServerProtocol
from nanofed_b200.utils import Logger, get_current_time

from . import codec
from .types import GlobalModelResponse, ModelUpdateResponse, ServerModelUpdateRequest

if TYPE_CHECKING:
    from nanofed_b200.orchestration import Coordinator

_REQUIRED_UPDATE_KEYS = ("client_id", "round_number", "model_state", "metrics", "timestamp")


@dataclass(slots=True, frozen=True)
class ServerEndpoints:
    get_model: str = "/model"
    submit_update: str = "/update"
    get_status: str = "/status"


def _now() -> str:
    return get_current_time().isoformat()


def _error(message: str, status: int) -> web.Response:
    return web.json_response(
        {"status": "error", "message": message, "timestamp": _now()}, status=status
    )


class HTTPServer(ServerProtocol):
    def __init__(
        self,
        host: str,
        port: int,
        endpoints: ServerEndpoints | None = None,
        max_request_size: int = 100 * 1024 * 1024,
    ) -> None:
        self._host = host
        self._port = port
        self._endpoints = endpoints or ServerEndpoints()
        self._logger = Logger()
        self._app = web.Application(client_max_size=max_request_size)
        self._setup_routes()
        self._runner: web.AppRunner | None = None
        self._site: web.TCPSite | None = None
        self._coordinator: "Coordinator | None" = None

        self._current_round: int = 0
        self._updates: dict[str, ServerModelUpdateRequest] = {}
        self._lock = asyncio.Lock()
        self._is_training_done = False
        self._update_event: asyncio.Event | None = None
        # (version_id, round, binary?) -> encoded body
        self._model_cache: dict[tuple[str, int, bool], bytes] = {}

    # -- identity ---------------------------------------------------------
    @property
    def host(self) -> str:
        return self._host

    @property
    def port(self) -> int:
        return self._port

    @property
    def url(self) -> str:
        return f"http://{self._host}:{self._port}"

    @property
    def current_round(self) -> int:
        return self._current_round

    @property
    def num_updates(self) -> int:
        return len(self._updates)

    def set_coordinator(self, coordinator: "Coordinator") -> None:
        self._coordinator = coordinator

    # -- coordinator-facing update queue -------------------------------------
    def _event(self) -> asyncio.Event:
        if self._update_event is None:
            self._update_event = asyncio.Event()
        return self._update_event

    async def wait_for_updates(self, count: int, timeout: float) -> bool:
        """Return True as soon as ``count`` updates are stored (no 1 Hz polling)."""
        loop = asyncio.get_running_loop()
        deadline = loop.time() + timeout
        while len(self._updates) < count:
            remaining = deadline - loop.time()
            if remaining <= 0:
                return False
            event = self._event()
            event.clear()
            if len(self._updates) >= count:
                break
            try:
                await asyncio.wait_for(event.wait(), timeout=min(remaining, 1.0))
            except asyncio.TimeoutError:
                pass
        return True

    async def take_updates(self) -> dict[str, ServerModelUpdateRequest]:
        """Atomically remove and return everything submitted so far."""
        async with self._lock:
            taken, self._updates = self._updates, {}
            return taken

    async def clear_updates(self) -> None:
        async with self._lock:
            self._updates.clear()

    async def advance_round(self, round_number: int | None = None) -> int:
        """Move to the next round; pending updates for the old round are dropped."""
        async with self._lock:
            self._current_round = (
                self._current_round + 1 if round_number is None else int(round_number)
            )
            self._updates.clear()
            self._model_cache.clear()
            return self._current_round

    def invalidate_model_cache(self) -> None:
        self._model_cache.clear()

    # -- routing ----------------------------------------------------------------
    def _setup_routes(self) -> None:
        self._logger.debug("Setting up routes for HTTP server.")
        router = self._app.router
        router.add_get(self._endpoints.get_model, self._handle_get_model)
        router.add_post(self._endpoints.submit_update, self._handle_submit_update)
        router.add_get(self._endpoints.get_status, self._handle_get_status)
        router.add_get("/test", self._handle_test)

    def _convert_tensor(self, value: Any) -> Any:
        return codec.tensor_to_jsonable(value)

    async def _handle_test(self, request: web.Request) -> web.Response:
        return web.Response(text="Server is running")

    # -- GET /model ----------------------------------------------------------------
    @staticmethod
    def _wants_binary(request: web.Request) -> bool:
        return codec.BINARY_CONTENT_TYPE in request.headers.get("Accept", "")

    def _snapshot_model(self) -> dict[str, Any]:
        """Copy of the global state taken ON THE LOOP THREAD: aggregation and ``load_state_dict``
        run there too, so the executor thread that serialises the model can never see a
        half-updated tensor (and cache a torn body for this version)."""
        assert self._coordinator is not None
        state = self._coordinator.model_manager.model.state_dict()
        return {k: (v.detach().clone() if isinstance(v, torch.Tensor) else copy.deepcopy(v))
                for k, v in state.items()}

    def _encode_model(self, version_id: str, binary: bool,
                      state: dict[str, Any] | None = None) -> bytes:
        if state is None:
            state = self._snapshot_model()
        envelope = {
            "status": "success",
            "message": "Global model retrieved",
            "timestamp": _now(),
            "round_number": self._current_round,
            "version_id": version_id,
        }
        if binary:
            return codec.encode_frame(envelope, state)
        body: GlobalModelResponse = {
            **envelope,  # type: ignore[typeddict-item]
            "model_state": {k: self._convert_tensor(v) for k, v in state.items()},
        }
        # Same key order as the reference response.
        ordered = {k: body[k] for k in  # type: ignore[literal-required]
                   ("status", "message", "timestamp", "model_state", "round_number", "version_id")}
        return json.dumps(ordered).encode("utf-8")

    async def _handle_get_model(self, request: web.Request) -> web.Response:
        if not self._coordinator:
            return _error("Server not initialized with coordinator", 500)
        with self._logger.context("server.http", "get_model"):
            try:
                if self._is_training_done:
                    self._logger.info("Training complete. Sending termination signal.")
                    return web.json_response(
                        {
                            "status": "terminated",
                            "message": "Training is complete",
                            "timestamp": _now(),
                            "model_state": None,
                            "round_number": -1,
                        }
                    )
                manager = self._coordinator.model_manager
                version = manager.current_version
                if version is None:
                    version = manager.load_model()

                binary = self._wants_binary(request)
                key = (version.version_id, self._current_round, binary)
                body = self._model_cache.get(key)
                if body is None:
                    snapshot = self._snapshot_model()
                    body = await asyncio.get_running_loop().run_in_executor(
                        None, self._encode_model, version.version_id, binary, snapshot
                    )
                    self._model_cache = {key: body} if len(self._model_cache) > 4 else {
                        **self._model_cache, key: body}
                self._logger.debug(f"Serving model version {version.version_id}")
                return web.Response(
                    body=body,
                    content_type=codec.BINARY_CONTENT_TYPE if binary else "application/json",
                )
            except Exception as exc:
                self._logger.error(f"Error serving model: {exc}")
                return _error(str(exc), 500)

    # -- POST /update -----------------------------------------------------------------
    async def _read_update(self, request: web.Request) -> dict[str, Any]:
        if request.content_type == codec.BINARY_CONTENT_TYPE:
            blob = await request.read()
            header, state = await asyncio.get_running_loop().run_in_executor(
                None, codec.decode_frame, blob
            )
            if state is not None:
                header["model_state"] = state
            return header
        raw = await request.read()
        if len(raw) > (1 << 20):
            return await asyncio.get_running_loop().run_in_executor(None, json.loads, raw)
        return json.loads(raw)

    async def _handle_submit_update(self, request: web.Request) -> web.Response:
        with self._logger.context("server.http", "submit_update"):
            try:
                data = await self._read_update(request)
                missing = [k for k in _REQUIRED_UPDATE_KEYS if k not in data]
                if missing:
                    return _error(f"Missing keys: {', '.join(missing)}", 400)

                update: ServerModelUpdateRequest = {
                    "client_id": data["client_id"],
                    "round_number": data["round_number"],
                    "model_state": data["model_state"],
                    "metrics": data["metrics"],
                    "timestamp": data["timestamp"],
                    "status": data.get("status", "success"),
                    "message": data.get("message", data.get("mesage", "")),
                    "accepted": data.get("accepted", True),
                    "privacy_spent": data.get("privacy_spent"),
                }

                async with self._lock:
                    if update["round_number"] != self._current_round:
                        self._logger.warning(
                            f"Update round mismatch: expected {self._current_round}, "
                            f"got {update['round_number']} from client {update['client_id']}"
                        )
                        return _error("Invalid round number", 400)
                    client_id = update["client_id"]
                    self._updates[client_id] = update
                    self._logger.info(
                        f"Accepted update from client {client_id} for round {self._current_round}"
                    )
                    reply: ModelUpdateResponse = {
                        "status": "success",
                        "message": "Updated accepted",
                        "timestamp": _now(),
                        "update_id": f"update_{client_id}_{self._current_round}",
                        "accepted": True,
                    }
                self._event().set()
                return web.json_response(reply)
            except Exception as exc:
                self._logger.error(f"Error handling update: {exc}")
                return _error(str(exc), 500)

    # -- GET /status ------------------------------------------------------------------
    async def _handle_get_status(self, request: web.Request) -> web.Response:
        self._logger.debug("Processing /status request.")
        return web.json_response(
            {
                "status": "success",
                "message": "Server is running",
                "timestamp": _now(),
                "current_round": self._current_round,
                "num_updates": len(self._updates),
                "is_training_done": self._is_training_done,
            }
        )

    # -- lifecycle ------------------------------------------------------------------
    async def stop_training(self) -> None:
        self._is_training_done = True
        self._logger.info("Training completed. Broadcasting termination signal to clients.")

    async def start(self) -> None:
        self._logger.info("Starting HTTP server...")
        self._runner = web.AppRunner(self._app)
        await self._runner.setup()
        self._site = web.TCPSite(
            self._runner, self._host, self._port, reuse_address=True, reuse_port=True
        )
        await self._site.start()
        if self._port == 0 and self._runner.addresses:
            self._port = int(self._runner.addresses[0][1])
        self._logger.info(f"HTTP server started on {self._host}:{self._port}")

    async def stop(self) -> None:
        if self._site:
            await self._site.stop()
            self._site = None
        if self._runner:
            await self._runner.cleanup()
            self._runner = None
        self._logger.info("Server stopped")
