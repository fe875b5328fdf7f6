"""Wire-level TypedDicts (reference ``communication/http/types.py:6-50``)."""

from __future__ import annotations

from typing import Any, Literal, TypedDict

from nanofed_b200.privacy.accountant.base import PrivacySpent  # noqa: F401  (reference re-export)


class BaseResponse(TypedDict):
    status: Literal["success", "error"]
    message: str
    timestamp: str


class ClientModelUpdateRequest(TypedDict):
    client_id: str
    round_number: int
    model_state: dict[str, Any]
    metrics: dict[str, float]
    timestamp: str


class ServerModelUpdateRequest(TypedDict, total=False):
    client_id: str
    round_number: int
    model_state: dict[str, Any]
    metrics: dict[str, float]
    timestamp: str
    status: Literal["success", "error"]
    message: str
    accepted: bool
    privacy_spent: Any


class ModelUpdateResponse(BaseResponse):
    update_id: str
    accepted: bool


class GlobalModelResponse(BaseResponse):
    model_state: dict[str, Any]
    round_number: int
    version_id: str
