from .client import ClientEndpoints, HTTPClient
from .server import HTTPServer, ServerEndpoints
from .types import (
    ClientModelUpdateRequest,
    GlobalModelResponse,
    ModelUpdateResponse,
    ServerModelUpdateRequest,
)

__all__ = [
    "HTTPClient",
    "HTTPServer",
    "ClientEndpoints",
    "ServerEndpoints",
    "ClientModelUpdateRequest",
    "ServerModelUpdateRequest",
    "ModelUpdateResponse",
    "GlobalModelResponse",
]
