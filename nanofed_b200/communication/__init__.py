"""Layer L3: cross-host transport (reference ``nanofed/communication``)."""

from .http import (
    ClientEndpoints,
    ClientModelUpdateRequest,
    GlobalModelResponse,
    HTTPClient,
    HTTPServer,
    ModelUpdateResponse,
    ServerEndpoints,
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
