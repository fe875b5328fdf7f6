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

from .peer import ClientSpec, PeerClient, PeerHub

__all__ = [
    "PeerClient",
    "PeerHub",
    "ClientSpec",
    "HTTPClient",
    "HTTPServer",
    "ClientEndpoints",
    "ServerEndpoints",
    "ClientModelUpdateRequest",
    "ServerModelUpdateRequest",
    "ModelUpdateResponse",
    "GlobalModelResponse",
]
