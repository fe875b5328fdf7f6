from .base import BasePrivacyAccountant, PrivacyAccountant, PrivacySpent
from .gaussian import GaussianAccountant
from .rdp import RDPAccountant

__all__ = [
    "PrivacyAccountant",
    "PrivacySpent",
    "GaussianAccountant",
    "RDPAccountant",
    "BasePrivacyAccountant",
]
