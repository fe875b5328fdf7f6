from .base import AggregationResult, BaseAggregator
from .fedavg import FedAvgAggregator
from .privacy import (
    PrivacyAwareAggregationConfig,
    PrivacyAwareAggregator,
    SecureAggregationType,
    ThresholdSecureAggregation,
)
from .secure import (
    BaseSecureAggregator,
    HomomorphicSecureAggregator,
    SecureAggregationConfig,
    SecureMaskingAggregator,
)

__all__ = [
    "BaseAggregator",
    "AggregationResult",
    "FedAvgAggregator",
    "PrivacyAwareAggregator",
    "PrivacyAwareAggregationConfig",
    "SecureAggregationType",
    "ThresholdSecureAggregation",
    "SecureAggregationConfig",
    "SecureMaskingAggregator",
    "BaseSecureAggregator",
    "HomomorphicSecureAggregator",
]
