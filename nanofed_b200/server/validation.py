"""Update screening (shape / range / outlier) and RSA-PSS signing.

Parity: reference ``server/validation.py:15-212`` — ``ValidationResult`` members,
``ValidationConfig`` defaults, ``DefaultModelValidator.validate_shape`` /
``validate_range`` (all finite and **per-tensor** L2 ≤ ``max_norm``) /
``validate_statistics`` (z-score of the update's global L2 norm against the
reference updates' norms, skipped below ``min_clients_for_stats``), and
``SecurityManager`` (RSA-2048, PSS/SHA-256 over ``name:bytes`` of the
name-sorted state) — ``test_validation.py:65-166``.

Additions: ``validate_signature`` (declared in the reference's Protocol,
``validation.py:48-50``, but never implemented there) and ``max_update_size`` is
enforced by ``validate_range``.  On CUDA tensors the finite check and the norms
come from one fused K-STATS launch (``ops.update_stats`` / ``ops.segment_sumsq``)
instead of ``2·|keys|`` library kernels plus a host sync each.
"""

from __future__ import annotations

from dataclasses import dataclass
from enum import Enum, auto
from typing import Protocol, Sequence

import torch
from cryptography.exceptions import InvalidSignature
from cryptography.hazmat.primitives import hashes, serialization
from cryptography.hazmat.primitives.asymmetric import padding, rsa
from cryptography.hazmat.primitives.asymmetric.rsa import RSAPublicKey

from nanofed_b200.core import ModelUpdate
from nanofed_b200.utils import Logger


class ValidationResult(Enum):
    VALID = auto()
    INVALID_SHAPE = auto()
    INVALID_RANGE = auto()
    INVALID_SIGNATURE = auto()
    ANOMALOUS = auto()


@dataclass(frozen=True)
class ValidationConfig:
    max_norm: float = 10.0
    max_update_size: int = 1024 * 1024 * 100
    min_clients_for_stats: int = 5
    z_score_threshold: float = 2.0
    signature_required: bool = True


class ModelValidator(Protocol):
    def validate_shape(
        self, update: ModelUpdate, reference: dict[str, torch.Size]
    ) -> ValidationResult: ...

    def validate_range(
        self, update: ModelUpdate, config: ValidationConfig
    ) -> ValidationResult: ...

    def validate_statistics(
        self, update: ModelUpdate, reference_updates: Sequence[ModelUpdate]
    ) -> ValidationResult: ...

    def validate_signature(
        self, update: ModelUpdate, public_key: bytes
    ) -> ValidationResult: ...


def _update_norm(state: dict[str, torch.Tensor]) -> float:
    from nanofed_b200.privacy.mechanisms import global_l2_norm

    return float(global_l2_norm(state))


class DefaultModelValidator:
    def __init__(self, config: ValidationConfig) -> None:
        self._config = config
        self._logger = Logger()

    def validate_shape(
        self, update: ModelUpdate, reference: dict[str, torch.Size]
    ) -> ValidationResult:
        try:
            state = update["model_state"]
            for name, shape in reference.items():
                if name not in state:
                    self._logger.warning(f"Missing parameter: {name}")
                    return ValidationResult.INVALID_SHAPE
                if state[name].shape != shape:
                    self._logger.warning(
                        f"Shape mismatch for {name}: got {state[name].shape}, expected {shape}"
                    )
                    return ValidationResult.INVALID_SHAPE
            return ValidationResult.VALID
        except Exception as exc:
            self._logger.error(f"Shape validation failed: {exc}")
            return ValidationResult.INVALID_SHAPE

    def validate_range(self, update: ModelUpdate, config: ValidationConfig) -> ValidationResult:
        try:
            tensors = list(update["model_state"].values())
            nbytes = sum(t.numel() * t.element_size() for t in tensors)
            if nbytes > config.max_update_size:
                return ValidationResult.INVALID_RANGE
            if tensors and all(t.is_cuda for t in tensors):
                from nanofed_b200 import ops

                sumsq, bad = ops.per_tensor_stats(tensors)
                if bool(bad.any()) or bool((sumsq.sqrt() > config.max_norm).any()):
                    return ValidationResult.INVALID_RANGE
                return ValidationResult.VALID
            for t in tensors:
                if not torch.all(torch.isfinite(t)):
                    return ValidationResult.INVALID_RANGE
                if torch.linalg.vector_norm(t.to(torch.float32)) > config.max_norm:
                    return ValidationResult.INVALID_RANGE
            return ValidationResult.VALID
        except Exception as exc:
            self._logger.error(f"Range validation failed: {exc}")
            return ValidationResult.INVALID_RANGE

    def validate_statistics(
        self, update: ModelUpdate, reference_updates: Sequence[ModelUpdate]
    ) -> ValidationResult:
        if len(reference_updates) < self._config.min_clients_for_stats:
            return ValidationResult.VALID
        try:
            norms = torch.tensor([_update_norm(r["model_state"]) for r in reference_updates])
            mine = _update_norm(update["model_state"])
            z = abs(mine - norms.mean()) / (norms.std() + 1e-8)
            if z > self._config.z_score_threshold:
                return ValidationResult.ANOMALOUS
            return ValidationResult.VALID
        except Exception as exc:
            self._logger.error(f"Statistical validation failed: {exc}")
            return ValidationResult.ANOMALOUS

    def validate_signature(
        self, update: ModelUpdate, public_key: bytes, signature: bytes | None = None
    ) -> ValidationResult:
        """Check ``signature`` (or ``update["signature"]``) against ``public_key``."""
        sig = signature if signature is not None else update.get("signature")  # type: ignore[attr-defined]
        if sig is None:
            return (
                ValidationResult.INVALID_SIGNATURE
                if self._config.signature_required
                else ValidationResult.VALID
            )
        ok = SecurityManager.verify_with_key(update, sig, public_key)
        return ValidationResult.VALID if ok else ValidationResult.INVALID_SIGNATURE


def _pss() -> padding.PSS:
    return padding.PSS(mgf=padding.MGF1(hashes.SHA256()), salt_length=padding.PSS.MAX_LENGTH)


def _canonical_bytes(update: ModelUpdate) -> bytes:
    parts = []
    for name, tensor in sorted(update["model_state"].items()):
        parts.append(name.encode("utf-8") + b":" + tensor.detach().cpu().contiguous().numpy().tobytes())
    return b"".join(parts)


class SecurityManager:
    """Holds an RSA-2048 key pair; signs and verifies whole updates."""

    def __init__(self) -> None:
        self._private_key = rsa.generate_private_key(public_exponent=65537, key_size=2048)
        self._public_key = self._private_key.public_key()
        self._logger = Logger()

    def get_public_key(self) -> bytes:
        return self._public_key.public_bytes(
            encoding=serialization.Encoding.PEM,
            format=serialization.PublicFormat.SubjectPublicKeyInfo,
        )

    def sign_update(self, update: ModelUpdate) -> bytes:
        try:
            return self._private_key.sign(_canonical_bytes(update), _pss(), hashes.SHA256())
        except Exception as exc:
            self._logger.error(f"Failed to sign update: {exc}")
            raise

    @staticmethod
    def verify_with_key(update: ModelUpdate, signature: bytes, public_key: bytes) -> bool:
        try:
            key = serialization.load_pem_public_key(public_key)
            if not isinstance(key, RSAPublicKey):
                Logger().error("Unsupported public key type.")
                return False
            key.verify(signature, _canonical_bytes(update), _pss(), hashes.SHA256())
            return True
        except InvalidSignature:
            return False
        except Exception as exc:
            Logger().error(f"Signature verification failed: {exc}")
            return False

    def verify_signature(self, update: ModelUpdate, signature: bytes, public_key: bytes) -> bool:
        return self.verify_with_key(update, signature, public_key)
