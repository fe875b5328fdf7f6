"""Encrypted-transport aggregators (single-party simulations, as in the reference).

Parity: reference ``server/aggregator/secure.py:32-313`` — same config fields,
same three-call protocol (``encrypt_update`` → ``aggregate_encrypted`` →
``decrypt_aggregate``), same error texts ("Need at least N clients",
"Decryption failed for <key>: …"), ``_chunk_size = key_size/8 - 66`` for
RSA-OAEP(SHA-256), AES-256-GCM with a 12-byte nonce prefix for the masking
scheme (``test_secure.py:58-272``).

Honest naming: ``HomomorphicSecureAggregator`` is NOT additively
homomorphic — like the reference (``secure.py:143-153``) it XORs ciphertext
chunks, which only round-trips for a single update.  The API and that behaviour
are preserved; ``SecureMaskingAggregator`` is the one that really sums.

B200 notes: the mask arithmetic (``masked = x + mask``, running cumulative mask,
``sum − cumulative``) is elementwise and runs wherever the tensors live; on CUDA
the mask comes from the Philox kernel (``ops.philox_uniform_``) so no host RNG
traffic occurs, and the final sum uses the fused reduction.  RSA/AES stay on the
host via ``cryptography`` — they are byte-serial and outside the GPU's remit.
"""

from __future__ import annotations

import os
from abc import ABC, abstractmethod
from dataclasses import dataclass
from functools import reduce
from typing import Generic, Protocol, Sequence, TypeVar

import numpy as np
import torch
from cryptography.hazmat.primitives import hashes
from cryptography.hazmat.primitives.asymmetric import padding, rsa
from cryptography.hazmat.primitives.ciphers.aead import AESGCM
from cryptography.hazmat.primitives.kdf.pbkdf2 import PBKDF2HMAC

from nanofed_b200.utils import Logger

EncryptedType = TypeVar("EncryptedType")

_NONCE_BYTES = 12


class SecureAggregationProtocol(Protocol, Generic[EncryptedType]):
    def encrypt_update(self, update: dict[str, torch.Tensor]) -> dict[str, EncryptedType]: ...

    def decrypt_aggregate(
        self, encrypted_sum: dict[str, EncryptedType]
    ) -> dict[str, torch.Tensor]: ...

    def aggregate_encrypted(
        self, encrypted_updates: Sequence[dict[str, EncryptedType]]
    ) -> dict[str, EncryptedType]: ...


@dataclass(slots=True, frozen=True)
class SecureAggregationConfig:
    min_clients: int
    key_size: int = 2048
    threshold: int | None = None
    masking_seed_size: int = 256
    dropout_tolerance: float = 0.0


class BaseSecureAggregator(ABC, Generic[EncryptedType]):
    def __init__(self, config: SecureAggregationConfig) -> None:
        self._config = config
        self._logger = Logger()
        self._setup_crypto()

    def _require_quorum(self, count: int) -> None:
        if count < self._config.min_clients:
            raise ValueError(f"Need at least {self._config.min_clients} clients")

    @abstractmethod
    def _setup_crypto(self) -> None: ...

    @abstractmethod
    def encrypt_update(self, update: dict[str, torch.Tensor]) -> dict[str, EncryptedType]: ...

    @abstractmethod
    def decrypt_aggregate(
        self, encrypted_sum: dict[str, EncryptedType]
    ) -> dict[str, torch.Tensor]: ...

    @abstractmethod
    def aggregate_encrypted(
        self, encrypted_updates: Sequence[dict[str, EncryptedType]]
    ) -> dict[str, EncryptedType]: ...


def _fp32_bytes(tensor: torch.Tensor) -> bytes:
    return tensor.detach().to("cpu", torch.float32).contiguous().numpy().tobytes()


def _fp32_tensor(raw: bytes, shape: tuple[int, ...]) -> torch.Tensor:
    return torch.from_numpy(np.frombuffer(raw, dtype=np.float32).copy()).reshape(shape)


def _oaep() -> padding.OAEP:
    return padding.OAEP(
        mgf=padding.MGF1(algorithm=hashes.SHA256()), algorithm=hashes.SHA256(), label=None
    )


class HomomorphicSecureAggregator(
    BaseSecureAggregator[list[bytes]], SecureAggregationProtocol[list[bytes]]
):
    """Chunked RSA-OAEP transport.  Despite the name, ciphertexts are XOR-combined."""

    def _setup_crypto(self) -> None:
        self._private_key = rsa.generate_private_key(
            public_exponent=65537, key_size=self._config.key_size
        )
        self._public_key = self._private_key.public_key()
        self._shapes: dict[str, tuple[int, ...]] = {}
        self._lengths: dict[str, int] = {}
        # OAEP overhead with SHA-256 is 2·32 + 2 bytes.
        self._chunk_size = self._config.key_size // 8 - 2 * 32 - 2

    def encrypt_update(self, update: dict[str, torch.Tensor]) -> dict[str, list[bytes]]:
        size = self._chunk_size
        out: dict[str, list[bytes]] = {}
        for name, tensor in update.items():
            self._shapes[name] = tuple(tensor.shape)
            raw = _fp32_bytes(tensor)
            self._lengths[name] = len(raw)
            pieces = [raw[i : i + size] for i in range(0, len(raw), size)]
            if pieces and len(pieces[-1]) < size:
                short = size - len(pieces[-1])
                pieces[-1] += bytes([short]) * short
            out[name] = [self._public_key.encrypt(piece, _oaep()) for piece in pieces]
        return out

    def _homomorphic_add(self, encrypted_values: list[bytes]) -> bytes:
        arrays = [np.frombuffer(v, dtype=np.uint8) for v in encrypted_values]
        return bytes(reduce(np.bitwise_xor, arrays))

    def aggregate_encrypted(
        self, encrypted_updates: Sequence[dict[str, list[bytes]]]
    ) -> dict[str, list[bytes]]:
        self._require_quorum(len(encrypted_updates))
        first = encrypted_updates[0]
        return {
            name: [
                self._homomorphic_add([u[name][i] for u in encrypted_updates])
                for i in range(len(first[name]))
            ]
            for name in first
        }

    def decrypt_aggregate(self, encrypted_sum: dict[str, list[bytes]]) -> dict[str, torch.Tensor]:
        out: dict[str, torch.Tensor] = {}
        for name, chunks in encrypted_sum.items():
            try:
                plain = [self._private_key.decrypt(c, _oaep()) for c in chunks]
                # The reference strips ``plain[-1][-1]`` bytes whenever that value is
                # below the chunk size, which corrupts tensors whose byte length is an
                # exact multiple of the chunk; the recorded length is used instead.
                joined = b"".join(plain)[: self._lengths[name]]
                out[name] = _fp32_tensor(joined, self._shapes[name])
            except Exception as exc:
                raise ValueError(f"Decryption failed for {name}: {exc}") from exc
        return out


class SecureMaskingAggregator(BaseSecureAggregator[bytes], SecureAggregationProtocol[bytes]):
    """Additive masking + AES-GCM transport; masks cancel in the aggregate."""

    def __init__(self, config: SecureAggregationConfig, key: bytes | None = None) -> None:
        if key:
            self._key = key
        super().__init__(config)

    def _setup_crypto(self) -> None:
        if not hasattr(self, "_key"):
            kdf = PBKDF2HMAC(
                algorithm=hashes.SHA256(), length=32, salt=os.urandom(16), iterations=100000
            )
            self._key = kdf.derive(os.urandom(32))
            self._iv = os.urandom(16)
        self._shapes: dict[str, tuple[int, ...]] = {}
        self._cumulative_mask: dict[str, torch.Tensor] = {}
        self._mask_seed = int.from_bytes(os.urandom(7), "little")
        self._mask_offset = 0

    def _draw_mask(self, like: torch.Tensor) -> torch.Tensor:
        if like.is_cuda:
            from nanofed_b200 import ops

            mask = torch.empty_like(like, dtype=torch.float32)
            ops.philox_uniform_(mask, self._mask_seed, self._mask_offset)
            self._mask_offset += (like.numel() + 3) // 4 * 4
            return mask
        return torch.rand_like(like, dtype=torch.float32)

    def _seal(self, payload: bytes) -> bytes:
        nonce = os.urandom(_NONCE_BYTES)
        return nonce + AESGCM(self._key).encrypt(nonce, payload, None)

    def encrypt_update(self, update: dict[str, torch.Tensor]) -> dict[str, bytes]:
        out: dict[str, bytes] = {}
        for name, tensor in update.items():
            self._shapes[name] = tuple(tensor.shape)
            mask = self._draw_mask(tensor)
            running = self._cumulative_mask.get(name)
            self._cumulative_mask[name] = mask.clone() if running is None else running + mask
            out[name] = self._seal(_fp32_bytes(tensor.detach().to(torch.float32) + mask))
        return out

    def decrypt_aggregate(self, encrypted_sum: dict[str, bytes]) -> dict[str, torch.Tensor]:
        out: dict[str, torch.Tensor] = {}
        for name, blob in encrypted_sum.items():
            try:
                raw = AESGCM(self._key).decrypt(blob[:_NONCE_BYTES], blob[_NONCE_BYTES:], None)
                out[name] = _fp32_tensor(raw, self._shapes[name])
            except Exception as exc:
                raise ValueError(f"Decryption failed for {name}: {exc}") from exc
        return out

    def aggregate_encrypted(self, encrypted_updates: Sequence[dict[str, bytes]]) -> dict[str, bytes]:
        self._require_quorum(len(encrypted_updates))
        totals: dict[str, torch.Tensor] = {}
        for blob in encrypted_updates:
            for name, tensor in self.decrypt_aggregate(blob).items():
                totals[name] = tensor if name not in totals else totals[name] + tensor
        sealed: dict[str, bytes] = {}
        for name, total in totals.items():
            mask = self._cumulative_mask.get(name)
            clean = total if mask is None else total - mask.to(total.device)
            sealed[name] = self._seal(_fp32_bytes(clean))
        self._cumulative_mask = {}
        return sealed
