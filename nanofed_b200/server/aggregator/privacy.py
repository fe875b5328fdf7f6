"""Differentially-private aggregation (central / local DP, optional threshold sum).

Parity: reference ``server/aggregator/privacy.py:20-346`` —

* ``PrivacyAwareAggregationConfig`` = ``PrivacyConfig`` + ``privacy_type``,
  ``secure_aggregation``, ``min_clients``, ``dropout_tolerance``, ``clip_norm``
  (``:28-57``);
* ``_validate_updates`` raises ``ValueError`` (not ``AggregationError``) with
  "No updates provided" / "Not enough clients" / "Updates from different rounds" /
  "Inconsistent model architectures" / "Missing privacy budget for client …"
  (``:141-171``; matched by ``test_privacy_aggregation.py:181-216``);
* CENTRAL: every update is clipped to ``max_gradient_norm`` and noised with std
  ``σ·C/K``; K accountant events per round (``:179-194`` → ``mechanisms.py:106-129``);
* LOCAL: updates pass through; sample weights are multiplied by each client's
  normalised ε and renormalised (``:213-246``);
* with a secure-aggregation object: ``verify_shares`` then an **unweighted** sum
  (``:311-319``); ``THRESHOLD`` is auto-wired (``:135-139``);
* result metrics are the weighted **sum** of numeric metrics plus
  ``privacy_epsilon`` / ``privacy_delta`` (``:251-297``); ``result.model is model``;
  ``current_round`` is not advanced (``:342``).

Fix: a ``PrivacySpent`` instance contributes its ``delta_spent`` as
``delta`` (the reference copies ``epsilon_spent`` there, ``:222``); the value is
not used by the weighting, so behaviour is unchanged.

B200 path (central DP, stock Gaussian mechanism): the per-client work collapses
algebraically to

    out = Σ_k (w_k·c_k)·x_k + N(0, (σC/K)²·Σ_k w_k²),   c_k = min(C/(‖x_k‖+1e-6), 1)

so ONE fused launch (``ops.fedavg_reduce``: K-STATS norms → clip scalars → weighted
fp32 reduce → single Philox draw → cast) replaces K clip passes, K·|keys| noise
tensors and the weighted-sum loop.  It is distributionally identical to the
reference, not bitwise.  A user-supplied mechanism, a non-Gaussian generator, or
``fused=False`` selects the literal per-update path instead (Laplace noise does
not fold: a sum of Laplacians is not Laplacian).  Noise is applied to
floating-point entries only; integer entries (BatchNorm counters, incl. 0-d) are
clipped and averaged but not noised — the reference crashes on those
(``generators.py:22-23``).
"""

from __future__ import annotations

import math
from enum import Enum, auto
from typing import Protocol, Sequence, cast

import torch
from pydantic import ConfigDict, Field

from nanofed_b200 import ops
from nanofed_b200.core import ModelProtocol, ModelUpdate
from nanofed_b200.privacy.accountant import PrivacySpent
from nanofed_b200.privacy.config import PrivacyConfig
from nanofed_b200.privacy.mechanisms import (
    BasePrivacyMechanism,
    CentralPrivacyMechanism,
    PrivacyMechanismFactory,
    PrivacyType,
)
from nanofed_b200.privacy.noise import GaussianNoiseGenerator
from nanofed_b200.utils import Logger

from .base import AggregationResult, BaseAggregator, sample_count_of


class SecureAggregationType(Enum):
    NONE = auto()
    THRESHOLD = auto()
    HOMOMORPHIC = auto()


class PrivacyAwareAggregationConfig(PrivacyConfig):
    """Privacy budget plus aggregation-side knobs."""

    model_config = ConfigDict(arbitrary_types_allowed=True, frozen=True)

    privacy_type: PrivacyType = Field(default=PrivacyType.CENTRAL)
    secure_aggregation: SecureAggregationType = Field(default=SecureAggregationType.NONE)
    min_clients: int = Field(default=1, ge=1)
    dropout_tolerance: float = Field(default=0.0, ge=0.0, le=1.0)
    clip_norm: float = Field(default=1.0, gt=0.0)


class SecureAggregationProtocol(Protocol):
    def aggregate_shares(
        self, shares: Sequence[dict[str, torch.Tensor]]
    ) -> dict[str, torch.Tensor]: ...

    def verify_shares(self, shares: Sequence[dict[str, torch.Tensor]]) -> bool: ...


class ThresholdSecureAggregation:
    """Sum of shares, refused below ``min_clients`` (reference ``privacy.py:72-110``).

    The sum runs through the same fused reduction as FedAvg with unit weights
    and no normalisation (one launch, fp32 accumulation).
    """

    def __init__(self, min_clients: int) -> None:
        self._min_clients = min_clients
        self._logger = Logger()

    def aggregate_shares(
        self, shares: Sequence[dict[str, torch.Tensor]]
    ) -> dict[str, torch.Tensor]:
        if len(shares) < self._min_clients:
            raise ValueError(f"Not enough clients: {len(shares)} < {self._min_clients}")
        from nanofed_b200.runtime.flat import FlatLayout

        layout = FlatLayout.from_state(shares[0])
        slab = layout.stack(shares)
        ones = torch.ones(len(shares), dtype=torch.float32, device=slab.device)
        total = ops.fedavg_reduce(slab, ones, normalize=False)
        return layout.unflatten(total)

    def verify_shares(self, shares: Sequence[dict[str, torch.Tensor]]) -> bool:
        if len(shares) < self._min_clients:
            return False
        first = shares[0]
        for share in shares:
            if share.keys() != first.keys():
                return False
            if any(share[name].shape != first[name].shape for name in first):
                return False
        return True


def _epsilon_of(spent: object) -> float:
    if isinstance(spent, PrivacySpent):
        return spent.epsilon_spent
    if isinstance(spent, dict):
        return spent.get("epsilon", 1.0)
    raise TypeError(
        f"privacy_spent should be a dict or PrivacySpent instance, got {type(spent)}"
    )


class PrivacyAwareAggregator(BaseAggregator):
    """FedAvg with DP clipping/noise and optional threshold secure sum."""

    def __init__(
        self,
        config: PrivacyAwareAggregationConfig,
        privacy_mechanism: BasePrivacyMechanism | None = None,
        secure_aggregation: SecureAggregationProtocol | None = None,
        fused: bool | None = None,
    ) -> None:
        super().__init__()
        self._config = config
        self._user_mechanism = privacy_mechanism is not None
        self._privacy_mech = privacy_mechanism or PrivacyMechanismFactory.create(
            config.privacy_type, config=config
        )
        self._secure_agg = secure_aggregation
        if (
            config.secure_aggregation == SecureAggregationType.THRESHOLD
            and secure_aggregation is None
        ):
            self._secure_agg = ThresholdSecureAggregation(config.min_clients)
        self._fused = fused

    # -- validation ---------------------------------------------------------
    def _validate_updates(self, updates: Sequence[ModelUpdate]) -> None:
        if not updates:
            raise ValueError("No updates provided")
        if len(updates) < self._config.min_clients:
            raise ValueError(
                f"Not enough clients: {len(updates)} < {self._config.min_clients}"
            )
        first_round = updates[0].get("round_number")
        if any(u.get("round_number") != first_round for u in updates):
            raise ValueError("Updates from different rounds")
        names = updates[0]["model_state"].keys()
        if any(u["model_state"].keys() != names for u in updates):
            raise ValueError("Inconsistent model architectures")
        if self._config.privacy_type == PrivacyType.LOCAL:
            for u in updates:
                if u.get("privacy_spent") is None:
                    raise ValueError(f"Missing privacy budget for client {u['client_id']}")

    # -- per-update (literal) processing -----------------------------------
    def _process_local_updates(self, updates: Sequence[ModelUpdate]) -> Sequence[ModelUpdate]:
        return list(updates)

    def _process_central_updates(self, updates: Sequence[ModelUpdate]) -> Sequence[ModelUpdate]:
        k = len(updates)
        noised = []
        for u in updates:
            state = {
                n: (v if isinstance(v, torch.Tensor) else torch.as_tensor(v))
                for n, v in u["model_state"].items()
            }
            private = self._privacy_mech.add_noise(state, batch_size=k)
            noised.append(cast(ModelUpdate, {**u, "model_state": private}))
        return noised

    # -- weights / metrics -----------------------------------------------
    def _compute_weights(self, updates: Sequence[ModelUpdate]) -> list[float]:
        counts = self._sample_counts(updates)
        total = sum(counts)
        weights = [c / total for c in counts]
        if self._config.privacy_type == PrivacyType.LOCAL:
            eps = [
                _epsilon_of(u.get("privacy_spent", {"epsilon": 1.0, "delta": 1e-5}))
                for u in updates
            ]
            eps_total = sum(eps)
            if eps_total > 0:
                # Larger ε spent ⇒ less noise in that update ⇒ more say.
                weights = [w * (e / eps_total) for w, e in zip(weights, eps)]
                norm = sum(weights)
                weights = [w / norm for w in weights]
        self._logger.debug(f"Computed weights: {weights}")
        return weights

    def _aggregate_metrics(
        self, updates: Sequence[ModelUpdate], weights: list[float] | None = None
    ) -> dict[str, float]:
        if not updates:
            return {}
        if weights is None:
            counts = [u["metrics"].get("samples_processed", 1) for u in updates]
            total = sum(counts)
            weights = [c / total for c in counts]
        names: set[str] = set()
        for u in updates:
            names.update(
                k for k, v in u.get("metrics", {}).items() if isinstance(v, (int, float))
            )
        out: dict[str, float] = {
            name: sum(u["metrics"].get(name, 0) * w for u, w in zip(updates, weights))
            for name in names
        }
        if self._privacy_mech is not None:
            spent = self._privacy_mech.get_privacy_spent()
            out["privacy_epsilon"] = spent.epsilon_spent
            out["privacy_delta"] = spent.delta_spent
        return out

    # -- fused central path --------------------------------------------------
    def _can_fuse(self) -> bool:
        if self._fused is False:
            return False
        if self._config.privacy_type != PrivacyType.CENTRAL or self._secure_agg is not None:
            return False
        mech = self._privacy_mech
        stock = type(mech) is CentralPrivacyMechanism and type(
            getattr(mech, "_noise_gen", None)
        ) is GaussianNoiseGenerator
        return stock

    def _aggregate_fused(
        self, model: ModelProtocol, updates: Sequence[ModelUpdate]
    ) -> None:
        mech = self._privacy_mech
        cfg = mech.config
        k = len(updates)
        layout, slab = self._pack(model, updates)
        counts = torch.tensor(self._sample_counts(updates), dtype=torch.float32,
                              device=slab.device)
        sumsq, _ = ops.update_stats(slab)
        base_std = cfg.noise_multiplier * cfg.max_gradient_norm / k
        gen = mech._noise_gen
        if slab.is_cuda:
            seed, offset = gen._advance(layout.total)
            noise = None
        else:
            # Draw through the mechanism's (seedable) CPU generator so runs are
            # reproducible the same way the reference's are.
            seed = offset = 0
            noise = torch.zeros(layout.total, dtype=torch.float32)
            for e in layout.entries:
                if e.is_float and e.numel:
                    noise[e.offset : e.offset + e.numel] = gen.generate(
                        (e.numel,), scale=1.0
                    )
        flat = ops.fedavg_reduce(
            slab,
            counts,
            sumsq=sumsq,
            clip_norm=float(cfg.max_gradient_norm),
            noise_std=base_std,
            noise_mask=layout.float_mask(slab.device).to(torch.uint8),
            seed=seed,
            offset=offset,
            unit_noise=noise,
        )
        model.load_state_dict(layout.unflatten(flat, clone=False))
        for _ in range(k):
            mech.accountant.add_noise_event(sigma=cfg.noise_multiplier, samples=k)

    # -- entry point -----------------------------------------------------------
    def aggregate(self, model: ModelProtocol, updates: Sequence[ModelUpdate]) -> AggregationResult:
        self._validate_updates(updates)

        if self._can_fuse():
            self._aggregate_fused(model, updates)
            processed: Sequence[ModelUpdate] = updates
        else:
            if self._config.privacy_type == PrivacyType.LOCAL:
                processed = self._process_local_updates(updates)
            else:
                processed = self._process_central_updates(updates)

            if self._secure_agg is not None:
                shares = [
                    {n: (v if isinstance(v, torch.Tensor) else torch.as_tensor(v))
                     for n, v in u["model_state"].items()}
                    for u in processed
                ]
                if not self._secure_agg.verify_shares(shares):
                    raise ValueError("Invalid shares for secure aggregation")
                state = self._secure_agg.aggregate_shares(shares)
            else:
                weights = self._compute_weights(processed)
                layout, slab = self._pack(model, processed)
                coef = torch.tensor(weights, dtype=torch.float32, device=slab.device)
                flat = ops.fedavg_reduce(slab, coef, normalize=False)
                state = layout.unflatten(flat, clone=False)
            model.load_state_dict(state)

        return AggregationResult(
            model=model,
            round_number=self._current_round,
            num_clients=len(updates),
            timestamp=self._get_timestamp(),
            metrics=self._aggregate_metrics(processed),
        )
