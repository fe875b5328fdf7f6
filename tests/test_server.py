"""Server-side contract: aggregators, model manager, fault tolerance, validation, secure agg."""

import json

import pytest
import torch
import torch.nn as nn

from nanofed_b200.core import AggregationError, ModelManagerError
from nanofed_b200.privacy.mechanisms import PrivacyType
from nanofed_b200.server import (FaultTolerantCoordinator, FedAvgAggregator, FileStateStore,
                                 ModelManager, RoundState, SimpleRecoveryStrategy)
from nanofed_b200.server.aggregator import (HomomorphicSecureAggregator,
                                            PrivacyAwareAggregationConfig, PrivacyAwareAggregator,
                                            SecureAggregationConfig, SecureAggregationType,
                                            SecureMaskingAggregator, ThresholdSecureAggregation)
from nanofed_b200.server.validation import (DefaultModelValidator, SecurityManager,
                                            ValidationConfig, ValidationResult)
from nanofed_b200.utils import get_current_time


class Tiny(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(4, 3)
        self.bn = nn.BatchNorm1d(3)

    def forward(self, x):
        return self.bn(self.fc(x))


def _update(cid, state, n=None, rnd=0, **extra):
    metrics = {"loss": 0.5, "accuracy": 0.5}
    if n is not None:
        metrics["samples_processed"] = n
    return {"client_id": cid, "round_number": rnd, "model_state": state, "metrics": metrics,
            "timestamp": get_current_time(), **extra}


def test_fedavg_weighted_mean_includes_buffers_and_lists():
    model = Tiny()
    states = []
    for k in range(3):
        m = Tiny()
        with torch.no_grad():
            for p in m.parameters():
                p.fill_(float(k + 1))
            m.bn.running_mean.fill_(10.0 * (k + 1))
            m.bn.num_batches_tracked.fill_(3 * (k + 1))
        states.append(m.state_dict())
    updates = [_update(f"c{k}", s, n) for k, (s, n) in enumerate(zip(states, (100, 200, 700)))]
    updates[1]["model_state"] = {k: v.tolist() for k, v in states[1].items()}  # nested lists
    agg = FedAvgAggregator()
    result = agg.aggregate(model, updates)
    w = torch.tensor([0.1, 0.2, 0.7])
    want = float((w * torch.tensor([1.0, 2.0, 3.0])).sum())
    assert torch.allclose(model.fc.weight, torch.full((3, 4), want))
    assert torch.allclose(model.bn.running_mean, torch.full((3,), 10 * want))
    assert model.bn.num_batches_tracked.dtype == torch.int64
    assert int(model.bn.num_batches_tracked) == int(3 * want)       # averaged, truncated on load
    assert result.model is model and result.num_clients == 3 and agg.current_round == 1
    assert abs(result.metrics["samples_processed"] - (10 + 40 + 490)) < 1e-6


def test_fedavg_missing_counts_and_errors():
    model = nn.Linear(2, 1)
    s = lambda v: {"weight": torch.full((1, 2), v), "bias": torch.full((1,), v)}  # noqa: E731
    FedAvgAggregator().aggregate(model, [_update("a", s(1.0)), _update("b", s(3.0))])
    assert torch.allclose(model.weight, torch.full((1, 2), 2.0))
    with pytest.raises(AggregationError, match="Updates from different rounds: {1, 2}"):
        FedAvgAggregator().aggregate(model, [_update("a", s(1.0), 1, 1), _update("b", s(1.0), 1, 2)])
    with pytest.raises(AggregationError, match="Inconsistent model architectures in updates"):
        FedAvgAggregator().aggregate(model, [_update("a", s(1.0), 1),
                                             _update("b", {"weight": torch.ones(1, 2)}, 1)])
    with pytest.raises(AggregationError):
        FedAvgAggregator().aggregate(model, [])


def _dp_config(**kw):
    base = dict(epsilon=5.0, noise_multiplier=1.0, max_gradient_norm=1.0,
                privacy_type=PrivacyType.CENTRAL, min_clients=2)
    base.update(kw)
    return PrivacyAwareAggregationConfig(**base)


def test_privacy_aggregator_fused_central_matches_algebra():
    model = Tiny()   # has a 0-d int64 buffer: the reference crashes here
    states = [Tiny().state_dict() for _ in range(4)]
    updates = [_update(f"c{k}", s, 50 * (k + 1)) for k, s in enumerate(states)]
    agg = PrivacyAwareAggregator(_dp_config(noise_multiplier=1e-6))   # ~no noise → exact algebra
    result = agg.aggregate(model, updates)
    w = torch.tensor([50.0, 100, 150, 200]) / 500
    norms = [torch.sqrt(sum(v.float().pow(2).sum() for v in s.values())) for s in states]
    c = [min(1.0 / (float(n) + 1e-6), 1.0) for n in norms]
    want = sum(wi * ci * s["fc.weight"] for wi, ci, s in zip(w, c, states))
    assert torch.allclose(model.fc.weight, want, atol=1e-5)
    assert result.model is model and agg.current_round == 0
    assert {"privacy_epsilon", "privacy_delta", "accuracy"} <= set(result.metrics)
    assert agg._privacy_mech.accountant._event_count == 4           # K events per round


def test_privacy_aggregator_validation_and_local_weights():
    model = nn.Linear(2, 1)
    s = {"weight": torch.ones(1, 2), "bias": torch.ones(1)}
    agg = PrivacyAwareAggregator(_dp_config(min_clients=3))
    with pytest.raises(ValueError, match="Not enough clients"):
        agg.aggregate(model, [_update("a", s, 1), _update("b", s, 1)])
    agg = PrivacyAwareAggregator(_dp_config())
    with pytest.raises(ValueError, match="Updates from different rounds"):
        agg.aggregate(model, [_update("a", s, 1, 0), _update("b", s, 1, 1)])
    with pytest.raises(ValueError, match="Inconsistent model architectures"):
        agg.aggregate(model, [_update("a", s, 1), _update("b", {"weight": s["weight"]}, 1)])
    local = PrivacyAwareAggregator(_dp_config(privacy_type=PrivacyType.LOCAL))
    with pytest.raises(ValueError, match="Missing privacy budget"):
        local.aggregate(model, [_update("a", s, 1), _update("b", s, 1)])
    ups = [_update("a", s, 100, privacy_spent={"epsilon": 0.1, "delta": 1e-5}),
           _update("b", s, 100, privacy_spent={"epsilon": 0.9, "delta": 1e-5})]
    w = local._compute_weights(ups)
    assert abs(sum(w) - 1) < 1e-9 and w[1] > w[0]
    local.aggregate(model, ups)
    assert torch.allclose(model.weight, torch.ones(1, 2))


def test_threshold_and_crypto_aggregators():
    shares = [{"w": torch.full((2, 2), float(k))} for k in range(1, 4)]
    th = ThresholdSecureAggregation(min_clients=3)
    assert th.verify_shares(shares) and torch.allclose(th.aggregate_shares(shares)["w"],
                                                       torch.full((2, 2), 6.0))
    with pytest.raises(ValueError, match="Not enough clients"):
        th.aggregate_shares(shares[:2])
    assert not th.verify_shares(shares[:2])
    model = nn.Linear(2, 2, bias=False)
    agg = PrivacyAwareAggregator(_dp_config(privacy_type=PrivacyType.LOCAL, min_clients=3,
                                            secure_aggregation=SecureAggregationType.THRESHOLD))
    ups = [_update(f"c{k}", {"weight": s["w"]}, 1, privacy_spent={"epsilon": 1.0})
           for k, s in enumerate(shares)]
    agg.aggregate(model, ups)
    assert torch.allclose(model.weight, torch.full((2, 2), 6.0))    # unweighted sum

    cfg = SecureAggregationConfig(min_clients=1, key_size=2048)
    he = HomomorphicSecureAggregator(cfg)
    t = {"w": torch.randn(95, 2)}                                    # 760 B = exactly 4 chunks
    enc = he.encrypt_update(t)
    assert len(enc["w"]) == -(-760 // he._chunk_size) and he._chunk_size == 190
    dec = he.decrypt_aggregate(he.aggregate_encrypted([enc]))
    assert torch.allclose(dec["w"], t["w"], atol=1e-4)
    enc["w"][0] = b"\0" * len(enc["w"][0])
    with pytest.raises(ValueError, match="Decryption failed for .*"):
        he.decrypt_aggregate(enc)

    mask = SecureMaskingAggregator(SecureAggregationConfig(min_clients=2))
    a, b = {"w": torch.randn(3, 3)}, {"w": torch.randn(3, 3)}
    total = mask.decrypt_aggregate(mask.aggregate_encrypted([mask.encrypt_update(a),
                                                             mask.encrypt_update(b)]))
    assert torch.allclose(total["w"], a["w"] + b["w"], atol=1e-4)
    with pytest.raises(ValueError, match="Need at least .* clients"):
        mask.aggregate_encrypted([mask.encrypt_update(a)])


def test_model_manager_layout_and_roundtrip(tmp_path):
    model = Tiny()
    mm = ModelManager(model)
    with pytest.raises(ModelManagerError):
        mm.save_model({})
    mdir, cdir = tmp_path / "models" / "models", tmp_path / "models" / "configs"
    mdir.mkdir(parents=True), cdir.mkdir(parents=True)
    mm.set_dirs(mdir, cdir)                                  # seeds version _001
    (first,) = mm.list_versions()
    assert first.version_id.startswith("model_v_") and first.version_id.endswith("_001")
    assert first.config == {"name": "default", "version": "1.0"}
    with torch.no_grad():
        model.fc.weight.fill_(7.0)
    v2 = mm.save_model({"round_id": 0, "when": get_current_time()}, metrics={"loss": 0.1})
    record = json.loads((cdir / f"{v2.version_id}.json").read_text())
    assert set(record) == {"version_id", "timestamp", "config", "metrics"}
    state = torch.load(mdir / f"{v2.version_id}.pt", weights_only=True)
    assert set(state) == set(model.state_dict()) and all(t.device.type == "cpu" for t in state.values())
    with torch.no_grad():
        model.fc.weight.zero_()
    assert mm.load_model().version_id == v2.version_id       # latest = lexicographically last
    assert torch.allclose(model.fc.weight, torch.full((3, 4), 7.0))
    mm.load_model(first.version_id)
    with pytest.raises(ModelManagerError, match="not found"):
        mm.load_model("nope")


def test_fault_tolerance_store_and_recovery(tmp_path):
    store = FileStateStore(tmp_path)
    ft = FaultTolerantCoordinator(tmp_path, state_store=store)
    upd = {"c1": _update("c1", {"w": torch.ones(2)}, 5)}
    for rid, st in ((2, RoundState.COMPLETED), (10, RoundState.COMPLETED), (11, RoundState.FAILED)):
        ft.checkpoint_round(rid, upd, f"v{rid}", {"w": torch.full((2,), float(rid))}, st)
    assert (tmp_path / "checkpoints" / "round_10" / "metadata.json").exists()
    assert [c.round_id for c in store.list_checkpoints()] == [2, 10, 11]   # numeric order
    meta, state = ft.restore_round(2)
    assert meta.global_model_version == "v2" and torch.equal(state["w"], torch.full((2,), 2.0))
    assert isinstance(meta.client_updates["c1"]["model_state"]["w"], torch.Tensor)
    assert ft.restore_round(99) is None
    meta, state = ft.handle_failure(TimeoutError("x"), 12)
    assert meta.round_id == 10
    assert ft.handle_failure(ValueError("x"), 12) is None
    assert SimpleRecoveryStrategy().get_recovery_point([]) is None


def test_validation_and_signatures():
    v = DefaultModelValidator(ValidationConfig(max_norm=10.0, min_clients_for_stats=3))
    good = _update("a", {"w": torch.ones(4, 4) * 0.1})
    assert v.validate_shape(good, {"w": torch.Size([4, 4])}) == ValidationResult.VALID
    assert v.validate_shape(good, {"w": torch.Size([4, 5])}) == ValidationResult.INVALID_SHAPE
    assert v.validate_shape(good, {"z": torch.Size([1])}) == ValidationResult.INVALID_SHAPE
    cfg = ValidationConfig(max_norm=10.0)
    assert v.validate_range(good, cfg) == ValidationResult.VALID
    assert v.validate_range(_update("b", {"w": torch.tensor([float("nan")])}), cfg) \
        == ValidationResult.INVALID_RANGE
    assert v.validate_range(_update("c", {"w": torch.ones(4, 4) * 100}), cfg) \
        == ValidationResult.INVALID_RANGE
    refs = [_update(f"r{i}", {"w": torch.ones(4, 4) * (0.1 + 0.001 * i)}) for i in range(5)]
    assert v.validate_statistics(good, refs) == ValidationResult.VALID
    assert v.validate_statistics(_update("d", {"w": torch.ones(4, 4) * 5}), refs) \
        == ValidationResult.ANOMALOUS
    assert v.validate_statistics(good, refs[:2]) == ValidationResult.VALID   # too few for stats
    sm, other = SecurityManager(), SecurityManager()
    sig = sm.sign_update(good)
    assert sm.verify_signature(good, sig, sm.get_public_key())
    assert not sm.verify_signature(good, sig, other.get_public_key())
    assert not sm.verify_signature(_update("e", {"w": torch.zeros(4, 4)}), sig, sm.get_public_key())
    assert v.validate_signature(good, sm.get_public_key(), sig) == ValidationResult.VALID
    assert v.validate_signature(good, sm.get_public_key()) == ValidationResult.INVALID_SIGNATURE
