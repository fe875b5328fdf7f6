"""Privacy layer contract: config bounds, noise, accountants, mechanisms."""

import math

import pytest
import torch
from pydantic import ValidationError

from nanofed_b200.privacy import (GaussianNoiseGenerator, LaplacianNoiseGenerator, NoiseType,
                                  PrivacyConfig, PrivacyError)
from nanofed_b200.privacy.accountant import GaussianAccountant, PrivacySpent, RDPAccountant
from nanofed_b200.privacy.mechanisms import (CentralPrivacyMechanism, LocalPrivacyMechanism,
                                             PrivacyMechanismFactory, PrivacyType)


def test_config_defaults_bounds_and_frozen():
    cfg = PrivacyConfig()
    assert (cfg.epsilon, cfg.delta, cfg.noise_type) == (1.0, 1e-5, NoiseType.GAUSSIAN)
    PrivacyConfig(epsilon=0.01, delta=1e-10)
    PrivacyConfig(epsilon=10.0, delta=0.1)
    for bad in ({"epsilon": -1.0}, {"delta": 2.0}, {"max_gradient_norm": -1.0},
                {"noise_multiplier": 0.0}):
        with pytest.raises(ValidationError):
            PrivacyConfig(**bad)
    with pytest.raises(ValidationError):
        cfg.epsilon = 2.0


@pytest.mark.parametrize("gen_cls,stat", [(GaussianNoiseGenerator, "std"),
                                          (LaplacianNoiseGenerator, "mean_abs")])
def test_noise_generators(gen_cls, stat):
    a, b = gen_cls(seed=7), gen_cls(seed=7)
    assert torch.equal(a.generate((100,), 1.0), b.generate((100,), 1.0))
    x = gen_cls(seed=1).generate((200_000,), 2.0)
    value = x.std() if stat == "std" else x.abs().mean()
    assert abs(float(value) - 2.0) < 0.05
    assert gen_cls().generate((), 1.0).shape == ()          # 0-d accepted (reference rejects)
    for shape, scale in ((None, 1.0), ([3], 1.0), ((0,), 1.0), ((3,), 0.0), ((3,), "x")):
        with pytest.raises(ValueError):
            gen_cls().generate(shape, scale)


def test_gaussian_accountant_closed_form():
    cfg = PrivacyConfig(epsilon=5.0, delta=1e-5, max_gradient_norm=1.0)
    acc = GaussianAccountant(cfg)
    assert acc.get_privacy_spent() == PrivacySpent(0.0, 0.0)
    c = math.sqrt(2 * math.log(1.25 / 1e-5))
    acc.add_noise_event(1.1, 32)
    acc.add_noise_event(2.0, 32)
    spent = acc.get_privacy_spent()
    assert math.isclose(spent.epsilon_spent, c / 1.1 + c / 2.0, rel_tol=1e-9)
    assert spent.delta_spent == 1e-5
    # q = min(samples / max_gradient_norm, 1): the reference's (odd) sampling rate
    big = GaussianAccountant(PrivacyConfig(max_gradient_norm=64.0))
    big.add_noise_event(1.0, 16)
    assert math.isclose(big.get_privacy_spent().epsilon_spent,
                        math.sqrt(2 * math.log(1.25 / 1e-5)) * 0.25, rel_tol=1e-9)
    with pytest.raises(ValueError, match="Number of samples must be positive"):
        acc.add_noise_event(1.0, 0)
    with pytest.raises(ValueError, match="Noise multiplier must be positive"):
        acc.add_noise_event(0.0, 1)
    assert acc.validate_budget() is (spent.epsilon_spent <= 5.0)


def test_rdp_accountant():
    cfg = PrivacyConfig(epsilon=10.0)
    with pytest.raises(PrivacyError):
        RDPAccountant(cfg, orders=[1.0, 2.0])
    acc = RDPAccountant(cfg)
    assert list(acc._orders) == [1.5, 2.0, 2.5, 3.0, 4.0, 8.0, 16.0, 32.0, 64.0]
    e0 = acc.get_privacy_spent().epsilon_spent
    acc.add_noise_event(1.0, 10)
    e1 = acc.get_privacy_spent().epsilon_spent
    acc.add_noise_event(1.0, 10)
    e2 = acc.get_privacy_spent().epsilon_spent
    assert e0 < e1 < e2
    want = min(2 * a / 2 + math.log(1 / 1e-5) / (a - 1) for a in acc._orders)
    assert math.isclose(e2, want, rel_tol=1e-9)


def test_mechanisms_clip_noise_and_accounting():
    cfg = PrivacyConfig(max_gradient_norm=1.0, noise_multiplier=1.0)
    params = {"w": torch.full((10, 10), 3.0), "b": torch.full((10,), -2.0)}
    mech = CentralPrivacyMechanism(cfg, noise_generator=GaussianNoiseGenerator(seed=0))
    clipped, meta = mech._clip_update(params, 1.0)
    total = math.sqrt(sum(float(v.pow(2).sum()) for v in clipped.values()))
    assert total <= 1.0 + 1e-6 and meta.num_parameters == 110
    assert math.isclose(meta.total_norm, math.sqrt(900 + 40), rel_tol=1e-5)
    out = mech.add_noise(params, batch_size=4)
    assert out.keys() == params.keys() and mech.accountant._event_count == 1
    resid = torch.cat([(out[k] - clipped[k]).flatten() for k in out])
    assert abs(float(resid.std()) - 0.25) < 0.06               # σ·C / batch
    local = LocalPrivacyMechanism(cfg, noise_generator=GaussianNoiseGenerator(seed=0))
    resid = torch.cat([(local.add_noise(params, 99)[k] - clipped[k]).flatten() for k in params])
    assert abs(float(resid.std()) - 1.0) < 0.2                 # batch size ignored → σ·C
    assert PrivacyMechanismFactory.create(PrivacyType.LOCAL, cfg).privacy_type == PrivacyType.LOCAL
    with pytest.raises(ValueError, match="Unknown privacy type: invalid"):
        PrivacyMechanismFactory.create("invalid", cfg)
    assert math.isclose(mech.folded_noise_std([0.5, 0.5]), 0.5 * math.sqrt(0.5))
