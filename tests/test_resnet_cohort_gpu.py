"""ResNet-18 cohort engine (csrc/rn_conv.cu, csrc/rn_elt.cu, models/resnet_cohort.py) against
PyTorch fp32 autograd with TF32 off."""

import sys
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))


@pytest.fixture(autouse=True)
def _no_tf32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("spec", [(64, 64, 32, 1, 3), (64, 128, 32, 2, 3), (64, 128, 32, 2, 1),
                                  (128, 128, 16, 1, 3), (256, 256, 8, 1, 3), (256, 512, 8, 2, 3),
                                  (512, 512, 4, 1, 3)])
def test_conv_kernels_match_fp32(spec):
    """Forward (+ BatchNorm batch sums), dgrad (residual + mask, parity classes) and wgrad (plain
    and split-K) of one layer geometry; bf16 outputs within 1.2e-2, fp32 outputs (wgrad) 1e-3."""
    import rn_check

    assert rn_check.check_layer(*spec)


def test_fused_projection_shortcut_dgrad():
    import rn_check

    assert rn_check.check_fused_shortcut()


def test_step_matches_fp32_autograd_as_closely_as_bf16_allows():
    """Loss within 0.5 %, BatchNorm running statistics within 1 %, parameter updates as close to
    fp32 as PyTorch's own bf16 autocast run of the same step (mean over parameters +15 %, worst
    parameter +50 %)."""
    import rn_step_check

    w = rn_step_check.run(clients=2, batch=16, steps=1, verbose=False)
    assert w["loss_err"] < 5e-3, w
    assert w["stat_rel"] < 1e-2, w
    assert rn_step_check.acceptable(w), w


def test_bn_and_head_kernels_match_torch():
    """The elementwise kernels alone, fp32 oracle: BatchNorm forward (+ residual, ReLU, running
    statistics), backward (reduce + apply) and the pool/fc/cross-entropy head."""
    from nanofed_b200.ops import rn

    dev = torch.device("cuda")
    g, b, h, c = 2, 8, 8, 128
    gen = torch.Generator(device="cpu").manual_seed(5)
    x = (torch.randn(g * b, h, h, c, generator=gen) * 1.5 + 0.3).to(dev).bfloat16()
    res = torch.randn(g * b, h, h, c, generator=gen).to(dev).bfloat16()
    dout = torch.randn(g * b, h, h, c, generator=gen).to(dev).bfloat16()
    n_row = 5 * c + 64
    slots = torch.zeros(g, n_row, device=dev)
    offs = (0, c, 2 * c, 3 * c, 4 * c)
    slots[:, 0:c] = torch.rand(g, c, generator=gen).to(dev) + 0.5       # gamma
    slots[:, c : 2 * c] = torch.randn(g, c, generator=gen).to(dev)     # beta
    slots[:, 3 * c : 4 * c] = 1.0                                        # running_var
    xf = x.float().view(g, -1, c)
    stats = torch.stack([xf.sum(1), (xf * xf).sum(1)], dim=1).contiguous()
    y = torch.empty_like(x)
    rn.bn_apply(x, y, slots, stats, offs, clients=g, relu=True, res=res)
    sgrad = torch.zeros(g, 2 * c, device=dev)
    rn.bn_bwd(False, dout, y, x, slots, sgrad, (0, c, 0, 0), stats, offs, clients=g)
    dx = torch.empty_like(x)
    rn.bn_bwd(True, dout, y, x, slots, sgrad, (0, c, 0, 0), stats, offs, clients=g, dx=dx)
    torch.cuda.synchronize()
    for k in range(g):
        xi = x.float().view(g, b, h, h, c)[k].permute(0, 3, 1, 2).clone().requires_grad_(True)
        gamma = slots[k, 0:c].clone().requires_grad_(True)
        beta = slots[k, c : 2 * c].clone().requires_grad_(True)
        rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
        out = F.relu(F.batch_norm(xi, rm, rv, gamma, beta, training=True, momentum=0.1, eps=1e-5)
                     + res.float().view(g, b, h, h, c)[k].permute(0, 3, 1, 2))
        out.backward(dout.float().view(g, b, h, h, c)[k].permute(0, 3, 1, 2))
        got_y = y.float().view(g, b, h, h, c)[k].permute(0, 3, 1, 2)
        got_dx = dx.float().view(g, b, h, h, c)[k].permute(0, 3, 1, 2)
        assert (got_y - out).abs().max() <= 1e-2 * out.abs().max()          # bf16 output
        assert (got_dx - xi.grad).norm() <= 1e-2 * xi.grad.norm()           # bf16 output
        # fp32 outputs; the ReLU mask is taken from the bf16-rounded output, so a handful of
        # elements within rounding distance of zero may differ from the fp32 oracle
        assert (sgrad[k, 0:c] - gamma.grad).norm() <= 5e-3 * gamma.grad.norm()
        assert (sgrad[k, c:] - beta.grad).norm() <= 5e-3 * beta.grad.norm()
        torch.testing.assert_close(slots[k, 2 * c : 3 * c], rm, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(slots[k, 3 * c : 4 * c], rv, rtol=1e-3, atol=1e-4)
        assert float(slots[k, 4 * c]) == 1.0
    # ---- head ----
    feat = torch.randn(g * b, 4, 4, 512, generator=gen).to(dev).bfloat16()
    labels = torch.randint(0, 10, (g, b), generator=gen).to(dev)
    hs = torch.zeros(g, 10 * 512 + 64, device=dev)
    hs[:, : 10 * 512] = (torch.randn(g, 10 * 512, generator=gen) * 0.05).to(dev)
    hs[:, 10 * 512 : 10 * 512 + 10] = torch.randn(g, 10, generator=gen).to(dev)
    hg = torch.zeros(g, 10 * 512 + 64, device=dev)
    dfeat = torch.empty_like(feat)
    st = torch.zeros(g, 3, device=dev)
    rn.head(feat, labels, hs, 0, 10 * 512, hg, 0, 10 * 512, dfeat, st, clients=g)
    torch.cuda.synchronize()
    for k in range(g):
        f = feat.float().view(g, b, 16, 512)[k].clone().requires_grad_(True)
        wgt = hs[k, : 10 * 512].view(10, 512).clone().requires_grad_(True)
        bias = hs[k, 10 * 512 : 10 * 512 + 10].clone().requires_grad_(True)
        logits = f.mean(1) @ wgt.t() + bias
        loss = F.cross_entropy(logits, labels[k])
        loss.backward()
        torch.testing.assert_close(st[k, 0] / b, loss.detach(), rtol=1e-4, atol=1e-5)
        assert float(st[k, 1]) == float((logits.argmax(1) == labels[k]).sum())
        torch.testing.assert_close(hg[k, : 10 * 512].view(10, 512), wgt.grad, rtol=1e-3, atol=1e-6)
        torch.testing.assert_close(hg[k, 10 * 512 : 10 * 512 + 10], bias.grad, rtol=1e-3, atol=1e-6)
        # the head hands back the gradient w.r.t. the PRE-ReLU feature map: masked where feat <= 0
        want_df = f.grad * (feat.float().view(g, b, 16, 512)[k] > 0)
        assert (dfeat.float().view(g, b, 16, 512)[k] - want_df).norm() <= 1e-2 * want_df.norm()


def _sim(**kw):
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    cfg = CohortConfig(model="resnet18", clients_per_gpu=2, samples_per_client=32, batch_size=16,
                       learning_rate=0.01, device="cuda", **kw)
    return FedSimulation(cfg)


def test_round_uses_the_grouped_engine_and_graph_replay_equals_eager():
    from nanofed_b200.ops import tc

    eager = _sim(use_cuda_graph=False)
    graph = _sim(use_cuda_graph=True)
    assert eager.engine_name == graph.engine_name == "grouped"
    assert type(eager.engine).__name__ == "ResNetCohortNet"
    outs = []
    m0 = eager.arena.master.clone()
    for sim in (eager, graph):
        for _ in range(2):
            out = sim.run_round()
        outs.append((out, sim.arena.master.clone()))
        tc.check_status(sim.device)
    (oe, me), (og, mg) = outs
    assert torch.isfinite(me).all() and torch.isfinite(mg).all()
    # fp32 atomics reorder sums between runs and bf16 roundings tip over, which a batch-16
    # BatchNorm network amplifies: compare the two runs' UPDATES in relative L2
    moved = float((me - m0).norm())
    assert moved > 0 and float((me - mg).norm()) < 0.25 * moved, (float((me - mg).norm()), moved)
    assert abs(oe.agg_metrics["loss"] - og.agg_metrics["loss"]) < 2e-2 * abs(oe.agg_metrics["loss"])
    # the state_dict round-trips through the reference's layout, BatchNorm counters included
    state = graph.global_state()
    assert int(state["bn1.num_batches_tracked"]) == 2 * 2      # 2 rounds x 2 local steps, averaged
    from nanofed_b200.models import ResNet18

    ResNet18().load_state_dict(state)


def test_momentum_and_weight_decay_match_torch_sgd():
    """The unfused optimizer path (tap_sgd + small_sgd with momentum buffers): two steps against
    torch.optim.SGD(momentum=0.9, weight_decay=5e-4) on the fp32 model; BatchNorm buffers must
    not be touched by weight decay."""
    from nanofed_b200 import ops
    from nanofed_b200.models import ResNet18
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    cfg = CohortConfig(model="resnet18", clients_per_gpu=1, samples_per_client=32, batch_size=16,
                       learning_rate=0.05, momentum=0.9, weight_decay=5e-4, use_cuda_graph=False,
                       device="cuda")
    sim = FedSimulation(cfg)
    a = sim.arena
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    sim.engine.after_broadcast()
    before = sim.layout.unflatten(a.slots[0].clone())
    for s in range(2):
        sim.engine.step(s)
    sim.engine.before_aggregate()
    torch.cuda.synchronize()
    got = sim.layout.unflatten(a.slots[0].clone())
    model = ResNet18().cuda().train()
    model.load_state_dict(before)
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
    for s in range(2):
        x, y = sim.shards.images[0, s * 16 : (s + 1) * 16], sim.shards.labels[0, s * 16 : (s + 1) * 16]
        opt.zero_grad()
        F.cross_entropy(model(x), y).backward()
        opt.step()
    want = model.state_dict()
    for k in ("layer4.1.conv2.weight", "layer1.0.conv1.weight", "fc.weight", "layer3.0.bn1.weight"):
        d_got = got[k].float() - before[k].float().cuda()
        d_want = want[k].float() - before[k].float().cuda()
        cos = float(F.cosine_similarity(d_got.flatten(), d_want.flatten(), dim=0))
        ratio = float(d_got.norm() / d_want.norm())
        # direction: bf16 activations (PyTorch's own autocast sits at the same distance from fp32);
        # length: the optimizer arithmetic (momentum, weight decay, learning rate)
        assert cos > 0.7 and 0.9 < ratio < 1.1, (k, cos, ratio)
    for k in ("bn1.running_var", "layer2.0.shortcut.1.running_mean"):
        torch.testing.assert_close(got[k].float(), want[k].float(), rtol=2e-2, atol=2e-3)


def test_bn_sums_from_the_dgrad_epilogue_give_the_bn_gradients(monkeypatch):
    """Every BatchNorm backward of one step, checked in place: the gradient rows (dgamma, dbeta)
    the apply kernel leaves in ``sgrad`` - derived from the sums of the dgrad epilogue where the
    engine fuses them - against fp32 torch on the very tensors the kernel read.  (Two separate runs
    cannot be compared this tightly: fp32 atomics reorder the forward statistics and a batch-16
    network in bf16 amplifies that to per cents within a few layers.)"""
    import nanofed_b200.models.resnet_cohort as rc
    from nanofed_b200 import ops
    from nanofed_b200.ops import rn

    seen = []
    real = rn.bn_bwd

    def spy(apply, dout, out, x, slots, sgrad, grad_offs, stats, bn_offs, **kw):
        if apply:
            torch.cuda.synchronize()
            g, c = stats.shape[0], stats.shape[2]
            d, xx = dout.float().reshape(g, -1, c), x.float().reshape(g, -1, c)
            n = d.shape[1]
            mean = stats[:, 0] / n
            rstd = (stats[:, 1] / n - mean * mean + 1e-5).rsqrt()
            want_g = (d * (xx - mean[:, None]) * rstd[:, None]).sum(1)
            want_b = d.sum(1)
        real(apply, dout, out, x, slots, sgrad, grad_offs, stats, bn_offs, **kw)
        if apply:
            torch.cuda.synchronize()
            got_g = sgrad[:, grad_offs[0] : grad_offs[0] + c]
            got_b = sgrad[:, grad_offs[1] : grad_offs[1] + c]
            seen.append((kw.get("raw_sums") is not None,
                         float((got_g - want_g).norm() / want_g.norm()),
                         float((got_b - want_b).norm() / want_b.norm())))

    monkeypatch.setattr(rc.rn, "bn_bwd", spy)
    sim = _sim(use_cuda_graph=False)
    assert sim.engine.fuse_bn_sums
    a = sim.arena
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    sim.engine.after_broadcast()
    sim.engine.step(0)
    torch.cuda.synchronize()
    lanes = len(sim.engine.lanes)
    assert len(seen) == 17 * lanes               # stem + 8 blocks x (bn1, bn2 [+ shortcut])
    assert sum(fused for fused, _, _ in seen) == 13 * lanes, seen   # all but the last bn2 and 3 projections
    for fused, eg, eb in seen:
        assert eg < 1e-5 and eb < 1e-5, seen


@pytest.mark.parametrize("c,hw,two", [(64, 32, False), (128, 16, True), (512, 4, True)])
def test_group_norm_kernels_match_torch(c, hw, two):
    """csrc/rn_gn.cu against fp32 torch (F.group_norm + autograd): statistics, forward with the
    residual variants, backward (parameter gradients, dx; with ``two`` also the projection
    shortcut's GroupNorm fed by the same gradient)."""
    from nanofed_b200.ops import rn

    dev = torch.device("cuda")
    gen = torch.Generator(device="cpu").manual_seed(c + hw)
    clients, batch = 2, 8
    imgs = clients * batch
    x = (torch.randn(imgs, hw, hw, c, generator=gen) * 1.5 + 0.3).to(dev).bfloat16()
    x2 = torch.randn(imgs, hw, hw, c, generator=gen).to(dev).bfloat16()
    dout = torch.randn(imgs, hw, hw, c, generator=gen).to(dev).bfloat16()
    row = 4 * c + 8
    slots = torch.randn(clients, row, generator=gen).to(dev)          # gamma, beta, gamma2, beta2
    offs, offs2 = (0, c), (2 * c, 3 * c)
    stats = torch.zeros(imgs, 8, 2, device=dev)
    stats2 = torch.zeros(imgs, 8, 2, device=dev)
    rn.gn_stats(x, stats)
    rn.gn_stats(x2, stats2)
    xg = x.float().view(imgs, hw * hw, 8, c // 8)
    torch.testing.assert_close(stats[:, :, 0], xg.sum((1, 3)), rtol=1e-4, atol=1e-2)
    torch.testing.assert_close(stats[:, :, 1], (xg * xg).sum((1, 3)), rtol=1e-4, atol=1e-2)

    def ref_forward(xf, x2f, gam, bet, gam2, bet2):
        outs = []
        for k in range(clients):
            sl = slice(k * batch, (k + 1) * batch)
            y = F.group_norm(xf[sl].permute(0, 3, 1, 2), 8, gam[k], bet[k], eps=1e-5)
            if two:
                y = y + F.group_norm(x2f[sl].permute(0, 3, 1, 2), 8, gam2[k], bet2[k], eps=1e-5)
            else:
                y = y + x2f[sl].permute(0, 3, 1, 2)
            outs.append(y.permute(0, 2, 3, 1))
        return torch.cat(outs)

    leaves = [t.clone().requires_grad_(True) for t in
              (x.float(), x2.float(), slots[:, :c], slots[:, c : 2 * c], slots[:, 2 * c : 3 * c],
               slots[:, 3 * c : 4 * c])]
    pre = ref_forward(*leaves)
    y = torch.empty_like(x)
    rn.gn_apply(x, y, slots, stats, offs, clients=clients, relu=True, res=x2,
                res_stats=stats2 if two else None, res_offs=offs2 if two else None)
    torch.cuda.synchronize()
    want = torch.relu(pre)
    assert float((y.float() - want).abs().max()) <= 2e-2 * float(want.abs().max()) + 1e-2
    # backward: dout is the gradient w.r.t. the pre-ReLU sum
    pre.backward(dout.float())
    sgrad = torch.zeros(clients, 4 * c, device=dev)
    bs, bs2 = torch.zeros(imgs, 8, 2, device=dev), torch.zeros(imgs, 8, 2, device=dev)
    dx, dx2 = torch.empty_like(x), torch.empty_like(x)
    kw = dict(clients=clients)
    if two:
        kw.update(x2=x2, stats2=stats2, gn2_offs=offs2, bsums2=bs2)
    args = (dout, x, slots, sgrad, (0, c, 2 * c, 3 * c), stats, offs, bs)
    rn.gn_bwd(False, *args, **kw)
    rn.gn_bwd(True, *args, dx=dx, dx2=dx2 if two else None, **kw)
    torch.cuda.synchronize()

    def close(got, ref, tol):
        assert float((got.float() - ref).norm()) <= tol * float(ref.norm()) + 1e-6, (
            float((got.float() - ref).norm()), float(ref.norm()))

    close(sgrad[:, :c], leaves[2].grad, 2e-3)          # dgamma
    close(sgrad[:, c : 2 * c], leaves[3].grad, 2e-3)   # dbeta
    close(dx, leaves[0].grad, 1e-2)                    # bf16 output
    if two:
        close(sgrad[:, 2 * c : 3 * c], leaves[4].grad, 2e-3)
        close(sgrad[:, 3 * c :], leaves[5].grad, 2e-3)
        close(dx2, leaves[1].grad, 1e-2)


def test_group_norm_step_matches_fp32_autograd():
    """resnet18_gn on the grouped engine: one step against fp32 autograd, judged like the
    BatchNorm network (as close to fp32 as PyTorch's own bf16 autocast run)."""
    import rn_step_check

    w = rn_step_check.run(clients=2, batch=16, steps=1, verbose=False, model="resnet18_gn")
    assert w["loss_err"] < 5e-3, w
    assert rn_step_check.acceptable(w), w


def test_group_norm_model_trains_on_the_grouped_engine():
    from nanofed_b200.models import ResNet18
    from nanofed_b200.ops import tc
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation

    sim = FedSimulation(CohortConfig(model="resnet18_gn", clients_per_gpu=2, samples_per_client=32,
                                     batch_size=16, learning_rate=0.02, device="cuda"))
    assert sim.engine_name == "grouped" and sim.engine.lanes[0].norm == "group"
    losses = [sim.run_round().agg_metrics["loss"] for _ in range(4)]
    tc.check_status(sim.device)
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses
    ResNet18(norm="group").load_state_dict(sim.global_state())


def _dp_sim(sigma, clip, **kw):
    from nanofed_b200.parallel.cohort import CohortConfig, FedSimulation
    from nanofed_b200.privacy import PrivacyConfig

    cfg = CohortConfig(model="resnet18_gn", clients_per_gpu=2, samples_per_client=32, batch_size=16,
                       learning_rate=0.1, device="cuda", dp_seed=1234,
                       local_privacy=PrivacyConfig(epsilon=8.0, delta=1e-5, max_gradient_norm=clip,
                                                   noise_multiplier=sigma), **kw)
    return FedSimulation(cfg)


def _param_delta(sim, before, k):
    """Update of client k's trainable parameters, flattened (buffers excluded)."""
    after = sim.layout.unflatten(sim.arena.slots[k].clone())
    names = [n for n, _ in sim.template.named_parameters()]
    return {n: after[n].float() - before[n].float().cuda() for n in names}


def test_client_side_dp_sgd_in_the_resnet_engine_clips_and_adds_noise():
    """``local_privacy`` on the grouped ResNet engine (reference recipe
    ``nanofed/trainer/private.py:54-86``): with sigma ~ 0 one step moves the parameters by exactly
    lr * C in total norm (the batch gradient is far larger than C); with sigma > 0 every parameter
    receives N(0, (sigma C)^2) noise, different per client and per step."""
    from nanofed_b200 import ops

    lr, clip = 0.1, 0.05
    sim = _dp_sim(1e-6, clip, use_cuda_graph=False)   # sigma must be > 0; 1e-6 adds nothing visible
    assert sim.engine_name == "grouped" and sim.engine.lanes[0].local_dp is not None
    a = sim.arena
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    sim.engine.after_broadcast()
    before = sim.layout.unflatten(a.slots[0].clone())
    sim.engine.step(0)
    sim.engine.before_aggregate()
    torch.cuda.synchronize()
    for k in range(2):
        d = _param_delta(sim, before, k)
        total = float(torch.sqrt(sum((v * v).sum() for v in d.values())))
        assert abs(total - lr * clip) < 2e-3 * lr * clip, (k, total, lr * clip)

    sigma = 2.0
    sim = _dp_sim(sigma, clip, use_cuda_graph=False)
    a = sim.arena
    ops.bcast_global(a.master, a.slots, a.slots_lowp, a.momentum)
    sim.engine.after_broadcast()
    before = sim.layout.unflatten(a.slots[0].clone())
    sim.engine.step(0)
    sim.engine.before_aggregate()
    torch.cuda.synchronize()
    d0, d1 = _param_delta(sim, before, 0), _param_delta(sim, before, 1)
    for name in ("layer4.1.conv2.weight", "layer1.0.conv1.weight", "fc.weight", "bn1.weight"):
        z0, z1 = d0[name].flatten() / (lr * sigma * clip), d1[name].flatten() / (lr * sigma * clip)
        n = z0.numel()
        tol = 5.0 / n ** 0.5
        assert abs(float(z0.std()) - 1.0) < max(tol, 0.01), (name, float(z0.std()))
        assert abs(float(z0.mean())) < tol + 0.02, (name, float(z0.mean()))
        # different clients draw different noise
        corr = float((z0 * z1).mean())
        assert abs(corr) < tol + 0.02, (name, corr)


def test_client_side_dp_sgd_noise_advances_inside_a_replayed_graph():
    """Two rounds through the CUDA graph: the second replay must not repeat the first one's noise
    (the step counter lives on the device and moves once per round)."""
    sim = _dp_sim(1.0, 0.05)
    m0 = sim.arena.master.clone()
    sim.run_round()
    m1 = sim.arena.master.clone()
    sim.arena.master.copy_(m0)          # same starting model: only the noise stream differs
    sim._slots_fresh = False
    sim.run_round()
    m2 = sim.arena.master.clone()
    moved = float((m1 - m0).norm())
    assert moved > 0 and float((m2 - m1).norm()) > 0.5 * moved, (float((m2 - m1).norm()), moved)
