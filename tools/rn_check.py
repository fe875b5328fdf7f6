"""Numerics check of the generic tcgen05 convolution kernels (csrc/rn_conv.cu) against PyTorch fp32
(``torch.backends.cudnn.allow_tf32 = False``): forward (+ BatchNorm batch sums), data gradient
(stride 1, stride 2 by parity classes, fused projection shortcut, residual + ReLU mask) and weight
gradient, for every layer geometry of the CIFAR ResNet-18.

    python tools/rn_check.py [--quick]

Every output sits inside a canary-filled buffer (compute-sanitizer is closed on the GPU pool).
"""

from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from nanofed_b200.ops import rn, tc  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
DEV = torch.device("cuda")
CANARY = 12345.0


def guarded(shape, dtype):
    """Tensor view in the middle of a canary-filled allocation."""
    n = 1
    for s in shape:
        n *= s
    pad = 4096
    raw = torch.full((n + 2 * pad,), CANARY, dtype=dtype, device=DEV)
    return raw, raw[pad : pad + n].view(shape)


def canaries_ok(raw, n):
    pad = 4096
    return bool((raw[:pad] == CANARY).all() and (raw[pad + n :] == CANARY).all())


def rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


def tap_major(w):  # [G, co, ci, kh, kw] -> [G, taps*co, ci]
    g, co, ci, kh, kw = w.shape
    return w.permute(0, 3, 4, 1, 2).reshape(g, kh * kw * co, ci).contiguous()


def check_layer(cin, cout, h, stride, k, clients=2, batch=16, tol=1.2e-2):
    ho = h // stride
    name = f"{k}x{k} s{stride} {cin}->{cout} @{h}x{h}"
    gen = torch.Generator(device="cpu").manual_seed(cin * 7 + cout + h + stride * 3 + k)
    x = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
    w = (torch.randn(clients, cout, cin, k, k, generator=gen) / (cin * k * k) ** 0.5).to(DEV).bfloat16()
    dy = torch.randn(clients * batch, ho, ho, cout, generator=gen).to(DEV).bfloat16()
    wt_rows = tap_major(w)
    wt_buf = torch.zeros(clients, wt_rows[0].numel() + 128, dtype=torch.bfloat16, device=DEV)
    wt_buf[:, 64 : 64 + wt_rows[0].numel()] = wt_rows.reshape(clients, -1)
    wref = rn.WeightRef(wt_buf, 64, k * k * cout, cin)
    ok = True
    # ---------------- references (fp32) ----------------
    xs = x.float().view(clients, batch, h, h, cin).permute(0, 1, 4, 2, 3)
    dys = dy.float().view(clients, batch, ho, ho, cout).permute(0, 1, 4, 2, 3)
    ref_y, ref_dx, ref_dw = [], [], []
    for g in range(clients):
        xi = xs[g].clone().requires_grad_(True)
        wi = w[g].float().clone().requires_grad_(True)
        y = F.conv2d(xi, wi, stride=stride, padding=k // 2)
        y.backward(dys[g])
        ref_y.append(y.detach().permute(0, 2, 3, 1))
        ref_dx.append(xi.grad.permute(0, 2, 3, 1))
        ref_dw.append(wi.grad)
    ref_y = torch.cat(ref_y).contiguous()
    ref_dx = torch.cat(ref_dx).contiguous()
    ref_dw = torch.stack(ref_dw)
    # ---------------- forward ----------------
    raw, y = guarded((clients * batch, ho, ho, cout), torch.bfloat16)
    stats = torch.zeros(clients, 2, cout, device=DEV)
    gstats = torch.zeros(clients * batch, 8, 2, device=DEV)      # GroupNorm(8) sums per image
    rn.conv_gemm(False, x, wref, y, clients=clients, grid_hw=(ho, ho),
                 taps=rn.fwd_taps(k, stride, cout), a_parity=stride == 2, stats=stats, n_out=cout,
                 gn_stats=gstats)
    torch.cuda.synchronize()
    e = rel(y, ref_y)
    yf = y.float().view(clients, -1, cout)
    es = rel(stats[:, 0], yf.sum(1))
    eq = rel(stats[:, 1], (yf * yf).sum(1))
    yg = y.float().view(clients * batch, ho * ho, 8, cout // 8)
    eg = max(rel(gstats[:, :, 0], yg.sum((1, 3))), rel(gstats[:, :, 1], (yg * yg).sum((1, 3))))
    good = e < tol and es < 1e-3 and eq < 1e-3 and eg < 1e-3 and canaries_ok(raw, y.numel())
    ok &= good
    print(f"  fwd   {name}: rel {e:.2e}  sum {es:.1e}  sumsq {eq:.1e}  group sums {eg:.1e}  "
          f"{'ok' if good else 'FAIL'}")
    # ---------------- dgrad ----------------
    raw, dx = guarded((clients * batch, h, h, cin), torch.bfloat16)
    if stride == 1:
        res = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
        msk = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
        om = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
        bx = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
        bsum = torch.zeros(clients, 2, cin, device=DEV)
        rn.conv_gemm(True, dy, wref, dx, clients=clients, grid_hw=(h, h),
                     taps=rn.dgrad_taps(k, 1, cout), residual=res, res_mask=msk, out_mask=om,
                     n_out=cin, stats=bsum, bn_x=bx)
        want = (ref_dx + res.float() * (msk.float() > 0)) * (om.float() > 0)
    else:
        dx.zero_()
        om = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
        for cls in ((0, 0), (0, 1), (1, 0), (1, 1)):
            taps = rn.dgrad_taps(k, 2, cout, cls)
            if not taps:
                continue     # 1x1 stride 2: only the even/even pixels receive a gradient
            rn.conv_gemm(True, dy, wref, dx, clients=clients, grid_hw=(ho, ho), taps=taps,
                         out_parity=cls, out_mask=om, n_out=cin)
        want = ref_dx * (om.float() > 0)
    torch.cuda.synchronize()
    e = rel(dx, want)
    good = e < tol and canaries_ok(raw, dx.numel())
    extra = ""
    if stride == 1:   # BatchNorm backward sums of the stored (rounded) gradient
        gf = dx.float().view(clients, -1, cin)
        e1 = rel(bsum[:, 0], gf.sum(1))
        e2 = rel(bsum[:, 1], (gf * bx.float().view(clients, -1, cin)).sum(1))
        good = good and e1 < 1e-3 and e2 < 1e-3
        extra = f"  sum g {e1:.1e}  sum gx {e2:.1e}"
    ok &= good
    print(f"  dgrad {name}: rel {e:.2e}{extra}  {'ok' if good else 'FAIL'}")
    # ---------------- wgrad ----------------
    for k_split, posk in ((1, 128), (3, 128), (1, 64)):
        n = k * k * cout * cin
        raw, dwt = guarded((clients, n + 128), torch.float32)
        dwt.zero_()
        rn.conv_wgrad(x, dy, dwt, 64, clients=clients, taps=rn.fwd_taps(k, stride, cout),
                      x_parity=stride == 2, k_split=k_split, tile=(min(cout, 256), posk))
        torch.cuda.synchronize()
        got = dwt[:, 64 : 64 + n].view(clients, k, k, cout, cin).permute(0, 3, 4, 1, 2)
        e = rel(got, ref_dw)
        untouched = bool((dwt[:, :64] == 0).all() and (dwt[:, 64 + n :] == 0).all())
        good = e < tol and canaries_ok(raw, dwt.numel()) and untouched
        ok &= good
        print(f"  wgrad {name} split {k_split} k-block {posk}: rel {e:.2e}  {'ok' if good else 'FAIL'}")
    tc.check_status(DEV)
    return ok


def check_fused_shortcut(cin=64, cout=128, h=32, clients=2, batch=16, tol=1.2e-2):
    """Stride-2 block input gradient: 3x3 stride-2 conv + 1x1 stride-2 projection in one k-loop."""
    ho = h // 2
    gen = torch.Generator(device="cpu").manual_seed(99)
    w3 = (torch.randn(clients, cout, cin, 3, 3, generator=gen) / (cin * 9) ** 0.5).to(DEV).bfloat16()
    w1 = (torch.randn(clients, cout, cin, 1, 1, generator=gen) / cin ** 0.5).to(DEV).bfloat16()
    dy3 = torch.randn(clients * batch, ho, ho, cout, generator=gen).to(DEV).bfloat16()
    dy1 = torch.randn(clients * batch, ho, ho, cout, generator=gen).to(DEV).bfloat16()
    n3, n1 = 9 * cout * cin, cout * cin
    buf = torch.zeros(clients, n3 + n1, dtype=torch.bfloat16, device=DEV)
    buf[:, :n3] = tap_major(w3).reshape(clients, -1)
    buf[:, n3:] = tap_major(w1).reshape(clients, -1)
    r3, r1 = rn.WeightRef(buf, 0, 9 * cout, cin), rn.WeightRef(buf, n3, cout, cin)
    raw, dx = guarded((clients * batch, h, h, cin), torch.bfloat16)
    # all four output parity classes in one launch (grid.z = class x client), with an output mask
    om = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
    bx = torch.randn(clients * batch, h, h, cin, generator=gen).to(DEV).bfloat16()
    bsum = torch.zeros(clients, 2, cin, device=DEV)
    rn.conv_gemm(True, dy3, r3, dx, clients=clients, grid_hw=(ho, ho),
                 taps=rn.dgrad_taps_all_classes(3, cout, cout), out_parity=(0, 0), all_classes=True,
                 a1=dy1, w1=r1, out_mask=om, n_out=cin, stats=bsum, bn_x=bx)
    torch.cuda.synchronize()
    want = []
    for g in range(clients):
        xi = torch.zeros(batch, cin, h, h, device=DEV, requires_grad=True)
        y3 = F.conv2d(xi, w3[g].float(), stride=2, padding=1)
        y1 = F.conv2d(xi, w1[g].float(), stride=2)
        d3 = dy3.float().view(clients, batch, ho, ho, cout)[g].permute(0, 3, 1, 2)
        d1 = dy1.float().view(clients, batch, ho, ho, cout)[g].permute(0, 3, 1, 2)
        (y3 * d3).sum().add((y1 * d1).sum()).backward()
        want.append(xi.grad.permute(0, 2, 3, 1))
    e = rel(dx, torch.cat(want) * (om.float() > 0))
    gf = dx.float().view(clients, -1, cin)
    e1 = rel(bsum[:, 0], gf.sum(1))
    e2 = rel(bsum[:, 1], (gf * bx.float().view(clients, -1, cin)).sum(1))
    good = e < tol and e1 < 1e-3 and e2 < 1e-3 and canaries_ok(raw, dx.numel())
    print(f"  dgrad 3x3 s2 + 1x1 s2 shortcut {cin}->{cout}: rel {e:.2e}  sum g {e1:.1e}  "
          f"sum gx {e2:.1e}  {'ok' if good else 'FAIL'}")
    tc.check_status(DEV)
    return good


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    layers = [(64, 64, 32, 1, 3), (64, 128, 32, 2, 3), (64, 128, 32, 2, 1), (128, 128, 16, 1, 3),
              (128, 256, 16, 2, 3), (256, 256, 8, 1, 3), (256, 512, 8, 2, 3), (256, 512, 8, 2, 1),
              (512, 512, 4, 1, 3)]
    if args.quick:
        layers = layers[:2]
    ok = True
    for spec in layers:
        try:
            ok &= check_layer(*spec)
        except Exception as exc:  # keep going: one broken geometry should not hide the others
            ok = False
            print(f"  {spec}: EXCEPTION {exc!r}")
    try:
        ok &= check_fused_shortcut()
    except Exception as exc:
        ok = False
        print(f"  fused shortcut: EXCEPTION {exc!r}")
    print("ALL OK" if ok else "FAILURES")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
