"""Numerics check of every tcgen05 GEMM/conv variant against plain PyTorch fp32.

    python tools/tc_check.py            # each variant in its own subprocess (60 s timeout)
    python tools/tc_check.py conv_fwd   # one variant in-process
    python tools/tc_check.py --inproc conv_pool fc_dgrad   # several in one process (sanitizer runs)

Used while bringing the kernels up (a bad descriptor can fault the context, so variants
are isolated); ``tests/test_tc_gpu.py`` runs the same functions under pytest.
"""
import subprocess
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

# the oracles are fp32: no TF32 in the library convolutions / matmuls they call
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

W = 26  # grid width

# kernels whose OUTPUT is fp32 (fp32 accumulation of exact bf16 products: only the summation order
# differs from the oracle) are held to 1e-3 of the output range, bf16 outputs to 1e-2
FP32_OUT = {"fc_fwd", "conv_wgrad", "conv_wgrad_p", "fc_wgrad", "fc_wgrad_sgd"}


def tolerance(name: str) -> float:
    return 1e-3 if name in FP32_OUT else 1e-2


def _bf(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).bfloat16()


_GUARDS = []


def _guarded(*shape, dtype=torch.bfloat16, pad=4096):
    """Zeroed tensor of ``shape`` inside a larger canary-filled buffer; ``_check_guards()``
    verifies afterwards that nothing outside the tensor was written (compute-sanitizer is not
    available on the GPU pool, so out-of-bounds stores are caught this way)."""
    n = 1
    for d in shape:
        n *= d
    canary = 77 if dtype == torch.uint8 else 1234.0
    buf = torch.full((n + 2 * pad,), canary, dtype=dtype, device="cuda")
    view = buf[pad:pad + n].view(*shape)
    view.zero_()
    _GUARDS.append((buf, pad, n, canary))
    return view


def _check_guards():
    torch.cuda.synchronize()
    for buf, pad, n, canary in _GUARDS:
        assert bool((buf[:pad] == canary).all()) and bool((buf[pad + n:] == canary).all()), \
            "a kernel wrote outside its output tensor"
    _GUARDS.clear()


def _shift_rows(x, s):
    """y[:, q] = x[:, q + s] with zero fill outside (per client)."""
    c, r, ch = x.shape
    y = torch.zeros_like(x)
    if s >= 0:
        y[:, : r - s] = x[:, s:]
    else:
        y[:, -s:] = x[:, : r + s]
    return y


def check_conv_fwd():
    from nanofed_b200.ops import tc
    c, b, cin, cout = 3, 2, 32, 64
    r = b * W * W
    x, w = _bf(c, r, cin), _bf(c, cout, 9 * cin, scale=0.05)
    bias = torch.randn(c, cout, device="cuda")
    out = torch.zeros(c, r, cout, device="cuda", dtype=torch.bfloat16)
    tc.conv3x3_fwd(x, w, bias, out, W, relu=True)
    ref = torch.zeros(c, r, cout, device="cuda")
    for t, s in enumerate(tc._shifts(W)):
        ref += torch.einsum("crk,cok->cro", _shift_rows(x, s).float(),
                            w[:, :, t * cin:(t + 1) * cin].float())
    ref = (ref + bias[:, None, :]).relu()
    return out.float(), ref


def check_conv_dgrad():
    from nanofed_b200.ops import tc
    c, b, cin, cout = 2, 2, 32, 64
    r = b * W * W
    dy, w = _bf(c, r, cout), _bf(c, cout, 9 * cin, scale=0.05)
    a1 = _bf(c, r, cin)
    out = torch.zeros(c, r, cin, device="cuda", dtype=torch.bfloat16)
    tc.conv3x3_dgrad(dy, w, out, W, relu_mask=a1)
    ref = torch.zeros(c, r, cin, device="cuda")
    for t, s in enumerate(tc._shifts(W)):
        ref += torch.einsum("cro,cok->crk", _shift_rows(dy, -s).float(),
                            w[:, :, t * cin:(t + 1) * cin].float())
    ref = ref * (a1.float() > 0)
    return out.float(), ref


def check_conv_fwd_p():
    from nanofed_b200.ops import tc
    c, b, cin, cout = 3, 5, 32, 64
    r = b * W * W
    x, w = _bf(c, r, cin), _bf(c, cout, 9 * cin, scale=0.05)
    bias = torch.randn(c, cout, device="cuda")
    out = _guarded(c, r, cout)
    tc.conv3x3_fwd_persistent(x, w, bias, out, W)
    ref = torch.zeros(c, r, cout, device="cuda")
    for t, s in enumerate(tc._shifts(W)):
        ref += torch.einsum("crk,cok->cro", _shift_rows(x, s).float(),
                            w[:, :, t * cin:(t + 1) * cin].float())
    return out.float(), (ref + bias[:, None, :]).relu()


def check_conv_pool():
    """Fused conv+bias+ReLU+maxpool (no dropout) vs F.conv2d + F.max_pool2d on NCHW."""
    import torch.nn.functional as F
    from nanofed_b200.ops import tc
    c, b = 3, 5
    r = b * W * W
    x = _bf(c, r, 32)
    w = _bf(c, 64, 288, scale=0.05)
    bias = torch.randn(c, 64, device="cuda")
    p = _guarded(c, b, 9216)
    idx = _guarded(c, b, 144, 64, dtype=torch.uint8)
    tc.conv3x3_pool_fwd(x, w, bias, p, idx, b, W, 0.0, None, 0)
    outs = []
    for k in range(c):
        xin = x[k].float().view(b, W, W, 32).permute(0, 3, 1, 2)            # NCHW
        wk = w[k].float().view(64, 9, 32).permute(0, 2, 1).reshape(64, 32, 3, 3)
        y = F.relu(F.conv2d(xin, wk, bias[k]))
        y = y.bfloat16().float()        # the kernel pools bf16-rounded activations
        outs.append(F.max_pool2d(y, 2).flatten(1))
    ref = torch.stack(outs)
    # argmax codes must point at a maximal element and carry the ReLU flag
    assert int((idx & 8).min()) == 8, "keep bit must be set without dropout"
    return p.float(), ref


def check_conv_dgrad_p():
    from nanofed_b200.ops import tc
    c, b, cin, cout = 2, 5, 32, 64
    r = b * W * W
    dy, w = _bf(c, r, cout), _bf(c, cout, 9 * cin, scale=0.05)
    a1 = _bf(c, r, cin)
    out = _guarded(c, r, cin)
    tc.conv3x3_dgrad_persistent(dy, w, out, W, relu_mask=a1)
    ref = torch.zeros(c, r, cin, device="cuda")
    for t, s in enumerate(tc._shifts(W)):
        ref += torch.einsum("cro,cok->crk", _shift_rows(dy, -s).float(),
                            w[:, :, t * cin:(t + 1) * cin].float())
    return out.float(), ref * (a1.float() > 0)


def check_conv_wgrad():
    from nanofed_b200.ops import tc
    c, b, cin, cout = 2, 3, 32, 64
    r = b * W * W
    dy, x = _bf(c, r, cout, scale=0.1), _bf(c, r, cin)
    n_total = 64 * 32 * 9 + 128
    grad = torch.zeros(c, n_total, device="cuda")
    tc.conv3x3_wgrad(dy, x, grad[:, 64:], W, n_total, k_split=4)
    got = grad[:, 64:64 + 64 * 32 * 9].reshape(c, cout, cin, 9)
    ref = torch.stack([torch.einsum("cro,crk->cok", dy.float(), _shift_rows(x, s).float())
                       for s in tc._shifts(W)], dim=-1)
    return got, ref


def check_conv_wgrad_p():
    from nanofed_b200.ops import tc
    c, b, cin, cout = 3, 7, 32, 64
    r = b * W * W
    dy, x = _bf(c, r, cout, scale=0.1), _bf(c, r, cin)
    n_total = 64 * 32 * 9 + 128
    grad = torch.zeros(c, n_total, device="cuda")
    tc.conv3x3_wgrad_persistent(dy, x, grad[:, 64:], W, n_total)
    got = grad[:, 64:64 + 64 * 32 * 9].reshape(c, cout, cin, 9)
    ref = torch.stack([torch.einsum("cro,crk->cok", dy.float(), _shift_rows(x, s).float())
                       for s in tc._shifts(W)], dim=-1)
    return got, ref


def check_fc_fwd():
    from nanofed_b200.ops import tc
    c, b, j, i = 3, 64, 128, 9216
    w, x = _bf(c, j, i, scale=0.02), _bf(c, b, i)
    out = torch.zeros(c, b, j, device="cuda")
    tc.linear_fwd(w, x, out, k_split=8)
    return out, torch.einsum("cbi,cji->cbj", x.float(), w.float())


def check_fc_dgrad():
    from nanofed_b200.ops import tc
    c, b, j, i = 2, 64, 128, 9216
    w, dy = _bf(c, j, i, scale=0.02), _bf(c, b, j)
    out = _guarded(c, b, i)
    tc.linear_dgrad(w, dy, out)
    return out.float(), torch.einsum("cbj,cji->cbi", dy.float(), w.float())


def check_fc_dgrad_ragged():
    """Persistent fc dgrad with a batch that is not a multiple of 32 and strided weights."""
    from nanofed_b200.ops import tc
    c, b, j, i = 5, 100, 128, 1152
    wbuf = _bf(c, j * i + 512, scale=0.02)
    w = wbuf[:, 256:256 + j * i].unflatten(1, (j, i))
    dy = _bf(c, b, j)
    out = _guarded(c, b, i)
    tc.linear_dgrad(w, dy, out)
    return out.float(), torch.einsum("cbj,cji->cbi", dy.float(), w.float())


def check_fc_wgrad():
    from nanofed_b200.ops import tc
    c, b, j, i = 2, 64, 128, 9216
    dy, x = _bf(c, b, j, scale=0.1), _bf(c, b, i)
    n_total = j * i + 256
    grad = torch.zeros(c, n_total, device="cuda")
    tc.linear_wgrad(dy, x, grad[:, 128:], n_total)
    got = grad[:, 128:128 + j * i].reshape(c, j, i)
    return got, torch.einsum("cbj,cbi->cji", dy.float(), x.float())


def check_fc_wgrad_sgd():
    """Weight gradient consumed by the SGD epilogue: compare the updated fp32 weights."""
    from nanofed_b200.ops import tc
    c, b, j, i = 3, 64, 128, 9216
    lr, wd = 0.5, 0.01
    dy, x = _bf(c, b, j, scale=0.1), _bf(c, b, i)
    n_total = j * i + 256                      # weights live inside longer arena rows
    w = torch.randn(c, n_total, device="cuda")
    lowp = torch.zeros(c, n_total, device="cuda", dtype=torch.bfloat16)
    w0 = w.clone()
    wv = w[:, 128:128 + j * i].unflatten(1, (j, i))
    lv = lowp[:, 128:128 + j * i].unflatten(1, (j, i))
    ssq = torch.full((c,), 3.0, device="cuda")           # accumulated, not overwritten
    tc.linear_wgrad_sgd(dy, x, wv, lv, lr, wd, sumsq_out=ssq)
    torch.cuda.synchronize()
    want_ssq = 3.0 + (wv.double() ** 2).sum((1, 2))
    assert float(((ssq.double() - want_ssq) / want_ssq).abs().max()) < 1e-5, "sum-of-squares side output"
    old = w0[:, 128:128 + j * i].unflatten(1, (j, i))
    grad = torch.einsum("cbj,cbi->cji", dy.float(), x.float())
    want = old - lr * (grad + wd * old)
    # nothing outside the weight block may change, and the bf16 copy is the rounded result
    assert torch.equal(w[:, :128], w0[:, :128]) and torch.equal(w[:, 128 + j * i:], w0[:, 128 + j * i:])
    assert torch.equal(lv, wv.bfloat16()), "bf16 working copy is not the rounded fp32 weights"
    assert float(lowp[:, :128].abs().sum()) == 0 and float(lowp[:, 128 + j * i:].abs().sum()) == 0
    return wv.clone(), want


def check_conv1_fwd():
    """conv1 + bias + ReLU on the tensor core vs F.conv2d (fp32)."""
    import torch.nn.functional as F
    from nanofed_b200.ops import tc
    c, b = 3, 7
    x1 = torch.randn(c, b + 1, 28 * 28, device="cuda")[:, 1:]
    n = 1024
    slots = torch.randn(c, n, device="cuda") * 0.3
    w_off, b_off = 64, 512
    a1 = _guarded(c, b * W * W, 32)
    tc.conv1_fwd(x1, slots, w_off, b_off, a1, W)
    outs = []
    for k in range(c):
        wk = slots[k, w_off:w_off + 288].view(32, 1, 3, 3)
        y = F.relu(F.conv2d(x1[k].reshape(b, 1, 28, 28), wk, slots[k, b_off:b_off + 32]))
        outs.append(y.permute(0, 2, 3, 1).reshape(b * W * W, 32))
    return a1.float(), torch.stack(outs)


def check_conv_dgrad_wgrad1():
    """dgrad of conv2 fused with the weight/bias gradient of conv1 (da1 never stored)."""
    from nanofed_b200.ops import tc
    c, b, cin, cout = 3, 5, 32, 64
    r = b * W * W
    dy, w = _bf(c, r, cout, scale=0.2), _bf(c, cout, 9 * cin, scale=0.05)
    a1 = _bf(c, r, cin)
    x1 = torch.randn(c, b + 2, 28 * 28, device="cuda")[:, 1:b + 1]     # strided client dimension
    n = 1024
    grads = torch.zeros(c, n, device="cuda")
    w_off, b_off = 64, 512
    tc.conv3x3_dgrad_wgrad1(dy, w, a1, x1, grads, w_off, b_off, W)
    da1 = torch.zeros(c, r, cin, device="cuda")
    for t, s in enumerate(tc._shifts(W)):
        da1 += torch.einsum("cro,cok->crk", _shift_rows(dy, -s).float(),
                            w[:, :, t * cin:(t + 1) * cin].float())
    da1 = (da1 * (a1.float() > 0)).view(c, b, W, W, cin)
    img = x1.reshape(c, b, 28, 28)
    want = torch.zeros(c, cin, 10, device="cuda")
    for kh in range(3):
        for kw in range(3):
            win = img[:, :, kh:kh + W, kw:kw + W]                       # [c, b, 26, 26]
            want[:, :, kh * 3 + kw] = torch.einsum("cbhwk,cbhw->ck", da1, win)
    want[:, :, 9] = da1.sum(dim=(1, 2, 3))
    got = torch.cat([grads[:, w_off:w_off + cin * 9].view(c, cin, 9),
                     grads[:, b_off:b_off + cin].view(c, cin, 1)], dim=2)
    untouched = grads.clone()
    untouched[:, w_off:w_off + cin * 9] = 0
    untouched[:, b_off:b_off + cin] = 0
    assert float(untouched.abs().sum()) == 0.0, "wrote outside the conv1 gradient entries"
    return got, want


CHECKS = {"conv1_fwd": check_conv1_fwd, "conv_dgrad_wgrad1": check_conv_dgrad_wgrad1, "fc_wgrad_sgd": check_fc_wgrad_sgd, "conv_pool": check_conv_pool, "conv_wgrad_p": check_conv_wgrad_p, "conv_fwd_p": check_conv_fwd_p, "conv_dgrad_p": check_conv_dgrad_p,
          "conv_fwd": check_conv_fwd, "conv_dgrad": check_conv_dgrad,
          "conv_wgrad": check_conv_wgrad, "fc_fwd": check_fc_fwd, "fc_dgrad": check_fc_dgrad, "fc_dgrad_ragged": check_fc_dgrad_ragged,
          "fc_wgrad": check_fc_wgrad}


def run_one(name: str) -> bool:
    from nanofed_b200.ops import tc
    torch.manual_seed(0)
    got, ref = CHECKS[name]()
    torch.cuda.synchronize()
    tc.check_status(got.device)
    _check_guards()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item()
    rel = err / max(scale, 1e-6)
    tol = tolerance(name)
    bad = int(((got - ref).abs() > tol * scale).sum())
    ok = rel < tol and bad == 0 and torch.isfinite(got).all().item()
    print(f"{name:12s} max_abs_err={err:.4e} ref_max={scale:.3e} rel={rel:.3e} "
          f"mismatches={bad}/{got.numel()} {'OK' if ok else 'FAIL'}", flush=True)
    if not ok:
        d = (got - ref).abs()
        idx = torch.nonzero(d > tol * scale)[:5].tolist()
        for ix in idx:
            print("   at", ix, "got", got[tuple(ix)].item(), "want", ref[tuple(ix)].item())
    return ok


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--inproc":
        # several variants in ONE process (for compute-sanitizer runs)
        results = [run_one(name) for name in sys.argv[2:]]
        sys.exit(0 if all(results) else 1)
    if len(sys.argv) > 1:
        sys.exit(0 if run_one(sys.argv[1]) else 1)
    failed = []
    for name in CHECKS:
        try:
            p = subprocess.run([sys.executable, __file__, name], timeout=90, capture_output=True,
                               text=True)
            print(p.stdout.strip() or f"{name}: no output")
            if p.returncode != 0:
                failed.append(name)
                print(p.stderr.strip()[-800:])
        except subprocess.TimeoutExpired:
            failed.append(name)
            print(f"{name}: TIMEOUT")
    print("FAILED:", failed if failed else "none")
    sys.exit(1 if failed else 0)
