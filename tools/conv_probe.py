"""Bottleneck probe for the persistent conv kernel: time fwd/dgrad with parts disabled."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from nanofed_b200.ops import tc
c, b = 32, 64
r = b * 676
x = torch.randn(c, r, 32, device="cuda").bfloat16()
dy = torch.randn(c, r, 64, device="cuda").bfloat16()
w = (torch.randn(c, 64, 288, device="cuda") * 0.05).bfloat16()
bias = torch.randn(c, 64, device="cuda")
o64 = torch.empty(c, r, 64, device="cuda", dtype=torch.bfloat16)
o32 = torch.empty(c, r, 32, device="cuda", dtype=torch.bfloat16)
def t(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): fn()
    e.record(); torch.cuda.synchronize()
    return a.elapsed_time(e) / 10 * 1000
print("debug", os.environ.get("NFB_CONV_DEBUG", "0"),
      "fwd %.1f us" % t(lambda: tc.conv3x3_fwd_persistent(x, w, bias, o64, 26)),
      "dgrad %.1f us" % t(lambda: tc.conv3x3_dgrad_persistent(dy, w, o32, 26, relu_mask=x)))
p = torch.zeros(c, b, 9216, device="cuda", dtype=torch.bfloat16)
idx = torch.zeros(c, b, 144, 64, device="cuda", dtype=torch.uint8)
rng = torch.zeros(1, dtype=torch.int64, device="cuda")
x1 = torch.randn(c, b, 784, device="cuda")
g1 = torch.zeros(c, 1024, device="cuda")
print("debug", os.environ.get("NFB_CONV_DEBUG", "0"),
      "fwd+pool %.1f us" % t(lambda: tc.conv3x3_pool_fwd(x, w, bias, p, idx, b, 26, 0.25, rng, 1)),
      "dgrad+wgrad1 %.1f us" % t(lambda: tc.conv3x3_dgrad_wgrad1(dy, w, x, x1, g1, 64, 512, 26)))
n_total = 64 * 32 * 9 + 64
grad = torch.zeros(c, n_total, device="cuda")
print("debug", os.environ.get("NFB_CONV_DEBUG", "0"),
      "wgrad %.1f us" % t(lambda: tc.conv3x3_wgrad_persistent(dy, x, grad, 26, n_total)))
