import sys, torch
sys.path.insert(0, "/root/repo")
from nanofed_b200 import ops
c, n = 32, 1199936
p = torch.randn(c, n, device="cuda"); g = torch.randn(c, n, device="cuda") * 0.01
low = torch.zeros(c, n, dtype=torch.bfloat16, device="cuda")
mask = torch.ones(n, dtype=torch.uint8, device="cuda"); ctr = torch.ones(1, dtype=torch.int64, device="cuda")
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3
ss, _ = ops.update_stats(g)
print("update_stats us", t(lambda: ops.update_stats(g)))
print("sgd plain us", t(lambda: ops.fused_sgd_(p.view(-1), g.view(-1), 0.1, lowp_out=low.view(-1), row_len=n)))
print("sgd clip only us", t(lambda: ops.fused_sgd_(p.view(-1), g.view(-1), 0.1, lowp_out=low.view(-1), row_len=n, row_sumsq=ss, clip_norm=1.0)))
print("sgd clip+noise us", t(lambda: ops.fused_sgd_(p.view(-1), g.view(-1), 0.1, lowp_out=low.view(-1), row_len=n, row_sumsq=ss, clip_norm=1.0, noise_std=0.1, seed=3, step_counter=ctr)))
print("sgd clip+noise+mask us", t(lambda: ops.fused_sgd_(p.view(-1), g.view(-1), 0.1, lowp_out=low.view(-1), row_len=n, row_sumsq=ss, clip_norm=1.0, noise_std=0.1, seed=3, noise_mask=mask, step_counter=ctr)))
