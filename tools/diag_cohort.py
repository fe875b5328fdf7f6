import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import torch, torch.nn.functional as F
from test_mnist_cohort_gpu import _setup
for b in (16, 64):
    cfg, model, layout, arena, shards, eng = _setup(c=2, b=b, steps=1)
    eng.step(0); torch.cuda.synchronize()
    ref = model.cuda().eval()
    for k in range(2):
        ref.zero_grad()
        x, y = shards.images[k, :b], shards.labels[k, :b]
        F.cross_entropy(ref(x), y).backward()
        g32 = {n: p.grad.clone() for n, p in ref.named_parameters()}
        ref.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = ref(x)
        F.cross_entropy(out.float(), y).backward()
        got = layout.views(eng.grads[k])
        for n, p in ref.named_parameters():
            r = g32[n]; a = p.grad; g = got[n].float()
            print(f"B={b} k={k} {n:14s} ours-vs-fp32 {(g-r).norm()/r.norm():.4f}  autocast-vs-fp32 {(a-r).norm()/r.norm():.4f}  cos {F.cosine_similarity(g.flatten(), r.flatten(), dim=0):.5f}")
