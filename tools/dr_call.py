"""Distinct-root calls: python tools/dr_call.py [A|B] [S|P] [log2n]
state_seed_i = i, transition_seed_i = i * 0x9E3779B97F4A7C15 mod 2^64, root position 2*R*C*T."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from pinyon_b200 import Context, Engine

name = sys.argv[1] if len(sys.argv) > 1 else "A"
mode = sys.argv[2] if len(sys.argv) > 2 else "S"
n = 1 << (int(sys.argv[3]) if len(sys.argv) > 3 else 20)
ctx = Context(0)
dev = torch.device("cuda", 0)
payoff = torch.empty(n, dtype=torch.uint8, device=dev)
sums = torch.zeros(8, dtype=torch.int64, device=dev)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
cfgs = {"A": dict(rows=3, cols=3, transitions=1, depth=10), "B": dict(rows=9, cols=9, transitions=3, depth=20)}
cfg = cfgs[name]
idx = torch.arange(n, dtype=torch.int64, device=dev)
seeds = idx.clone()
tseeds = idx * torch.tensor(0x9E3779B97F4A7C15 - (1 << 64), dtype=torch.int64, device=dev)
kw = dict(state_seed=seeds, transition_seed=tseeds)
if mode == "P":
    kw["model_seeds"] = idx.clone()
for it in range(3):
    sums.zero_()
    if mode == "P":
        ctx.rollout_random_tree(n, payoff=payoff, sums=sums, **cfg, **kw)
    else:
        ctx.rollout_random_tree(n, engine=Engine(0), payoff=payoff, sums=sums, **cfg, **kw)
torch.cuda.synchronize()
ms = ctx.last_kernel_ms
print(name, mode, n, ctx.last_path, "kernel_ms", ms, "rollouts/s", n / (ms * 1e-3), sums.tolist()[:4])
