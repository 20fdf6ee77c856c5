"""One hot-path call (for ncu captures): python tools/one_call.py [A|B|M] [streams_per_sm] [log2n]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from pinyon_b200 import Context, Engine

name = sys.argv[1] if len(sys.argv) > 1 else "A"
sps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
n = 1 << (int(sys.argv[3]) if len(sys.argv) > 3 else 24)
ctx = Context(0)
ctx.set_option("streams_per_sm", sps)
import os
if os.environ.get("PCU_RB"): ctx.set_option("rb", int(os.environ["PCU_RB"]))
dev = torch.device("cuda", 0)
payoff = torch.empty(n, dtype=torch.uint8, device=dev)
sums = torch.zeros(8, dtype=torch.int64, device=dev)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
cfgs = {"A": dict(rows=3, cols=3, transitions=1, depth=10), "B": dict(rows=9, cols=9, transitions=3, depth=20)}
for _ in range(2):
    if name == "M":
        ctx.rollout_mold(n, n_actions=2, depth=10, engine=Engine(0), payoff=payoff, sums=sums)
    else:
        ctx.rollout_random_tree(n, engine=Engine(0), payoff=payoff, sums=sums, **cfgs[name])
torch.cuda.synchronize()
print(name, sps, ctx.last_kernel_ms, sums.tolist()[:4])
