"""One distinct-roots call of the general path (for ncu captures): python tools/one_call_dr.py [log2n] [warp2 0|1] [T]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from pinyon_b200 import Context, Engine

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 18
warp2 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
T = int(sys.argv[3]) if len(sys.argv) > 3 else 3
lock = int(sys.argv[4]) if len(sys.argv) > 4 else 0
ts = int(sys.argv[5]) if len(sys.argv) > 5 else 0  # 1: per-rollout transition seeds as well
n = 1 << log2n
ctx = Context(0)
ctx.set_option("warp2", warp2)
if lock: ctx.set_option("warp2_lock", lock)
dev = torch.device("cuda", 0)
payoff = torch.empty(n, dtype=torch.uint8, device=dev)
sums = torch.zeros(8, dtype=torch.int64, device=dev)
seeds = torch.arange(n, dtype=torch.int64, device=dev)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
tseeds = (seeds * 0x9E3779B97F4A7C15) if ts else 0
ms = []
for _ in range(3):
    ctx.rollout_random_tree(n, rows=9, cols=9, transitions=T, depth=20, state_seed=seeds, transition_seed=tseeds, engine=Engine(0), payoff=payoff, sums=sums)
    ms.append(round(ctx.last_kernel_ms, 4))
torch.cuda.synchronize()
print("warp2", warp2, "lock", lock, "tseeds", ts, "T", T, "n", n, ctx.last_path, ms, "rollouts/s %.3e" % (n / (ms[-1] * 1e-3)), sums.tolist()[:4])
