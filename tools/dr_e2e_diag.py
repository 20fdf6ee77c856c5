"""Where the DR end-to-end step goes: H2D of the seeds alone, the device-resident call, the host-buffer call."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from pinyon_b200 import Context, Engine
n = 1 << 24
ctx = Context(0)
dev = torch.device("cuda", 0)
seeds_h = ctx.pinned(n, np.uint64); seeds_h[:] = np.arange(n, dtype=np.uint64)
pay_h = ctx.pinned(n, np.uint8); sums_h = np.zeros(8, np.uint64)
seeds_d = torch.from_numpy(seeds_h.view(np.int64)).to(dev)
pay_d = torch.empty(n, dtype=torch.uint8, device=dev); sums_d = torch.zeros(8, dtype=torch.int64, device=dev)
t = torch.from_numpy(seeds_h.view(np.int64))
def timeit(f, k=5):
    f(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(k): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / k * 1e3
print("H2D 134 MB alone      %.2f ms" % timeit(lambda: seeds_d.copy_(t, non_blocking=True)))
print("device-resident call  %.2f ms" % timeit(lambda: ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, state_seed=seeds_d, engine=Engine(0), payoff=pay_d, sums=sums_d)))
print("host-buffer call      %.2f ms" % timeit(lambda: ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, state_seed=seeds_h, engine=Engine(0), payoff=pay_h, sums=sums_h)))
ctx.set_option("zero_copy", 0)
print("host, staged payoffs  %.2f ms" % timeit(lambda: ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, state_seed=seeds_h, engine=Engine(0), payoff=pay_h, sums=sums_h)))
