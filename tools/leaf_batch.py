"""Host-side latency of ONE leaf-batch call as a search makes it: 4096 mid-tree leaves of one root,
ragged depth and position, host buffers.  python tools/leaf_batch.py [n]"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from pinyon_b200 import Context, Engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ctx = Context(0)
rng = np.random.default_rng(0)
level = rng.integers(1, 6, n)
depth = (10 - level).astype(np.uint16)
pos = (18 + level * 20 + rng.integers(0, 9, n) * level // 2).astype(np.uint32)
bias = rng.integers(-2, 3, n).astype(np.int32)
payoff = np.zeros(n, np.uint8)
sums = np.zeros(8, np.uint64)
eng = Engine(0)
for it in range(3):
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=depth, state_pos=pos, payoff_bias=bias, engine=eng, payoff=payoff, sums=sums)
t0 = time.perf_counter()
K = 50
for it in range(K):
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=depth, state_pos=pos, payoff_bias=bias, engine=eng, payoff=payoff, sums=sums)
dt = (time.perf_counter() - t0) / K
print("path", ctx.last_path, "n", n, "ms/call", dt * 1e3, "kernel_ms", ctx.last_kernel_ms, "launches/call", ctx.launches / (K + 3))
# uniform depth for comparison (fused path)
for it in range(3):
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=7, state_pos=pos, payoff_bias=bias, engine=eng, payoff=payoff, sums=sums)
t0 = time.perf_counter()
for it in range(K):
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=7, state_pos=pos, payoff_bias=bias, engine=eng, payoff=payoff, sums=sums)
dt = (time.perf_counter() - t0) / K
print("path", ctx.last_path, "n", n, "ms/call", dt * 1e3, "kernel_ms", ctx.last_kernel_ms)
