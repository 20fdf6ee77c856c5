"""Stress of k_warp2's seeder/player hand-off (spin waits on shared-memory counters): many calls over random batch sizes,
shapes, depths and lock-step widths, each compared with k_seed_pre + k_warp (option warp2 = 0) byte for byte.
python tools/stress_warp2.py [iterations] [seed]"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from pinyon_b200 import Context, Engine

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 300
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
ctx = Context(0)
t0 = time.time()
rollouts = 0
for it in range(iters):
    small = rng.random() < 0.5
    n = int(rng.integers(33, 700)) if small else int(rng.integers(700, 60000))
    rows, cols = int(rng.integers(1, 10)), int(rng.integers(1, 10))
    T = int(rng.integers(1, 5))
    dmax = int(rng.integers(1, 45))
    ragged = rng.random() < 0.5
    depth = rng.integers(1, dmax + 1, n).astype(np.uint16) if ragged else dmax
    seeds = rng.integers(0, 2**63, n).astype(np.uint64)
    tseeds = rng.integers(0, 2**63, n).astype(np.uint64) if rng.random() < 0.6 else int(rng.integers(0, 2**31))
    pre = rng.integers(0, 4000, n).astype(np.uint32) + np.uint32(2 * rows * cols * T) if rng.random() < 0.5 else None
    lock = int(rng.choice([1, 2, 4]))
    res = []
    for warp2 in (1, 0):
        ctx.set_option("warp2", warp2)
        ctx.set_option("warp2_lock", lock)
        ctx.set_option("force_warp", 1)
        pay = np.full(n, 255, np.uint8)
        th = np.zeros(n, np.uint32)
        sums = np.zeros(8, np.uint64)
        ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, depth=depth, state_seed=seeds, state_pos=pre,
                                transition_seed=tseeds, engine=Engine(it), payoff=pay, traj_hash=th, sums=sums)
        assert ctx.last_path == "warp_private_S", ctx.last_path
        res.append((pay, th, sums))
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2]), \
        (it, n, rows, cols, T, dmax, lock)
    rollouts += n
ctx.set_option("warp2", 1); ctx.set_option("warp2_lock", 1); ctx.set_option("force_warp", 0)
print("stress ok: %d calls, %d rollouts, %.1f s" % (iters, rollouts, time.time() - t0))
