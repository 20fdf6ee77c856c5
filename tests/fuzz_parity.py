"""Randomised differential test: libpinyon_cuda against the oracle over random shapes, batch kinds and paths.
python tests/fuzz_parity.py [cases] [seed]   (GPU box; prints the path histogram, exits 1 on the first mismatch)"""
import collections, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from oracle import loader
from pinyon_b200 import Context, Engine

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
lane_bias = len(sys.argv) > 3 and sys.argv[3] == "lane"  # shapes the lane-per-rollout kernels accept
ctx = Context(0)
orc = loader.load_oracle()
hist = collections.Counter()
for case in range(cases):
    mold = rng.random() < (0.1 if lane_bias else 0.2)
    n = int(rng.choice([33, 100, 777, 3000, 20000] if lane_bias else [1, 5, 32, 33, 100, 777, 3000, 20000]))
    ragged = rng.random() < 0.5
    dmax = int(rng.integers(1, 9 if lane_bias else 41))
    depth = rng.integers(1, dmax + 1, n).astype(np.uint32) if ragged else dmax
    mode_p = rng.random() < 0.3
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64) if mode_p else None
    mseed, mpre = int(rng.integers(0, 2**40)), int(rng.integers(0, 2000))
    eng = None if mode_p else Engine(mseed).discard(mpre)
    d_arg = depth.astype(np.uint16) if ragged else depth
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    pay = np.full(n, 255, np.uint8)
    if mold:
        na = int(rng.integers(1, 20))
        exp = loader.mold_rollouts(orc, n, n_actions=na, depth=depth, model_seed=mseed, model_pre_discard=mpre,
                                   model_seeds=mseeds, want_traj=False)
        ctx.rollout_mold(n, n_actions=na, depth=d_arg, engine=eng, model_seeds=mseeds, payoff=pay, traj_hash=th, sums=sums)
        desc = f"mold na={na}"
        exp_pay = np.zeros(n, np.uint8)
    else:
        rows, cols, T = int(rng.integers(1, 13)), int(rng.integers(1, 13)), int(rng.choice([1, 1, 2, 3, 5, 8]))
        if lane_bias:
            rows, cols, T = int(rng.integers(1, 5)), int(rng.integers(1, 5)), 1
        thr = [(0, 1), (1, 2), (3, 10), (1, 1)][int(rng.integers(0, 4))]
        distinct = rng.random() < 0.6
        premax = int(rng.choice([0, 20, 60])) if lane_bias else int(rng.choice([0, 50, 2000]))
        kw = dict(state_seed=rng.integers(0, 2**63, n).astype(np.uint64) if distinct else int(rng.integers(0, 2**50)),
                  state_pre_discard=rng.integers(0, premax + 1, n).astype(np.uint32) if rng.random() < 0.5 else premax,
                  transition_seed=rng.integers(0, 2**63, n).astype(np.uint64) if distinct and rng.random() < 0.7 else 3,
                  payoff_bias=rng.integers(-2, 3, n).astype(np.int32) if rng.random() < 0.5 else int(rng.integers(-1, 2)))
        exp = loader.random_tree_rollouts(orc, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth,
                                          model_seed=mseed, model_pre_discard=mpre, model_seeds=mseeds, want_traj=False, **kw)
        gen = 2 * rows * cols * T
        pre = kw["state_pre_discard"]
        pos = (pre + np.uint32(gen)).astype(np.uint32) if isinstance(pre, np.ndarray) else pre + gen
        ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=d_arg, state_seed=kw["state_seed"],
                                state_pos=pos, transition_seed=kw["transition_seed"], payoff_bias=kw["payoff_bias"],
                                engine=eng, model_seeds=mseeds, payoff=pay, traj_hash=th, sums=sums)
        desc = f"tree {rows}x{cols}x{T} thr={thr} distinct={distinct} premax={premax}"
        exp_pay = exp["payoff"]
    path = ctx.last_path
    hist[path] += 1
    ok = np.array_equal(pay, exp_pay) and np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])
    if ok and eng is not None:
        total = int(depth.sum()) if ragged else dmax * n
        ref = Engine(mseed).discard(mpre + 4 * total)
        ok = [eng.next() for _ in range(5)] == [ref.next() for _ in range(5)]
    if not ok:
        print(f"MISMATCH case {case}: {desc} n={n} ragged={ragged} dmax={dmax} mode={'P' if mode_p else 'S'} path={path}")
        sys.exit(1)
print("all", cases, "cases bit-exact; paths:", dict(hist))
