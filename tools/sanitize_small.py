"""Small invocations of every kernel path, for compute-sanitizer (memcheck / racecheck)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from pinyon_b200 import Context, Engine

ctx = Context(0)
n = 3000
seeds = np.arange(n, dtype=np.uint64) + np.uint64(5)
pay = np.zeros(n, np.uint8)
sums = np.zeros(8, np.uint64)
paths = []
def run(**kw):
    ctx.rollout_random_tree(n, payoff=pay, sums=sums, **kw)
    paths.append(ctx.last_path)
run(rows=3, cols=3, depth=10, engine=Engine(0))
run(rows=9, cols=9, transitions=3, depth=20, engine=Engine(0))
run(rows=3, cols=3, depth=10, state_seed=seeds, engine=Engine(0))
run(rows=3, cols=3, depth=10, state_seed=seeds, model_seeds=seeds)
ctx.set_option("lane_full_columns", 1)
run(rows=3, cols=3, depth=10, state_seed=seeds, engine=Engine(0))
ctx.set_option("lane_full_columns", 0)
run(rows=9, cols=9, transitions=3, depth=20, state_seed=seeds, transition_seed=seeds, engine=Engine(0))
run(rows=9, cols=9, transitions=3, depth=20, state_seed=seeds, transition_seed=seeds, model_seeds=seeds)
d = (np.arange(n) % 30).astype(np.uint16)
ctx.rollout_mold(n, n_actions=3, depth=d, engine=Engine(1), payoff=pay, sums=sums); paths.append(ctx.last_path)
ctx.rollout_mold(n, n_actions=3, depth=d, model_seeds=seeds, payoff=pay, sums=sums); paths.append(ctx.last_path)
ctx.rollout_mold(n, n_actions=2, depth=10, engine=Engine(1), payoff=pay, sums=sums); paths.append(ctx.last_path)
for m in (1, 7, 32):
    ctx.rollout_random_tree(m, rows=5, cols=4, transitions=3, depth=9, state_seed=seeds[:m].copy(), engine=Engine(2),
                            payoff=pay[:m], sums=sums); paths.append(ctx.last_path)
    ctx.rollout_mold(m, n_actions=2, depth=10, engine=Engine(1), payoff=pay[:m], sums=sums); paths.append(ctx.last_path)
# mode W (leaf batches of many searches), packed payoffs, variant kernels, a multi-GPU context sharing this GPU
from pinyon_b200 import Engine as E
depth = (np.arange(n) % 9 + 1).astype(np.uint16)
words = np.concatenate([E(int(s)).take_raw(4 * int(k)) for s, k in zip(seeds, depth)])
pos = (18 + np.arange(n) % 200).astype(np.uint32)
ctx.rollout_random_tree(n, rows=3, cols=3, depth=depth, state_pos=pos, model_words=words, payoff=pay, sums=sums); paths.append(ctx.last_path)
ctx.rollout_random_tree(n, rows=3, cols=4, transitions=2, depth=depth, state_seed=seeds, transition_seed=seeds, model_words=words, payoff=pay, sums=sums); paths.append(ctx.last_path)
packed = np.zeros((n + 3) // 4, np.uint8)
ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(0), payoff_packed=packed, sums=sums); paths.append(ctx.last_path)
ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, state_seed=seeds, engine=Engine(0), payoff_packed=packed, sums=sums); paths.append(ctx.last_path)
xs = np.array([12345], np.uint64)
ctx.rollout_random_tree(n, rows=4, cols=3, transitions=2, depth=depth, state_seed=seeds, transition_seed=seeds, xs_state=xs, xor_shift=True, payoff=pay, sums=sums); paths.append(ctx.last_path)
ctx.rollout_random_tree(n, rows=1, cols=3, transitions=2, depth=depth, state_seed=seeds, state_pos=np.full(n, 12, np.uint32), engine=Engine(3), growth=1, payoff=pay, sums=sums); paths.append(ctx.last_path)
ctx.rollout_mold(n, n_actions=3, depth=depth, xs_state=xs, xor_shift=True, payoff=pay, sums=sums); paths.append(ctx.last_path)
from pinyon_b200 import Context as C2
m = C2([0, 0]); m.set_option("multi_min_rollouts", 100)
m.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(0), payoff=pay, sums=sums); paths.append(m.last_path)
m.close()
print("paths:", paths)
ctx.close()
