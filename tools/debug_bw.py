"""Which rollouts differ from the oracle, by stream (debugging the forward/backward generator)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from oracle import loader
from pinyon_b200 import Context, Engine

ctx = Context(0)
orc = loader.load_oracle()
for n, sps in ((1000, 0), (5000, 1), (20000, 1), (60000, 1), (60000, 0), (300000, 0)):
    ctx.set_option("streams_per_sm", sps)
    exp = loader.random_tree_rollouts(orc, n, rows=3, cols=3, depth=10, want_traj=False)
    pay = np.full(n, 255, np.uint8)
    th = np.zeros(n, np.uint32)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(0), payoff=pay, traj_hash=th)
    chunk, streams = int(ctx.stat("last_chunk")), int(ctx.stat("last_streams"))
    bad = np.nonzero(th != exp["traj_hash"])[0]
    print(f"n={n} sps={sps} chunk={chunk} streams={streams} bad={len(bad)}")
    if len(bad):
        by = {}
        for i in bad:
            by.setdefault(int(i) // chunk, []).append(int(i) % chunk)
        for s, v in sorted(by.items())[:12]:
            print(f"   stream {s} ({'bwd' if s % 2 and s != streams - 1 else 'fwd'}): {len(v)} bad, local range {min(v)}..{max(v)}, first few {v[:6]}")
