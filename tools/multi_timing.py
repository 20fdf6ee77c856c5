"""Per-call time of a multi-GPU context against a single-GPU one (python tools/multi_timing.py [ids...])."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from pinyon_b200 import Context, Engine

ids = [int(a) for a in sys.argv[1:]] or [0, 0]
for name, c in (("single", Context(ids[0])), ("multi" + str(ids), Context(ids))):
    for log2n in (20, 24):
        n = (1 << log2n) * (len(ids) if name != "single" else 1)
        pay = c.pinned(n, np.uint8)
        sums = np.zeros(8, np.uint64)
        for out in (dict(), dict(payoff=pay)):
            ts = []
            for it in range(8):
                eng = Engine(0)
                t0 = time.perf_counter()
                c.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=eng, sums=sums, **out)
                ts.append((time.perf_counter() - t0) * 1e3)
            print(f"{name} n=2^{log2n}x{n >> log2n} out={list(out)} path={c.last_path}: ms per call {[round(t, 3) for t in ts]}", flush=True)
            if name != "single":
                print("     phases of the last call (ms): dispatch %.3f reduce %.3f sync %.3f; slowest shard kernel %.3f" % (
                    c.stat("multi_dispatch_ms"), c.stat("multi_reduce_ms"), c.stat("multi_sync_ms"), c.last_kernel_ms), flush=True)
    c.close()
