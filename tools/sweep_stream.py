"""Sweep streams_per_sm for the fused stream kernel: kernel-only ms (pcu_last_kernel_ms) and whole-call ms."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from pinyon_b200 import Context, Engine

def main():
    ctx = Context(0)
    dev = torch.device("cuda", 0)
    n = 1 << 24
    payoff = torch.empty(n, dtype=torch.uint8, device=dev)
    sums = torch.zeros(8, dtype=torch.int64, device=dev)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    cfgs = {"A": dict(rows=3, cols=3, transitions=1, depth=10), "B": dict(rows=9, cols=9, transitions=3, depth=20)}
    for name in sys.argv[1:] or ["A"]:
        for sps in (4, 8, 16):
            ctx.set_option("streams_per_sm", sps)
            ks, ts = [], []
            for it in range(6):
                eng = Engine(0)
                torch.cuda.synchronize(); t0 = time.perf_counter()
                if name == "M":
                    ctx.rollout_mold(n, n_actions=2, depth=10, engine=eng, payoff=payoff, sums=sums)
                else:
                    ctx.rollout_random_tree(n, engine=eng, payoff=payoff, sums=sums, **cfgs[name])
                torch.cuda.synchronize(); t1 = time.perf_counter()
                if it >= 2:
                    ks.append(ctx.last_kernel_ms); ts.append((t1 - t0) * 1e3)
            print(f"{name} streams_per_sm={sps:2d} kernel {min(ks):.3f} ms  call {min(ts):.3f} ms", flush=True)
main()
