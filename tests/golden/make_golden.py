#!/usr/bin/env python
"""Generate tests/golden/reference_vectors.json from the REFERENCE ITSELF.

Runs only where /root/reference exists (the build container): it compiles the reference's
own headers into oracle/_ref/libpinyon_ref.so (oracle/Makefile, oracle/ref_harness.cc) and
records what the reference's MonteCarloModel / RandomTree / MoldState / prng produce for a
set of seeded cases.  The committed JSON is what pins oracle/pinyon_oracle.c (and, on the
GPU box, libpinyon_cuda) when the reference is not around.

    python tests/golden/make_golden.py
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import loader  # noqa: E402

OUT = Path(__file__).resolve().parent / "reference_vectors.json"

# rows, cols, T, thr, depth, state_seed, state_pre_discard, transition_seed, payoff_bias, model_seed, model_pre_discard
TREE_CASES = [
    dict(rows=3, cols=3, transitions=1, thr=(0, 1), depth=10),  # BASELINE configs 1-2 (Appendix B1)
    dict(rows=9, cols=9, transitions=3, thr=(0, 1), depth=20),  # BASELINE config 3 (Appendix B2)
    dict(rows=2, cols=2, transitions=2, thr=(1, 2), depth=5, state_seed=7, state_pre_discard=13, transition_seed=1,
         payoff_bias=1, model_seed=99, model_pre_discard=7),
    dict(rows=4, cols=7, transitions=4, thr=(3, 10), depth=9, state_seed=12345678901234, state_pre_discard=700,
         transition_seed=0xDEADBEEFCAFE, payoff_bias=-2, model_seed=2**32 + 5, model_pre_discard=1001),
    dict(rows=1, cols=1, transitions=1, thr=(0, 1), depth=1, state_seed=5, model_seed=1, model_pre_discard=623),
    dict(rows=5, cols=2, transitions=9, thr=(1, 1), depth=3, state_seed=99, transition_seed=3, model_seed=3,
         model_pre_discard=624),
    dict(rows=16, cols=17, transitions=1, thr=(0, 1), depth=6, state_seed=3, model_seed=3, model_pre_discard=2),
]
MOLD_CASES = [dict(n_actions=2, depth=10), dict(n_actions=2, depth=50), dict(n_actions=3, depth=1),
              dict(n_actions=7, depth=4, model_seed=11, model_pre_discard=3), dict(n_actions=17, depth=3)]
N_TRAJ = 48  # rollouts recorded with full trajectories per case
COUNTS = {0: 10**6, 1: 20000}  # Appendix B counts: case index -> rollouts


def main():
    if loader.build_reference() is None:
        raise SystemExit("needs /root/reference")
    ref = loader.load_reference()
    doc = {"generator": "tests/golden/make_golden.py", "source": "oracle/_ref/libpinyon_ref.so (reference headers "
           "compiled unmodified, g++ 13, libstdc++ 13)", "tree": [], "mold": [], "prng": {}, "private": []}

    for idx, case in enumerate(TREE_CASES):
        res = loader.random_tree_rollouts(ref, N_TRAJ, prefix="ref", **case)
        entry = dict(case=case, n=N_TRAJ, payoff_x2=res["payoff"].tolist(), traj=res["traj"].reshape(N_TRAJ, -1).tolist(),
                     steps=int(res["steps"]))
        if idx in COUNTS:
            big = loader.random_tree_rollouts(ref, COUNTS[idx], prefix="ref", want_traj=False, **case)
            c = np.bincount(big["payoff"], minlength=3)
            entry["counts"] = dict(n=COUNTS[idx], wins=int(c[2]), draws=int(c[1]), losses=int(c[0]))
        strat = np.zeros(case["rows"] * case["cols"] * case["transitions"])
        ref.ref_random_tree_root_strategies(case.get("state_seed", 0), case.get("state_pre_discard", 0), case["depth"],
                                            case["rows"], case["cols"], case["transitions"], *case["thr"],
                                            strat.ctypes.data)
        entry["root_strategies_hex"] = [float(x).hex() for x in strat[:27]]
        doc["tree"].append(entry)

    # distinct roots, ragged depth, per-rollout transition seeds, both model-PRNG modes
    rng = np.random.default_rng(2024)
    n = 40
    for rows, cols, T, thr in [(3, 3, 1, (0, 1)), (5, 4, 3, (3, 10)), (9, 9, 3, (0, 1))]:
        arrays = dict(state_seed=rng.integers(0, 2**63, n).astype(np.uint64),
                      state_pre_discard=rng.integers(0, 1500, n).astype(np.uint32),
                      transition_seed=rng.integers(0, 2**63, n).astype(np.uint64),
                      depth=rng.integers(1, 14, n).astype(np.uint32),
                      payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
        mseeds = rng.integers(0, 2**63, n).astype(np.uint64)
        for mode in ("S", "P"):
            res = loader.random_tree_rollouts(ref, n, prefix="ref", rows=rows, cols=cols, transitions=T, thr=thr,
                                              model_seed=21, model_pre_discard=5,
                                              model_seeds=mseeds if mode == "P" else None, **arrays)
            doc["private"].append(dict(rows=rows, cols=cols, transitions=T, thr=thr, mode=mode, n=n, model_seed=21,
                                       model_pre_discard=5, model_seeds=[int(x) for x in mseeds],
                                       arrays={k: [int(x) for x in v] for k, v in arrays.items()},
                                       payoff_x2=res["payoff"].tolist(), traj=res["traj"].reshape(n, -1).tolist(),
                                       steps=int(res["steps"])))

    for case in MOLD_CASES:
        res = loader.mold_rollouts(ref, N_TRAJ, prefix="ref", **case)
        doc["mold"].append(dict(case=case, n=N_TRAJ, payoff_row=res["payoff_row"].tolist(),
                                payoff_col=res["payoff_col"].tolist(), traj=res["traj"].reshape(N_TRAJ, -1).tolist(),
                                steps=int(res["steps"])))

    # variants: the xor_shift PRNG as a type list (with the one member IsPRNG misses added) and
    # GrowLib::alternating_constant_actions as actions_func; mode S, ragged depths, per-rollout everything
    doc["variants"] = []
    rng = np.random.default_rng(77)
    n = 40
    for engine in ("mt", "xs"):
        for growth in (0, 1):
            if engine == "mt" and growth == 0:
                continue
            for rows, cols, T, thr in [(3, 3, 1, (0, 1)), (1, 3, 2, (3, 10)), (3, 1, 1, (0, 1)), (4, 5, 3, (1, 2))]:
                arrays = dict(state_seed=rng.integers(1, 2**63, n).astype(np.uint64),
                              state_pre_discard=rng.integers(0, 50, n).astype(np.uint32),
                              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64),
                              depth=rng.integers(1, 9, n).astype(np.uint32),
                              payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
                res = loader.random_tree_rollouts(ref, n, prefix="ref", rows=rows, cols=cols, transitions=T, thr=thr,
                                                  model_seed=12345, model_pre_discard=7, engine=engine, growth=growth, **arrays)
                e = dict(engine=engine, growth=growth, rows=rows, cols=cols, transitions=T, thr=thr, n=n, model_seed=12345,
                         model_pre_discard=7, arrays={k: [int(x) for x in v] for k, v in arrays.items()},
                         payoff_x2=res["payoff"].tolist(), traj=res["traj"].reshape(n, -1).tolist(), steps=int(res["steps"]))
                if engine == "xs":
                    e["state_after_ctor"] = [int(x) for x in res["state_after_ctor"]]
                doc["variants"].append(e)
    d = rng.integers(0, 12, n).astype(np.uint32)
    res = loader.mold_rollouts(ref, n, prefix="ref", n_actions=3, depth=d, model_seed=99, engine="xs")
    doc["variants_mold_xs"] = dict(n=n, depth=[int(x) for x in d], model_seed=99, traj=res["traj"].reshape(n, -1).tolist())

    prng = {}
    for seed, disc in [(0, 0), (7, 620), (2**32 + 5, 1), (12345678901234, 100000)]:
        u64 = np.zeros(16, np.uint64)
        ud = np.zeros(16, np.float64)
        ref.ref_prng_uniform_64(seed, disc, 16, u64.ctypes.data)
        ref.ref_prng_uniform(seed, disc, 16, ud.ctypes.data)
        ri = {}
        for m in (2, 3, 9, 10, 17, 255):
            a = np.zeros(16, np.int32)
            ref.ref_prng_random_int(seed, disc, m, 16, a.ctypes.data)
            ri[str(m)] = a.tolist()
        prng[f"{seed}:{disc}"] = dict(uniform_64=[int(x) for x in u64], uniform_hex=[float(x).hex() for x in ud],
                                      random_int=ri)
    doc["prng"] = prng
    doc["sizeof"] = dict(prng=int(ref.ref_sizeof_prng()), random_tree_state=int(ref.ref_sizeof_random_tree_state()))

    stats = np.zeros(2 + 3 + 3, np.int64)
    vt = np.zeros(1, np.float64)
    ref.ref_tree_bandit_exp3(0, 10, 3, 3, 1, 0, 0, 4096, stats.ctypes.data, vt.ctypes.data)
    doc["tree_bandit_exp3_b4"] = dict(stats=stats.tolist(), value_total_row=float(vt[0]))

    OUT.write_text(json.dumps(doc, separators=(",", ":")) + "\n")
    print(f"wrote {OUT} ({OUT.stat().st_size} bytes)")


if __name__ == "__main__":
    main()
