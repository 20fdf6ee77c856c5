"""TEST INFRASTRUCTURE ONLY — ctypes access to the parity checkers.

  load_oracle()     oracle/_ref/liboracle.so       the plain-C restatement (pinyon_oracle.c)
  load_reference()  oracle/_ref/libpinyon_ref*.so  the reference itself (ref_harness.cc), or None

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module.  Nothing under pinyon_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
REFERENCE = Path(os.environ.get("PINYON_REFERENCE", "/root/reference"))

u64, u32, i32, vp = C.c_uint64, C.c_uint32, C.c_int32, C.c_void_p

_RT_ARGS = [u64, vp, u64, vp, u32, vp, u64, vp, u32, vp, i32, u32, u32, u32, C.c_int, C.c_int, vp, u64, u64, vp, vp,
            u32]


def build_oracle(force: bool = False) -> Path:
    lib = OUT / "liboracle.so"
    src = HERE / "pinyon_oracle.c"
    if force or not lib.exists() or lib.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(HERE), "-B" if force else "-s", "all"], check=True, capture_output=True)
    return lib


def build_reference(force: bool = False):
    """Compile the reference's own headers where they lie (only possible where
    /root/reference exists, i.e. in the build container, not on the GPU box)."""
    if not (REFERENCE / "src" / "model" / "monte-carlo-model.hh").exists():
        return None
    args = ["make", "-C", str(HERE), f"REFERENCE={REFERENCE}", "ref"]
    if force:
        args.insert(3, "-B")
    subprocess.run(args, check=True, capture_output=True)
    return OUT / "libpinyon_ref.so"


def _has_avx2() -> bool:
    try:
        return " avx2 " in open("/proc/cpuinfo").read()
    except OSError:
        return False


def load_oracle() -> C.CDLL:
    lib = C.CDLL(str(build_oracle()))
    lib.orc_random_tree_rollouts.argtypes = _RT_ARGS + [vp, vp]
    lib.orc_random_tree_rollouts.restype = u64
    lib.orc_mold_rollouts.argtypes = [u64, u32, vp, u32, vp, u64, u64, vp, vp, u32, vp, vp]
    lib.orc_mold_rollouts.restype = u64
    for name in ("orc_prng_uniform_64", "orc_prng_uniform"):
        getattr(lib, name).argtypes = [u64, u64, u64, vp]
    lib.orc_prng_random_int.argtypes = [u64, u64, C.c_int, u64, vp]
    lib.orc_mt_words.argtypes = [u64, u64, u64, vp]
    lib.orc_mt_state.argtypes = [u64, u64, vp, vp]
    lib.orc_set_model_engine.argtypes = [vp, u32]
    lib.orc_set_model_engine.restype = None
    lib.orc_set_engine_kind.argtypes = [C.c_int]
    lib.orc_set_growth.argtypes = [C.c_int]
    lib.orc_set_state_after_ctor.argtypes = [vp]
    lib.orc_random_tree_root_strategies.argtypes = [u64, u64, u32, u32, u32, u32, C.c_int, C.c_int, vp]
    return lib


def load_reference():
    """The compiled reference, or None when it was never built (GPU box without a prebuilt copy)."""
    names = ["libpinyon_ref_v3.so", "libpinyon_ref.so"] if _has_avx2() else ["libpinyon_ref.so"]
    for name in names:
        path = OUT / name
        if path.exists():
            try:
                lib = C.CDLL(str(path))
            except OSError:
                continue
            lib.ref_random_tree_rollouts.argtypes = _RT_ARGS
            lib.ref_random_tree_rollouts.restype = u64
            lib.ref_mold_rollouts.argtypes = [u64, u32, vp, u32, vp, u64, u64, vp, vp, vp, u32]
            lib.ref_mold_rollouts.restype = u64
            if hasattr(lib, "ref_xs_random_tree_rollouts"):  # variants: xor_shift type list, grow-lib growth function
                lib.ref_xs_random_tree_rollouts.argtypes = _RT_ARGS + [vp]
                lib.ref_xs_random_tree_rollouts.restype = u64
                lib.ref_xs_mold_rollouts.argtypes = [u64, u32, vp, u32, vp, u64, u64, vp, vp, vp, u32]
                lib.ref_xs_mold_rollouts.restype = u64
                lib.ref_set_growth.argtypes = [C.c_int]
            for fn in ("ref_prng_uniform_64", "ref_prng_uniform"):
                getattr(lib, fn).argtypes = [u64, u64, u64, vp]
            lib.ref_prng_random_int.argtypes = [u64, u64, C.c_int, u64, vp]
            lib.ref_random_tree_root_strategies.argtypes = [u64, u64, u32, u32, u32, u32, C.c_int, C.c_int, vp]
            lib.ref_time_rollouts.argtypes = [C.c_int, u64, u32, u64, u32, u32, u32, u32, u64, vp, vp]
            lib.ref_time_rollouts.restype = C.c_double
            lib.ref_tree_bandit_exp3.argtypes = [u64, u32, u32, u32, u32, u64, u64, u64, vp, vp]
            lib.ref_sizeof_prng.restype = u64
            lib.ref_sizeof_random_tree_state.restype = u64
            lib._pinyon_path = str(path)
            return lib
    return None


def set_model_engine(lib, engine):
    """The NEXT orc_* rollout call starts its model engine from `engine` (a pinyon_b200.Engine image, e.g.
    Engine(0).discard(k)) instead of stepping model_pre_discard draws: full-size windows far into the stream."""
    lib.orc_set_model_engine(engine.x.ctypes.data, int(engine.p[0]))


def _arr(a, dtype):
    if a is None or np.isscalar(a) or isinstance(a, int):
        return None, (0 if a is None else int(a))
    a = np.ascontiguousarray(a, dtype=dtype)
    return a, 0


def random_tree_rollouts(lib, n, *, rows, cols, transitions=1, thr=(0, 1), depth, state_seed=0, state_pre_discard=0,
                         transition_seed=0, payoff_bias=0, model_seed=0, model_pre_discard=0, model_seeds=None,
                         want_traj=True, prefix="orc", engine="mt", growth=0):
    """Run n RandomTree rollouts on a checker library.  `prefix` "orc" = restatement,
    "ref" = compiled reference.  Per-rollout numpy arrays or batch-uniform scalars.
    NOTE: states are described the reference's way — seed plus draws discarded BEFORE
    the constructor runs; the CUDA library's state_pos is that plus 2*rows*cols*transitions."""
    ss, ss_u = _arr(state_seed, np.uint64)
    sp, sp_u = _arr(state_pre_discard, np.uint32)
    ts, ts_u = _arr(transition_seed, np.uint64)
    dp, dp_u = _arr(depth, np.uint32)
    pb, pb_u = _arr(payoff_bias, np.int32)
    ms, _ = _arr(model_seeds, np.uint64)
    dmax = int(dp.max()) if dp is not None and len(dp) else dp_u
    payoff = np.zeros(n, np.uint8)
    traj = np.zeros((n, max(dmax, 1), 3), np.uint8) if want_traj else None
    p = lambda a: a.ctypes.data if a is not None else None  # noqa: E731
    args = [n, p(ss), ss_u, p(sp), sp_u, p(ts), ts_u, p(dp), dp_u, p(pb), pb_u, rows, cols, transitions, thr[0], thr[1],
            p(ms), model_seed, model_pre_discard, p(payoff), p(traj), max(dmax, 1)]
    res = {"payoff": payoff, "traj": traj}
    # engine="xs": Pinyon's xor_shift PRNG (seeds are then engine STATES); growth=1: GrowLib::alternating_constant_actions
    state_after = np.zeros(n, np.uint64) if engine == "xs" else None
    if prefix == "orc":
        th = np.zeros(n, np.uint32)
        sums = np.zeros(8, np.uint64)
        lib.orc_set_engine_kind(1 if engine == "xs" else 0)
        lib.orc_set_growth(growth)
        lib.orc_set_state_after_ctor(p(state_after))
        try:
            steps = lib.orc_random_tree_rollouts(*args, p(th), p(sums))
        finally:
            lib.orc_set_engine_kind(0)
            lib.orc_set_growth(0)
            lib.orc_set_state_after_ctor(None)
        res.update(traj_hash=th, sums=sums)
    else:
        lib.ref_set_growth(growth)
        try:
            if engine == "xs":
                steps = lib.ref_xs_random_tree_rollouts(*args, p(state_after))
            else:
                steps = lib.ref_random_tree_rollouts(*args)
        finally:
            lib.ref_set_growth(0)
    if state_after is not None:
        res["state_after_ctor"] = state_after
    res["steps"] = steps
    return res


def mold_rollouts(lib, n, *, n_actions, depth, model_seed=0, model_pre_discard=0, model_seeds=None, want_traj=True,
                  prefix="orc", engine="mt"):
    dp, dp_u = _arr(depth, np.uint32)
    ms, _ = _arr(model_seeds, np.uint64)
    dmax = int(dp.max()) if dp is not None and len(dp) else dp_u
    traj = np.zeros((n, max(dmax, 1), 3), np.uint8) if want_traj else None
    p = lambda a: a.ctypes.data if a is not None else None  # noqa: E731
    res = {"traj": traj}
    if prefix == "orc":
        payoff = np.zeros(n, np.uint8)
        th = np.zeros(n, np.uint32)
        sums = np.zeros(8, np.uint64)
        lib.orc_set_engine_kind(1 if engine == "xs" else 0)
        try:
            steps = lib.orc_mold_rollouts(n, n_actions, p(dp), dp_u, p(ms), model_seed, model_pre_discard, p(payoff),
                                          p(traj), max(dmax, 1), p(th), p(sums))
        finally:
            lib.orc_set_engine_kind(0)
        res.update(payoff=payoff, traj_hash=th, sums=sums)
    else:
        prow = np.zeros(n, np.float64)
        pcol = np.zeros(n, np.float64)
        fn = lib.ref_xs_mold_rollouts if engine == "xs" else lib.ref_mold_rollouts
        steps = fn(n, n_actions, p(dp), dp_u, p(ms), model_seed, model_pre_discard, p(prow), p(pcol),
                   p(traj), max(dmax, 1))
        res.update(payoff_row=prow, payoff_col=pcol)
    res["steps"] = steps
    return res
