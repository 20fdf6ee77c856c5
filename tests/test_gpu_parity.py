"""GPU parity: libpinyon_cuda (through its C ABI) against the oracle on the same seeded
inputs — payoffs, (row, col, chance) trajectories, trajectory hashes, exact win/draw/loss
sums and the post-batch model engine, all bit-exact."""
import numpy as np
import pytest

from oracle import loader
from pinyon_b200 import Context, Engine, PcuError
from pinyon_b200.capi import unpack2

pytestmark = pytest.mark.gpu


def run_tree(ctx, n, *, rows, cols, T=1, thr=(0, 1), depth, state_seed=0, pre=0, tseed=0, bias=0, engine=None,
             model_seeds=None, want_traj=True):
    gen = 2 * rows * cols * T
    pos = (pre + gen) if np.isscalar(pre) else (pre.astype(np.uint32) + np.uint32(gen))
    dmax = int(depth) if np.isscalar(depth) else int(depth.max())
    out = dict(payoff=np.full(n, 255, np.uint8), traj_hash=np.zeros(n, np.uint32), sums=np.zeros(8, np.uint64))
    out["traj"] = np.full((n, max(dmax, 1), 3), 255, np.uint8) if want_traj else None
    d = depth if np.isscalar(depth) else depth.astype(np.uint16)
    ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=d, state_seed=state_seed,
                            state_pos=pos, transition_seed=tseed, payoff_bias=bias, engine=engine,
                            model_seeds=model_seeds, payoff=out["payoff"], traj=out["traj"], traj_stride=max(dmax, 1),
                            traj_hash=out["traj_hash"], sums=out["sums"])
    return out


def check_same(got, exp, n):
    assert np.array_equal(got["payoff"], exp["payoff"])
    if got.get("traj") is not None:
        assert np.array_equal(got["traj"], exp["traj"])
    assert np.array_equal(got["traj_hash"], exp["traj_hash"])
    assert np.array_equal(got["sums"], exp["sums"])
    assert int(got["sums"][:3].sum()) == n


def engine_matches(engine, seed, discarded, oracle):
    exp = np.zeros(700, np.uint32)
    oracle.orc_mt_words(seed, discarded, 700, exp.ctypes.data)
    got = np.array([engine.next() for _ in range(700)], np.uint32)
    return np.array_equal(got, exp)


CASES = [
    # rows, cols, T, thr, depth, state_seed, pre, tseed, bias, model_seed, model_pre
    (3, 3, 1, (0, 1), 10, 0, 0, 0, 0, 0, 0),  # BASELINE config 1/2
    (9, 9, 3, (0, 1), 20, 0, 0, 0, 0, 0, 0),  # BASELINE config 3
    (2, 2, 2, (1, 2), 5, 7, 13, 1, 1, 99, 7),
    (4, 7, 4, (3, 10), 9, 12345678901234, 700, 0xDEADBEEFCAFE, -2, 2**32 + 5, 1001),
    (1, 1, 1, (0, 1), 1, 5, 0, 0, 0, 1, 623),
    (5, 2, 8, (1, 1), 3, 99, 0, 3, 0, 3, 624),
    (16, 17, 1, (0, 1), 6, 3, 0, 0, 0, 3, 2),  # rows*cols > 256 -> 16-bit action codes
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("n", [1, 223, 1000])
def test_tree_shared_root_stream(ctx, oracle, case, n):
    rows, cols, T, thr, depth, sseed, pre, tseed, bias, mseed, mpre = case
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth,
                                      state_seed=sseed, state_pre_discard=pre, transition_seed=tseed, payoff_bias=bias,
                                      model_seed=mseed, model_pre_discard=mpre)
    eng = Engine(mseed).discard(mpre)
    ctx.set_option("small_path", 0)  # n = 1 through the fused kernel too; the one-launch path has its own test
    try:
        got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=depth, state_seed=sseed, pre=pre, tseed=tseed,
                       bias=bias, engine=eng)
    finally:
        ctx.set_option("small_path", 1)
    assert ctx.last_path == "stream_table"
    check_same(got, exp, n)
    assert engine_matches(eng, mseed, mpre + 4 * depth * n, oracle)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("n", [1, 2, 32])
def test_tree_small_batches_one_launch(ctx, oracle, case, n):
    """<= 32 rollouts from host memory in mode S (what TreeBandit's single-leaf inference does,
    monte-carlo-model.hh:41-56): ONE k_small launch, engine and state fields in the launch parameters."""
    rows, cols, T, thr, depth, sseed, pre, tseed, bias, mseed, mpre = case
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth,
                                      state_seed=sseed, state_pre_discard=pre, transition_seed=tseed, payoff_bias=bias,
                                      model_seed=mseed, model_pre_discard=mpre)
    eng = Engine(mseed).discard(mpre)
    before = ctx.launches
    got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=depth, state_seed=sseed, pre=pre, tseed=tseed,
                   bias=bias, engine=eng)
    assert ctx.last_path == "small_tree" and ctx.launches == before + 1
    check_same(got, exp, n)
    assert engine_matches(eng, mseed, mpre + 4 * depth * n, oracle)


def test_tree_small_batch_private_ragged(ctx, oracle):
    """The one-launch path with everything per rollout: seeds, positions, depths, biases, T > 1."""
    rng = np.random.default_rng(77)
    n = 29
    kw = dict(state_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              state_pre_discard=rng.integers(0, 3000, n).astype(np.uint32),
              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              depth=rng.integers(1, 40, n).astype(np.uint32),
              payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
    exp = loader.random_tree_rollouts(oracle, n, rows=5, cols=4, transitions=3, thr=(3, 10), model_seed=8,
                                      model_pre_discard=600, **kw)
    eng = Engine(8).discard(600)
    got = run_tree(ctx, n, rows=5, cols=4, T=3, thr=(3, 10), depth=kw["depth"], state_seed=kw["state_seed"],
                   pre=kw["state_pre_discard"], tseed=kw["transition_seed"], bias=kw["payoff_bias"], engine=eng)
    assert ctx.last_path == "small_tree"
    check_same(got, exp, n)
    assert engine_matches(eng, 8, 600 + 4 * int(kw["depth"].sum()), oracle)


def test_tree_known_answers_appendix_b(ctx):
    """vectors produced by the reference's own MonteCarloModel."""
    b1 = [1, 0, 1, .5, 0, 1, 0, 1, 0, 1, 1, 1, 0, 1, .5, 1, 0, .5, 1, .5, 1, 1, 1, 1, 1, 1, .5, .5, 0, 1, 0, 0]
    got = run_tree(ctx, 32, rows=3, cols=3, depth=10, engine=Engine(0))
    assert (got["payoff"] / 2.0).tolist() == b1
    assert got["traj"][0].tolist() == [[2, 2, 0], [1, 2, 0], [1, 1, 0], [0, 0, 0], [0, 2, 0], [1, 1, 0], [1, 0, 0],
                                       [0, 1, 0], [2, 1, 0], [0, 0, 0]]
    b2 = [0, 1, .5, .5, 1, 0, .5, 1, 1, 1, 0, 0, 0, 0, .5, 0, 1, 0, 0, 1, 1, 1, 1, 1, 1, .5, 0, 1, 0, 1, 0, 0]
    got = run_tree(ctx, 32, rows=9, cols=9, T=3, depth=20, engine=Engine(0))
    assert (got["payoff"] / 2.0).tolist() == b2
    assert got["traj"][0][:4].tolist() == [[5, 5, 1], [7, 2, 1], [4, 7, 0], [0, 6, 1]]
    # 10^6 rollouts of B1: wins 364521, draws 159742, losses 475737
    sums = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(10**6, rows=3, cols=3, depth=10, engine=Engine(0), sums=sums)
    assert sums[:4].tolist() == [364521, 159742, 475737, 10**7]
    sums = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(20000, rows=9, cols=9, transitions=3, depth=20, engine=Engine(0), sums=sums)
    assert sums[:3].tolist() == [8949, 2124, 8927]


def test_tree_shared_root_ragged_positions(ctx, oracle):
    """One state stream, per-rollout positions and biases (what search leaves look like)."""
    rng = np.random.default_rng(3)
    n = 777
    pre = rng.integers(0, 400, n).astype(np.uint32)
    bias = rng.integers(-3, 4, n).astype(np.int32)
    for rows, cols, T, thr in [(3, 3, 1, (0, 1)), (4, 3, 3, (3, 10))]:
        exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=8,
                                          state_seed=11, state_pre_discard=pre, transition_seed=5, payoff_bias=bias,
                                          model_seed=4)
        eng = Engine(4)
        got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=8, state_seed=11, pre=pre, tseed=5, bias=bias,
                       engine=eng)
        assert ctx.last_path == "stream_table"
        check_same(got, exp, n)


@pytest.mark.parametrize("mode", ["S", "P"])
@pytest.mark.parametrize("shape", [(3, 3, 1, (0, 1)), (9, 9, 3, (0, 1)), (5, 4, 3, (3, 10)), (2, 6, 2, (1, 2))])
def test_tree_private_roots(ctx, oracle, mode, shape):
    """Distinct roots: private state engines, ragged depth, per-rollout transition seeds."""
    rows, cols, T, thr = shape
    rng = np.random.default_rng(rows * 100 + cols)
    n = 500
    kw = dict(state_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              state_pre_discard=rng.integers(0, 1500, n).astype(np.uint32),
              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              depth=rng.integers(1, 14, n).astype(np.uint32),
              payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64) if mode == "P" else None
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, model_seed=21,
                                      model_pre_discard=5, model_seeds=mseeds, **kw)
    eng = Engine(21).discard(5)
    got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=kw["depth"], state_seed=kw["state_seed"],
                   pre=kw["state_pre_discard"], tseed=kw["transition_seed"], bias=kw["payoff_bias"],
                   engine=None if mode == "P" else eng, model_seeds=mseeds)
    assert ctx.last_path == "warp_private_" + mode  # draw positions beyond the first refill, or T > 1
    check_same(got, exp, n)
    if mode == "S":
        assert engine_matches(eng, 21, 5 + 4 * int(kw["depth"].sum()), oracle)


@pytest.mark.parametrize("mode", ["S", "P"])
@pytest.mark.parametrize("shape", [(3, 3, 10, 150), (2, 5, 12, 60), (4, 4, 7, 250), (1, 1, 40, 0), (8, 9, 2, 20)])
def test_tree_private_roots_lane(ctx, oracle, mode, shape):
    """Distinct roots with one chance transition: the lane-per-rollout kernel,
    ragged depths and positions, against the oracle and against the warp kernel."""
    rows, cols, dmax, premax = shape
    rng = np.random.default_rng(rows * 1000 + cols + dmax)
    n = 777
    kw = dict(state_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              state_pre_discard=rng.integers(0, premax + 1, n).astype(np.uint32),
              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              depth=rng.integers(1, dmax + 1, n).astype(np.uint32),
              payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64) if mode == "P" else None
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=1, model_seed=4,
                                      model_pre_discard=9, model_seeds=mseeds, **kw)
    for force_warp, full in ((0, 0), (0, 1), (1, 0)):
        eng = Engine(4).discard(9)
        ctx.set_option("force_warp", force_warp)
        ctx.set_option("lane_full_columns", full)  # every row kept vs the ring of live rows
        try:
            got = run_tree(ctx, n, rows=rows, cols=cols, T=1, depth=kw["depth"], state_seed=kw["state_seed"],
                           pre=kw["state_pre_discard"], tseed=kw["transition_seed"], bias=kw["payoff_bias"],
                           engine=None if mode == "P" else eng, model_seeds=mseeds)
        finally:
            ctx.set_option("force_warp", 0)
            ctx.set_option("lane_full_columns", 0)
        assert ctx.last_path == ("warp_private_" if force_warp else "lane_private_") + mode
        check_same(got, exp, n)
        if mode == "S":
            assert engine_matches(eng, 4, 9 + 4 * int(kw["depth"].sum()), oracle)


@pytest.mark.parametrize("extra,path", [(0, "lane_private_S"), (1, "warp_private_S")])
def test_tree_lane_boundary(ctx, oracle, extra, path):
    """The lane kernel serves draw indices up to 622 (all of the first refill but its last word, whose
    twist needs the refilled X[0]); one more pre-discarded draw sends the batch to the warp kernel."""
    rows, cols, depth = 3, 3, 10
    gen = 2 * rows * cols
    pre = 623 - depth * (gen + rows * cols - 1 + 2) + extra  # bound = pre + depth * per_step
    n = 300
    seeds = np.arange(n, dtype=np.uint64) + np.uint64(2**32 - 100)  # seeds wrap mod 2^32 (random.hh:16)
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=1, depth=depth, model_seed=1,
                                      state_seed=seeds, state_pre_discard=pre)
    got = run_tree(ctx, n, rows=rows, cols=cols, depth=depth, state_seed=seeds, pre=pre, engine=Engine(1))
    assert ctx.last_path == path
    check_same(got, exp, n)


@pytest.mark.parametrize("case", [
    dict(rows=9, cols=9, T=3, thr=(0, 1), n=148 * 96 + 77, dmax=20),   # config 3 with distinct roots: every ring reused
    dict(rows=4, cols=3, T=2, thr=(3, 10), n=5000, dmax=40),            # thresholds; depth > 32: two code fetches
    dict(rows=3, cols=3, T=1, thr=(0, 1), n=3000, dmax=70),             # one transition, draws far past the first refill
    dict(rows=2, cols=2, T=8, thr=(1, 1), n=45, dmax=9),                # a partial second seeder pass
])
def test_warp2_against_oracle_and_k_warp(ctx, oracle, case):
    """k_warp2 (engines seeded and played in shared memory) against the oracle and against k_seed_pre + k_warp
    (option warp2 = 0) on ragged depths, per-rollout positions, biases, state and transition seeds."""
    rows, cols, T, thr, n, dmax = (case[k] for k in ("rows", "cols", "T", "thr", "n", "dmax"))
    rng = np.random.default_rng(n + T)
    depth = rng.integers(1, dmax + 1, n).astype(np.uint32)
    depth[:3] = dmax
    pre = rng.integers(0, 3000, n).astype(np.uint32)
    bias = rng.integers(-2, 3, n).astype(np.int32)
    seeds = rng.integers(0, 2**63, n).astype(np.uint64)
    tseeds = rng.integers(0, 2**63, n).astype(np.uint64)
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, model_seed=6,
                                      state_seed=seeds, state_pre_discard=pre, transition_seed=tseeds, payoff_bias=bias)
    results = []
    for warp2, lock in ((1, 4), (1, 2), (1, 1), (0, 4)):  # lock: rollouts a player warp steps in lock-step
        ctx.set_option("warp2", warp2)
        ctx.set_option("warp2_lock", lock)
        ctx.set_option("force_warp", 1)
        try:
            eng = Engine(6)
            got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=depth, state_seed=seeds, pre=pre, tseed=tseeds,
                           bias=bias, engine=eng)
        finally:
            ctx.set_option("warp2", 1)
            ctx.set_option("warp2_lock", 1)
            ctx.set_option("force_warp", 0)
        assert ctx.last_path == "warp_private_S"
        check_same(got, exp, n)
        assert engine_matches(eng, 6, 4 * int(depth.sum()), oracle)
        results.append(got)
    assert all(np.array_equal(results[0]["traj"], r["traj"]) for r in results[1:])


def test_tree_general_path_equals_fused(ctx, oracle):
    """force_general routes a shared-root batch through the warp kernel: same answers."""
    n = 600
    exp = loader.random_tree_rollouts(oracle, n, rows=9, cols=9, transitions=3, depth=20, model_seed=0)
    ctx.set_option("force_general", 1)
    try:
        got = run_tree(ctx, n, rows=9, cols=9, T=3, depth=20, engine=Engine(0))
        assert ctx.last_path == "warp_private_S"
    finally:
        ctx.set_option("force_general", 0)
    check_same(got, exp, n)


@pytest.mark.parametrize("n_actions,depth", [(2, 10), (2, 50), (3, 1), (7, 4), (17, 3)])
@pytest.mark.parametrize("n", [1, 500])
def test_mold_stream(ctx, oracle, n_actions, depth, n):
    exp = loader.mold_rollouts(oracle, n, n_actions=n_actions, depth=depth, model_seed=0, model_pre_discard=3)
    eng = Engine(0).discard(3)
    payoff = np.full(n, 255, np.uint8)
    traj = np.full((n, depth, 3), 255, np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    ctx.rollout_mold(n, n_actions=n_actions, depth=depth, engine=eng, payoff=payoff, traj=traj, traj_stride=depth,
                     traj_hash=th, sums=sums)
    assert ctx.last_path == ("small_mold" if n <= 32 else "stream_mold")
    assert not payoff.any()  # MoldState never sets a payoff: (0, 0)
    assert np.array_equal(traj, exp["traj"])
    assert np.array_equal(th, exp["traj_hash"])
    assert np.array_equal(sums, exp["sums"])
    assert engine_matches(eng, 0, 3 + 4 * depth * n, oracle)


@pytest.mark.parametrize("mode", ["S", "P"])
def test_mold_ragged(ctx, oracle, mode):
    rng = np.random.default_rng(8)
    n = 400
    depth = rng.integers(0, 30, n).astype(np.uint32)  # depth 0 is a terminal root: no steps
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64) if mode == "P" else None
    exp = loader.mold_rollouts(oracle, n, n_actions=3, depth=depth, model_seed=6, model_seeds=mseeds)
    eng = Engine(6)
    traj = np.zeros((n, int(depth.max()), 3), np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    ctx.rollout_mold(n, n_actions=3, depth=depth.astype(np.uint16), engine=None if mode == "P" else eng,
                     model_seeds=mseeds, traj=traj, traj_stride=int(depth.max()), traj_hash=th, sums=sums)
    assert ctx.last_path == ("lane_mold_P" if mode == "P" else "warp_mold_S")
    assert np.array_equal(traj, exp["traj"])
    assert np.array_equal(th, exp["traj_hash"])
    assert np.array_equal(sums, exp["sums"])


def test_jump_ahead_checkpoints_match_host_stepping(ctx, oracle):
    """Many CTAs, each starting mid-stream: the device jump-ahead against plain stepping on
    the host, and both against the oracle."""
    n = 224 * 1300 + 57  # > 592 rounds -> several rounds per CTA, hundreds of checkpoints
    outs = []
    for host in (1, 0):
        ctx.set_option("host_checkpoints", host)
        eng = Engine(0)
        got = run_tree(ctx, n, rows=3, cols=3, depth=10, engine=eng, want_traj=False)
        outs.append((got, [eng.next() for _ in range(8)]))
    ctx.set_option("host_checkpoints", 0)
    assert np.array_equal(outs[0][0]["payoff"], outs[1][0]["payoff"])
    assert np.array_equal(outs[0][0]["traj_hash"], outs[1][0]["traj_hash"])
    assert outs[0][1] == outs[1][1]
    exp = loader.random_tree_rollouts(oracle, n, rows=3, cols=3, depth=10, want_traj=False)
    check_same(outs[1][0], exp, n)


def test_batch_continuation(ctx):
    """Mode S is one stream: a batch run in two calls equals the same batch in one."""
    n = 50000
    eng = Engine(9)
    whole = run_tree(ctx, n, rows=9, cols=9, T=3, depth=20, engine=eng, want_traj=False)
    eng2 = Engine(9)
    a = run_tree(ctx, 12345, rows=9, cols=9, T=3, depth=20, engine=eng2, want_traj=False)
    b = run_tree(ctx, n - 12345, rows=9, cols=9, T=3, depth=20, engine=eng2, want_traj=False)
    assert np.array_equal(np.concatenate([a["payoff"], b["payoff"]]), whole["payoff"])
    assert np.array_equal(np.concatenate([a["traj_hash"], b["traj_hash"]]), whole["traj_hash"])
    assert np.array_equal(a["sums"] + b["sums"], whole["sums"])
    assert [eng.next() for _ in range(4)] == [eng2.next() for _ in range(4)]


def test_device_resident_batch(ctx, oracle):
    torch = pytest.importorskip("torch")
    n = 3000
    exp = loader.random_tree_rollouts(oracle, n, rows=9, cols=9, transitions=3, depth=20, want_traj=False)
    payoff = torch.zeros(n, dtype=torch.uint8, device="cuda")
    th = torch.zeros(n, dtype=torch.int32, device="cuda")
    sums = torch.zeros(8, dtype=torch.int64, device="cuda")
    ctx.rollout_random_tree(n, rows=9, cols=9, transitions=3, depth=20, engine=Engine(0), payoff=payoff, traj_hash=th,
                            sums=sums)
    torch.cuda.synchronize()
    assert np.array_equal(payoff.cpu().numpy(), exp["payoff"])
    assert np.array_equal(th.cpu().numpy().view(np.uint32), exp["traj_hash"])
    assert np.array_equal(sums.cpu().numpy().view(np.uint64), exp["sums"])
    # per-rollout device arrays go through the general path
    seeds = torch.arange(n, dtype=torch.int64, device="cuda")
    exp = loader.random_tree_rollouts(oracle, n, rows=3, cols=3, depth=10, state_seed=np.arange(n, dtype=np.uint64),
                                      want_traj=False)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, state_seed=seeds, engine=Engine(0), payoff=payoff)
    torch.cuda.synchronize()
    assert np.array_equal(payoff.cpu().numpy(), exp["payoff"])


def test_empty_and_invalid(ctx):
    sums = np.ones(8, np.uint64)
    eng = Engine(1)
    before = eng.x.copy()
    ctx.rollout_random_tree(0, rows=3, cols=3, depth=10, engine=eng, sums=sums)
    assert not sums.any() and np.array_equal(eng.x, before)
    for bad in (dict(rows=3, cols=3, depth=0), dict(rows=0, cols=3, depth=1), dict(rows=3, cols=3, depth=1, transitions=0),
                dict(rows=3, cols=3, depth=2, state_pos=3)):
        with pytest.raises(PcuError) as e:
            ctx.rollout_random_tree(4, engine=Engine(0), **bad)
        assert e.value.code == -1
    with pytest.raises(PcuError):
        ctx.rollout_mold(4, n_actions=0, depth=3, engine=Engine(0))


def test_full_size_properties(ctx, oracle):
    """BASELINE config 2 size (2^24 rollouts, 3x3x1, depth 10): exact counts, sampled windows
    against the oracle, and the post-batch engine against the host jump-ahead."""
    n = 1 << 24
    eng = Engine(0)
    payoff = np.zeros(n, np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=eng, payoff=payoff, traj_hash=th, sums=sums)
    assert int(sums[:3].sum()) == n and int(sums[3]) == 10 * n
    counts = np.bincount(payoff, minlength=3)
    assert counts.tolist() == [int(sums[2]), int(sums[1]), int(sums[0])]
    assert int(th.astype(np.uint64).sum()) == int(sums[4])
    assert payoff[:10**6].tolist().count(2) == 364521  # Appendix B1 prefix
    for start in (0, 5_000_000, n - 4096):
        exp = loader.random_tree_rollouts(oracle, 4096, rows=3, cols=3, depth=10, model_pre_discard=40 * start,
                                          want_traj=False)
        assert np.array_equal(payoff[start:start + 4096], exp["payoff"])
        assert np.array_equal(th[start:start + 4096], exp["traj_hash"])
    ref = Engine(0).discard(40 * n)
    assert [eng.next() for _ in range(700)] == [ref.next() for _ in range(700)]


def test_zero_copy_payoffs_equal_staged(ctx):
    """Batches >= 2^20 into PINNED host buffers are written by the kernels straight over PCIe; pageable
    buffers take the staged copy.  Same bytes either way, and the option switches it off."""
    n = (1 << 21) + 12345
    staged = np.zeros(n, np.uint8)
    s0 = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(3), payoff=staged, sums=s0)
    pinned = ctx.pinned(n, np.uint8)
    for zero_copy in (1, 0):
        pinned[:] = 255
        s1 = np.zeros(8, np.uint64)
        ctx.set_option("zero_copy", zero_copy)
        try:
            ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(3), payoff=pinned, sums=s1)
        finally:
            ctx.set_option("zero_copy", 1)
        assert np.array_equal(pinned, staged)
        assert np.array_equal(s0, s1)


def test_distinct_roots_full_size_windows(ctx, oracle):
    """SURVEY 8(d) M1-DR at 2^22: state_seed_i = i; exact counts and sampled windows against the oracle."""
    n = 1 << 22
    seeds = np.arange(n, dtype=np.uint64)
    payoff = np.zeros(n, np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    eng = Engine(0)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, state_seed=seeds, engine=eng, payoff=payoff, traj_hash=th,
                            sums=sums)
    assert ctx.last_path == "lane_private_S"
    assert int(sums[:3].sum()) == n and int(sums[3]) == 10 * n
    assert np.bincount(payoff, minlength=3).tolist() == [int(sums[2]), int(sums[1]), int(sums[0])]
    assert int(th.astype(np.uint64).sum()) == int(sums[4])
    for start in (0, 1_234_567, n - 2048):
        exp = loader.random_tree_rollouts(oracle, 2048, rows=3, cols=3, depth=10, model_pre_discard=40 * start,
                                          state_seed=seeds[start:start + 2048].copy(), want_traj=False)
        assert np.array_equal(payoff[start:start + 2048], exp["payoff"])
        assert np.array_equal(th[start:start + 2048], exp["traj_hash"])
    ref = Engine(0).discard(40 * n)
    assert [eng.next() for _ in range(700)] == [ref.next() for _ in range(700)]


def test_distinct_roots_config3_full_size_windows(ctx, oracle):
    """SURVEY 8(d) M2-DR at the bench's size (2^20 rollouts of config 3, state_seed_i = i, transition_seed_i = i * phi):
    k_warp2 with every ring reused ~90 times per CTA; exact counts and sampled windows against the oracle."""
    n = 1 << 20
    seeds = np.arange(n, dtype=np.uint64)
    tseeds = seeds * np.uint64(0x9E3779B97F4A7C15)
    payoff = np.zeros(n, np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    eng = Engine(0)
    ctx.rollout_random_tree(n, rows=9, cols=9, transitions=3, depth=20, state_seed=seeds, transition_seed=tseeds, engine=eng,
                            payoff=payoff, traj_hash=th, sums=sums)
    assert ctx.last_path == "warp_private_S"
    assert int(sums[:3].sum()) == n and int(sums[3]) == 20 * n
    assert np.bincount(payoff, minlength=3).tolist() == [int(sums[2]), int(sums[1]), int(sums[0])]
    assert int(th.astype(np.uint64).sum()) == int(sums[4])
    per_cta = -(-n // 148)  # k_warp2 gives each CTA a contiguous share: sample across a share boundary too
    for start in (0, per_cta - 256, 543_210, n - 512):
        exp = loader.random_tree_rollouts(oracle, 512, rows=9, cols=9, transitions=3, depth=20, model_pre_discard=80 * start,
                                          state_seed=seeds[start:start + 512].copy(),
                                          transition_seed=tseeds[start:start + 512].copy(), want_traj=False)
        assert np.array_equal(payoff[start:start + 512], exp["payoff"])
        assert np.array_equal(th[start:start + 512], exp["traj_hash"])
    ref = Engine(0).discard(80 * n)
    assert [eng.next() for _ in range(700)] == [ref.next() for _ in range(700)]


@pytest.mark.parametrize("shape", [
    (255, 255, 1, (0, 1), 3),    # largest action sets: joint codes need 16 bits, one step is 195 076 state draws
    (2, 2, 8, (0, 1), 12),       # most chance transitions before the reference's Rational<int> sum overflows int
    (3, 3, 8, (1, 1), 8),        # threshold 1: only n_k = 10 survives, frequent all-zero rows -> n_0 = 1 fallback
    (1, 1, 1, (0, 1), 300),      # one action each: the model still draws four words per step
])
@pytest.mark.parametrize("distinct", [False, True])
def test_tree_extreme_shapes(ctx, oracle, shape, distinct):
    rows, cols, T, thr, depth = shape
    n = 33  # just above the one-launch path
    kw = {}
    if distinct:
        kw = dict(state_seed=np.arange(n, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15),
                  transition_seed=np.arange(n, dtype=np.uint64) + np.uint64(2**40))
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, model_seed=11,
                                      **kw)
    eng = Engine(11)
    got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=depth, engine=eng,
                   state_seed=kw.get("state_seed", 0), tseed=kw.get("transition_seed", 0))
    check_same(got, exp, n)
    assert engine_matches(eng, 11, 4 * depth * n, oracle)


def test_mold_deep(ctx, oracle):
    """MoldState far deeper than one engine refill per rollout (tests/cart/node.cc uses depth 50)."""
    n, depth = 40, 3000
    exp = loader.mold_rollouts(oracle, n, n_actions=5, depth=depth, model_seed=2, want_traj=False)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    eng = Engine(2)
    ctx.rollout_mold(n, n_actions=5, depth=depth, engine=eng, traj_hash=th, sums=sums)
    assert np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])
    assert engine_matches(eng, 2, 4 * depth * n, oracle)


def test_nine_transitions_are_refused(ctx):
    """random-tree.hh:165-184 sums n_k/10 as Rational<int> over 10^T: the numerator overflows int from T = 9 on
    (10^8 * sum(n_k) >= 2^31 when sum >= 22), so the reference's results are undefined behaviour; the library says so."""
    with pytest.raises(PcuError) as e:
        ctx.rollout_random_tree(64, rows=2, cols=2, transitions=9, depth=3, engine=Engine(0), payoff=np.zeros(64, np.uint8))
    assert e.value.code == -3


# ---------------------------------------------------------------------------------------------
# Fused-kernel geometry: every streams_per_sm x rb the library can run (the auto setting picks 8 streams per SM from
# 2.5e8 rollout-steps per call: every BASELINE configs[2..3] headline), with and without trajectories (the
# trajectory-free consumer interleaves the walks of a lane's rollouts and stores vectors).


_GEOMETRY_EXPECTED = {}


@pytest.mark.parametrize("case", [CASES[0], CASES[1], CASES[3], CASES[6]])
@pytest.mark.parametrize("per_sm", [1, 4, 8, 16])
@pytest.mark.parametrize("rb", [32, 64, 128])
def test_stream_geometry_matrix(ctx, oracle, case, per_sm, rb):
    rows, cols, T, thr, depth, sseed, pre, tseed, bias, mseed, mpre = case
    n = 20011 if rows * cols * T > 100 else 70037  # ragged tail: the last stream and its last batch are partial
    if case not in _GEOMETRY_EXPECTED:  # one oracle run per case serves the twelve geometries
        _GEOMETRY_EXPECTED[case] = loader.random_tree_rollouts(
            oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, state_seed=sseed, state_pre_discard=pre,
            transition_seed=tseed, payoff_bias=bias, model_seed=mseed, model_pre_discard=mpre)
    exp = _GEOMETRY_EXPECTED[case]
    ctx.set_option("streams_per_sm", per_sm)
    ctx.set_option("rb", rb)
    try:
        for want_traj in (False, True):
            eng = Engine(mseed).discard(mpre)
            got = run_tree(ctx, n, rows=rows, cols=cols, T=T, thr=thr, depth=depth, state_seed=sseed, pre=pre, tseed=tseed,
                           bias=bias, engine=eng, want_traj=want_traj)
            assert ctx.last_path == "stream_table"
            assert ctx.stat("last_streams") >= min(148 * per_sm, n // 128) * 0.5 and ctx.stat("last_chunk") % 128 == 0
            check_same(got, exp, n)
            assert engine_matches(eng, mseed, mpre + 4 * depth * n, oracle)
    finally:
        ctx.set_option("streams_per_sm", 0)
        ctx.set_option("rb", 0)


@pytest.mark.parametrize("per_sm", [1, 8, 16])
@pytest.mark.parametrize("rb", [32, 64, 128])
def test_mold_geometry_matrix(ctx, oracle, per_sm, rb):
    n, depth = 90001, 10
    exp = loader.mold_rollouts(oracle, n, n_actions=2, depth=depth, model_seed=3, want_traj=False)
    ctx.set_option("streams_per_sm", per_sm)
    ctx.set_option("rb", rb)
    try:
        eng = Engine(3)
        th = np.zeros(n, np.uint32)
        sums = np.zeros(8, np.uint64)
        ctx.rollout_mold(n, n_actions=2, depth=depth, engine=eng, traj_hash=th, sums=sums)
        assert ctx.last_path == "stream_mold"
    finally:
        ctx.set_option("streams_per_sm", 0)
        ctx.set_option("rb", 0)
    assert np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])
    assert engine_matches(eng, 3, 4 * depth * n, oracle)


def test_packed_payoffs_every_path(ctx, oracle):
    """pcu_out.payoff_packed2: 2 bits per rollout, written by the fused kernel itself (one byte per lane) and by a
    packing pass behind the general kernels; alone and next to the byte output."""
    rng = np.random.default_rng(5)
    for n in (33, 1001, 70003):
        seeds = rng.integers(0, 2**63, n).astype(np.uint64)
        cases = [
            ("stream_table", dict(rows=3, cols=3, transitions=1, depth=10), {}),
            ("stream_table", dict(rows=9, cols=9, transitions=3, depth=20), {}),
            ("lane_private_S", dict(rows=3, cols=3, transitions=1, depth=10), dict(state_seed=seeds)),
            ("warp_private_S", dict(rows=4, cols=3, transitions=2, depth=6), dict(state_seed=seeds, transition_seed=seeds)),
        ]
        for path, shape, extra in cases:
            if path == "warp_private_S" and n > 2000:
                continue
            exp = loader.random_tree_rollouts(oracle, n, model_seed=2, want_traj=False, **shape, **extra)
            for with_bytes in (False, True):
                packed = np.full((n + 3) // 4, 255, np.uint8)
                pay = np.full(n, 255, np.uint8) if with_bytes else None
                sums = np.zeros(8, np.uint64)
                ctx.rollout_random_tree(n, engine=Engine(2), payoff=pay, payoff_packed=packed, sums=sums, **shape, **extra)
                assert ctx.last_path == path
                assert np.array_equal(unpack2(packed, n), exp["payoff"]), (path, n, with_bytes)
                if n % 4:
                    assert packed[-1] >> (2 * (n % 4)) == 0  # unused bit pairs of the last byte are zero
                if with_bytes:
                    assert np.array_equal(pay, exp["payoff"])
                assert np.array_equal(sums, exp["sums"])
    packed = np.full(26, 255, np.uint8)
    ctx.rollout_mold(101, n_actions=3, depth=4, engine=Engine(0), payoff_packed=packed)
    assert not packed.any()


def test_packed_payoffs_zero_copy(ctx, oracle):
    """>= 2^20 rollouts into pinned memory: the packed bytes are stored straight over PCIe by the consumer warps."""
    n = (1 << 21) + 777
    ref = np.zeros(n, np.uint8)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(7), payoff=ref)
    packed = ctx.pinned((n + 3) // 4, np.uint8)
    packed[:] = 255
    sums = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(n, rows=3, cols=3, depth=10, engine=Engine(7), payoff_packed=packed, sums=sums)
    assert np.array_equal(unpack2(packed, n), ref)
    assert np.bincount(ref, minlength=3).tolist() == [int(sums[2]), int(sums[1]), int(sums[0])]


def test_trajectory_stride_shorter_than_depth(ctx, oracle):
    """A per-rollout depth array with depth[i] > traj_stride: the extra steps are played and hashed but not recorded
    (the oracle clamps the same way); nothing is written past a rollout's traj_stride slots."""
    rng = np.random.default_rng(11)
    for n, kw, path in ((20, dict(), "small_tree"),
                        (500, dict(state_seed=rng.integers(0, 2**63, 500).astype(np.uint64)), "lane_private_S"),
                        (500, dict(state_seed=rng.integers(0, 2**63, 500).astype(np.uint64), T=2), "warp_private_S")):
        T = kw.pop("T", 1)
        depth = rng.integers(1, 9, n).astype(np.uint32)
        stride = 4
        exp = loader.random_tree_rollouts(oracle, n, rows=3, cols=3, transitions=T, depth=depth, model_seed=6, **kw)
        exp_traj = exp["traj"][:, :stride, :]
        traj = np.full((n + 1, stride, 3), 77, np.uint8)  # one guard row behind the buffer
        th = np.zeros(n, np.uint32)
        ctx.rollout_random_tree(n, rows=3, cols=3, transitions=T, depth=depth.astype(np.uint16), engine=Engine(6), traj=traj,
                                traj_stride=stride, traj_hash=th, **kw)
        assert ctx.last_path == path
        assert np.array_equal(traj[:n], exp_traj), path
        assert (traj[n] == 77).all()
        assert np.array_equal(th, exp["traj_hash"])
    depth = rng.integers(1, 9, 300).astype(np.uint32)
    exp = loader.mold_rollouts(oracle, 300, n_actions=3, depth=depth, model_seed=6)
    traj = np.full((301, 4, 3), 77, np.uint8)
    ctx.rollout_mold(300, n_actions=3, depth=depth.astype(np.uint16), engine=Engine(6), traj=traj, traj_stride=4)
    assert np.array_equal(traj[:300], exp["traj"][:, :4, :]) and (traj[300] == 77).all()


def test_general_path_rejects_low_positions(ctx):
    """state_pos arrays are range-checked on every path, not only the fused one (k_warp would index its engine below 0)."""
    n = 200
    pos = np.full(n, 2 * 3 * 3 * 2, np.uint32)
    pos[77] = 5
    seeds = np.arange(n, dtype=np.uint64)
    with pytest.raises(PcuError) as e:
        ctx.rollout_random_tree(n, rows=3, cols=3, transitions=2, depth=4, state_seed=seeds, state_pos=pos, engine=Engine(0),
                                payoff=np.zeros(n, np.uint8))
    assert e.value.code == -1
    eng = Engine(0)
    before = [eng.copy().next() for _ in range(3)]
    with pytest.raises(PcuError):
        ctx.rollout_random_tree(n, rows=3, cols=3, transitions=2, depth=4, state_seed=seeds, state_pos=pos, engine=eng,
                                payoff=np.zeros(n, np.uint8))
    assert [eng.copy().next() for _ in range(3)] == before  # a refused call leaves the caller's engine alone


def _windows(n, chunk, count):
    """Window starts that straddle stream-chunk boundaries (first + k * chunk +- a little), plus the two ends."""
    starts = {0, n - count}
    streams = (n + chunk - 1) // chunk
    for k in (1, streams // 3, streams - 1):
        if 0 < k < streams:
            starts.add(max(0, min(n - count, k * chunk - count // 2)))
    return sorted(starts)


@pytest.mark.parametrize("log2n", [24, 26])
def test_full_size_config_b(ctx, oracle, log2n):
    """BASELINE configs[2]: RandomTree depth 20, 9x9 actions, 3 chance transitions — 2^24 (one GPU's bench batch) and
    2^26 (the config as written).  These sizes run 8 model streams per SM (run_stream's auto rule), the geometry every
    config-3 number rides on.  Exact counts, oracle windows straddling stream-chunk boundaries, and the post-batch
    engine against the host jump-ahead."""
    n = 1 << log2n
    depth = 20
    eng = Engine(0)
    payoff = np.zeros(n, np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(n, rows=9, cols=9, transitions=3, depth=depth, engine=eng, payoff=payoff, traj_hash=th, sums=sums)
    assert ctx.last_path == "stream_table"
    chunk, streams = int(ctx.stat("last_chunk")), int(ctx.stat("last_streams"))
    assert streams > 148 * 4  # more than four streams per SM: the 8-per-SM geometry
    assert int(sums[:3].sum()) == n and int(sums[3]) == depth * n
    assert np.bincount(payoff, minlength=3).tolist() == [int(sums[2]), int(sums[1]), int(sums[0])]
    assert int(th.astype(np.uint64).sum()) == int(sums[4])
    assert np.bincount(payoff[:20000], minlength=3).tolist() == [8927, 2124, 8949]  # Appendix B2 prefix
    count = 768
    for start in _windows(n, chunk, count):
        loader.set_model_engine(oracle, Engine(0).discard(4 * depth * start))
        exp = loader.random_tree_rollouts(oracle, count, rows=9, cols=9, transitions=3, depth=depth, want_traj=False)
        assert np.array_equal(payoff[start:start + count], exp["payoff"]), start
        assert np.array_equal(th[start:start + count], exp["traj_hash"]), start
    ref = Engine(0).discard(4 * depth * n)
    assert [eng.next() for _ in range(700)] == [ref.next() for _ in range(700)]


def test_full_size_mold_share_of_config_4(ctx, oracle):
    """BASELINE configs[3]: MoldState{2, 10}, 2^28 rollouts over 8 GPUs = 2^25 per GPU (8 streams per SM)."""
    n, depth = 1 << 25, 10
    eng = Engine(0)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    ctx.rollout_mold(n, n_actions=2, depth=depth, engine=eng, traj_hash=th, sums=sums)
    assert ctx.last_path == "stream_mold"
    chunk, streams = int(ctx.stat("last_chunk")), int(ctx.stat("last_streams"))
    assert streams > 148 * 4
    assert sums[:5].tolist() == [0, n, 0, depth * n, int(th.astype(np.uint64).sum())]
    count = 4096
    for start in _windows(n, chunk, count):
        loader.set_model_engine(oracle, Engine(0).discard(4 * depth * start))
        exp = loader.mold_rollouts(oracle, count, n_actions=2, depth=depth, want_traj=False)
        assert np.array_equal(th[start:start + count], exp["traj_hash"]), start
    ref = Engine(0).discard(4 * depth * n)
    assert [eng.next() for _ in range(700)] == [ref.next() for _ in range(700)]


def test_state_table_is_cached_by_covering_range(ctx, oracle):
    """Leaf batches of one search share the root's state seed but move their position range from call to call: the
    state-stream table is rebuilt only when a batch leaves the range the cached one covers."""
    rng = np.random.default_rng(2)
    n = 4096
    built = []
    for lo, hi in ((0, 200), (50, 150), (0, 200), (100, 260), (30, 250)):
        pre = rng.integers(lo, hi, n).astype(np.uint32)
        exp = loader.random_tree_rollouts(oracle, n, rows=3, cols=3, depth=6, state_seed=9, state_pre_discard=pre,
                                          model_seed=1, want_traj=False)
        got = run_tree(ctx, n, rows=3, cols=3, depth=6, state_seed=9, pre=pre, engine=Engine(1), want_traj=False)
        assert ctx.last_path == "stream_table"
        check_same(got, exp, n)
        built.append(ctx.stat("last_table_build_ms") > 0)
    assert built == [True, False, False, True, False]


# ---------------------------------------------------------------------------------------------
# Mode W: every rollout brings the raw engine words of ITS OWN model (single-leaf inferences of many search trees
# evaluated as one batch).  Reference semantics: Model{prng{s_i}}.inference(state_i), i.e. the oracle's mode P.


def _words_for(mseeds, depth):
    d = np.broadcast_to(np.asarray(depth, np.uint32), mseeds.shape)
    return np.concatenate([Engine(int(s)).take_raw(4 * int(k)) for s, k in zip(mseeds, d)] or [np.zeros(0, np.uint32)])


@pytest.mark.parametrize("shape", [(3, 3, 1, (0, 1)), (4, 3, 3, (3, 10)), (9, 9, 3, (0, 1)), (16, 17, 1, (0, 1))])
@pytest.mark.parametrize("ragged", [False, True])
def test_words_mode_leaves_of_one_root(ctx, oracle, shape, ragged):
    """Leaves of searches from one root: one state stream, per-leaf positions, depths and biases -> k_leaf_table."""
    rows, cols, T, thr = shape
    rng = np.random.default_rng(rows + 10 * cols + 100 * T + ragged)
    n = 4096 + 17
    pre = rng.integers(0, 500, n).astype(np.uint32)
    bias = rng.integers(-3, 4, n).astype(np.int32)
    depth = rng.integers(1, 11, n).astype(np.uint32) if ragged else 7
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64)
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, state_seed=21,
                                      state_pre_discard=pre, transition_seed=4, payoff_bias=bias, model_seeds=mseeds)
    words = _words_for(mseeds, depth)
    dmax = int(np.max(depth))
    out = dict(payoff=np.full(n, 255, np.uint8), traj=np.zeros((n, dmax, 3), np.uint8), traj_hash=np.zeros(n, np.uint32),
               sums=np.zeros(8, np.uint64))
    gen = 2 * rows * cols * T
    ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth.astype(np.uint16) if ragged else depth,
                            state_seed=21, state_pos=pre + np.uint32(gen), transition_seed=4, payoff_bias=bias,
                            model_words=words, traj_stride=dmax, **out)
    assert ctx.last_path == "leaf_table_W"
    check_same(out, exp, n)


@pytest.mark.parametrize("T,path", [(1, "lane_private_W"), (2, "warp_private_W")])
def test_words_mode_private_roots(ctx, oracle, T, path):
    rng = np.random.default_rng(40 + T)
    n = 1500
    seeds = rng.integers(0, 2**63, n).astype(np.uint64)
    depth = rng.integers(1, 9, n).astype(np.uint32)
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64)
    exp = loader.random_tree_rollouts(oracle, n, rows=3, cols=4, transitions=T, depth=depth, state_seed=seeds,
                                      transition_seed=seeds, model_seeds=mseeds)
    out = dict(payoff=np.full(n, 255, np.uint8), traj=np.zeros((n, 8, 3), np.uint8), traj_hash=np.zeros(n, np.uint32),
               sums=np.zeros(8, np.uint64))
    ctx.rollout_random_tree(n, rows=3, cols=4, transitions=T, depth=depth.astype(np.uint16), state_seed=seeds,
                            transition_seed=seeds, model_words=_words_for(mseeds, depth), traj_stride=8, **out)
    assert ctx.last_path == path
    check_same(out, exp, n)


def test_words_mode_mold_and_multi(ctx, mctx, oracle):
    rng = np.random.default_rng(50)
    n = 3000
    depth = rng.integers(0, 12, n).astype(np.uint32)
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64)
    words = _words_for(mseeds, depth)
    exp = loader.mold_rollouts(oracle, n, n_actions=3, depth=depth, model_seeds=mseeds, want_traj=False)
    for c in (ctx, mctx):
        th = np.zeros(n, np.uint32)
        sums = np.zeros(8, np.uint64)
        c.rollout_mold(n, n_actions=3, depth=depth.astype(np.uint16), model_words=words, traj_hash=th, sums=sums)
        assert c.last_path.startswith("leaf_mold_W")
        assert np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])
    d1 = np.maximum(depth, 1)
    words = _words_for(mseeds, d1)
    exp = loader.random_tree_rollouts(oracle, n, rows=3, cols=3, depth=d1, state_seed=8, model_seeds=mseeds, want_traj=False)
    got = dict(payoff=np.zeros(n, np.uint8), traj_hash=np.zeros(n, np.uint32), sums=np.zeros(8, np.uint64))
    mctx.rollout_random_tree(n, rows=3, cols=3, depth=d1.astype(np.uint16), state_seed=8, model_words=words, **got)
    assert mctx.last_path.startswith("leaf_table_W x")
    check_same(got, exp, n)


# ---------------------------------------------------------------------------------------------
# Variants: Pinyon's xor_shift PRNG as the type list's PRNG, and GrowLib::alternating_constant_actions as
# actions_func — the one-rollout-per-thread kernel k_thread.  The oracle's restatement of both is pinned to the compiled
# reference (tests/test_oracle.py::test_variant_golden, test_variants_oracle_vs_reference).


@pytest.mark.parametrize("shape", [(3, 3, 1, (0, 1)), (1, 3, 2, (3, 10)), (3, 1, 1, (0, 1)), (4, 5, 3, (1, 2)), (9, 9, 3, (0, 1))])
@pytest.mark.parametrize("growth", [0, 1])
@pytest.mark.parametrize("mode", ["S", "P"])
def test_variant_xor_shift_tree(ctx, oracle, shape, growth, mode):
    rows, cols, T, thr = shape
    rng = np.random.default_rng(rows * 7 + cols + growth)
    n = 3000
    depth = rng.integers(1, 10, n).astype(np.uint32)
    kw = dict(state_seed=rng.integers(1, 2**63, n).astype(np.uint64), state_pre_discard=rng.integers(0, 40, n).astype(np.uint32),
              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64), payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
    mseeds = rng.integers(1, 2**63, n).astype(np.uint64) if mode == "P" else None
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, model_seed=987654321,
                                      model_seeds=mseeds, engine="xs", growth=growth, **kw)
    out = dict(payoff=np.full(n, 255, np.uint8), traj=np.zeros((n, 9, 3), np.uint8), traj_hash=np.zeros(n, np.uint32),
               sums=np.zeros(8, np.uint64))
    xs_state = np.array([987654321], np.uint64) if mode == "S" else None
    ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth.astype(np.uint16),
                            state_seed=exp["state_after_ctor"], transition_seed=kw["transition_seed"], payoff_bias=kw["payoff_bias"],
                            model_seeds=mseeds, xs_state=xs_state, xor_shift=True, growth=growth, traj_stride=9, **out)
    assert ctx.last_path == "thread_tree_xs"
    check_same(out, exp, n)
    if mode == "S":  # the model engine ends where the reference's would: two draws per step
        assert int(xs_state[0]) == ctx.lib.pcu_xs_discard(987654321, 2 * int(depth.sum()))


@pytest.mark.parametrize("shape", [(3, 3, 1, (0, 1)), (1, 3, 2, (3, 10)), (3, 1, 1, (0, 1)), (4, 5, 3, (1, 2))])
@pytest.mark.parametrize("mode", ["S", "P", "W"])
def test_variant_growth_mt19937(ctx, oracle, shape, mode):
    rows, cols, T, thr = shape
    rng = np.random.default_rng(rows * 11 + cols)
    n = 2000
    depth = rng.integers(1, 10, n).astype(np.uint32) if mode != "S" else 7
    kw = dict(state_seed=rng.integers(0, 2**63, n).astype(np.uint64), state_pre_discard=rng.integers(0, 700, n).astype(np.uint32),
              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64), payoff_bias=rng.integers(-2, 3, n).astype(np.int32))
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64) if mode != "S" else None
    exp = loader.random_tree_rollouts(oracle, n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, model_seed=31,
                                      model_pre_discard=5, model_seeds=mseeds, growth=1, **kw)
    dmax = int(np.max(depth))
    out = dict(payoff=np.full(n, 255, np.uint8), traj=np.zeros((n, dmax, 3), np.uint8), traj_hash=np.zeros(n, np.uint32),
               sums=np.zeros(8, np.uint64))
    gen = 2 * rows * cols * T
    eng = Engine(31).discard(5) if mode == "S" else None
    ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth.astype(np.uint16) if mode != "S" else depth,
                            state_seed=kw["state_seed"], state_pos=kw["state_pre_discard"] + np.uint32(gen),
                            transition_seed=kw["transition_seed"], payoff_bias=kw["payoff_bias"], engine=eng,
                            model_seeds=mseeds if mode == "P" else None,
                            model_words=_words_for(mseeds, depth) if mode == "W" else None, growth=1, traj_stride=dmax, **out)
    assert ctx.last_path == "thread_tree_growth"
    check_same(out, exp, n)
    if mode == "S":
        assert engine_matches(eng, 31, 5 + 4 * 7 * n, oracle)


@pytest.mark.parametrize("mode", ["S", "P"])
def test_variant_xor_shift_mold(ctx, oracle, mode):
    rng = np.random.default_rng(3)
    n = 5000
    depth = rng.integers(0, 14, n).astype(np.uint32)
    mseeds = rng.integers(1, 2**63, n).astype(np.uint64) if mode == "P" else None
    exp = loader.mold_rollouts(oracle, n, n_actions=3, depth=depth, model_seed=4242, model_seeds=mseeds, engine="xs")
    traj = np.zeros((n, 13, 3), np.uint8)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    xs_state = np.array([4242], np.uint64) if mode == "S" else None
    ctx.rollout_mold(n, n_actions=3, depth=depth.astype(np.uint16), model_seeds=mseeds, xs_state=xs_state, xor_shift=True,
                     traj=traj, traj_stride=13, traj_hash=th, sums=sums)
    assert ctx.last_path == "thread_mold_xs"
    assert np.array_equal(traj, exp["traj"]) and np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])


# ---------------------------------------------------------------------------------------------
# One context over several GPUs (pcu_create_multi).  On a one-GPU box the shards share GPU 0 (listing a device twice is
# allowed for exactly this): the shard plan, the per-shard slices of the global jump family, the output offsets and the
# post-batch engine are the same code; only the reduction differs (a summing kernel instead of ncclAllReduce).


def _multi_ids():
    import torch

    k = torch.cuda.device_count()
    return list(range(k)) if k >= 2 else [0, 0, 0]


@pytest.fixture(scope="module")
def mctx():
    c = Context(_multi_ids())
    c.set_option("multi_min_rollouts", 1000)
    yield c
    c.close()


@pytest.mark.parametrize("case", [CASES[0], CASES[1], CASES[3]])
def test_multi_gpu_context_equals_single(ctx, mctx, oracle, case):
    rows, cols, T, thr, depth, sseed, pre, tseed, bias, mseed, mpre = case
    n = 200000 + 77
    kw = dict(rows=rows, cols=cols, T=T, thr=thr, depth=depth, state_seed=sseed, pre=pre, tseed=tseed, bias=bias)
    e1, e2 = Engine(mseed).discard(mpre), Engine(mseed).discard(mpre)
    one = run_tree(ctx, n, engine=e1, want_traj=False, **kw)
    many = run_tree(mctx, n, engine=e2, want_traj=False, **kw)
    assert mctx.last_path.startswith("stream_table x") and mctx.device_count >= 2
    check_same(many, one, n)
    assert np.array_equal(e1.x, e2.x) and int(e1.p[0]) == int(e2.p[0])  # the caller's engine: byte for byte
    exp = loader.random_tree_rollouts(oracle, 3000, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth,
                                      state_seed=sseed, state_pre_discard=pre, transition_seed=tseed, payoff_bias=bias,
                                      model_seed=mseed, model_pre_discard=mpre, want_traj=False)
    assert np.array_equal(many["payoff"][:3000], exp["payoff"])


def test_multi_gpu_context_general_paths(ctx, mctx, oracle):
    """Distinct roots (uniform depth: shards are slices of one family of code streams), ragged depths (per-shard
    engines advanced on the host), mode P, MoldState, packed output and trajectories across shard boundaries."""
    rng = np.random.default_rng(13)
    n = 30011
    seeds = rng.integers(0, 2**63, n).astype(np.uint64)
    depth = rng.integers(1, 12, n).astype(np.uint32)
    mseeds = rng.integers(0, 2**63, n).astype(np.uint64)
    for kw, engine_words in (
        (dict(rows=3, cols=3, depth=10, state_seed=seeds), 40 * n),
        (dict(rows=3, cols=3, depth=depth, state_seed=seeds), 4 * int(depth.sum())),
        (dict(rows=4, cols=3, transitions=2, depth=depth, state_seed=seeds, transition_seed=seeds), 4 * int(depth.sum())),
        (dict(rows=3, cols=3, depth=10, state_seed=seeds, model_seeds=mseeds), 0),
    ):
        ms = kw.pop("model_seeds", None)
        exp = loader.random_tree_rollouts(oracle, n, model_seed=5, model_seeds=ms, want_traj=True, **kw)
        d = kw.pop("depth")
        dmax = int(d) if np.isscalar(d) else int(d.max())
        payoff = np.full(n, 255, np.uint8)
        packed = np.zeros((n + 3) // 4, np.uint8)
        traj = np.zeros((n, dmax, 3), np.uint8)
        th = np.zeros(n, np.uint32)
        sums = np.zeros(8, np.uint64)
        eng = None if ms is not None else Engine(5)
        mctx.rollout_random_tree(n, depth=d if np.isscalar(d) else d.astype(np.uint16), engine=eng, model_seeds=ms,
                                 payoff=payoff, payoff_packed=packed, traj=traj, traj_stride=dmax, traj_hash=th, sums=sums, **kw)
        assert " x" in mctx.last_path, mctx.last_path
        assert np.array_equal(payoff, exp["payoff"]) and np.array_equal(unpack2(packed, n), exp["payoff"])
        assert np.array_equal(traj, exp["traj"]) and np.array_equal(th, exp["traj_hash"])
        assert np.array_equal(sums, exp["sums"])
        if eng is not None:
            assert engine_matches(eng, 5, engine_words, oracle)
    exp = loader.mold_rollouts(oracle, n, n_actions=3, depth=depth, model_seed=6, want_traj=False)
    th = np.zeros(n, np.uint32)
    sums = np.zeros(8, np.uint64)
    eng = Engine(6)
    mctx.rollout_mold(n, n_actions=3, depth=depth.astype(np.uint16), engine=eng, traj_hash=th, sums=sums)
    assert np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])
    assert engine_matches(eng, 6, 4 * int(depth.sum()), oracle)
    exp = loader.mold_rollouts(oracle, n, n_actions=2, depth=10, model_seed=6, want_traj=False)
    eng = Engine(6)
    mctx.rollout_mold(n, n_actions=2, depth=10, engine=eng, traj_hash=th, sums=sums)
    assert mctx.last_path.startswith("stream_mold x")
    assert np.array_equal(th, exp["traj_hash"]) and np.array_equal(sums, exp["sums"])
    assert engine_matches(eng, 6, 40 * n, oracle)


def test_multi_gpu_small_batches_stay_on_first_gpu(mctx, oracle):
    mctx.set_option("multi_min_rollouts", 1 << 16)
    try:
        exp = loader.random_tree_rollouts(oracle, 500, rows=3, cols=3, depth=10, want_traj=False)
        got = run_tree(mctx, 500, rows=3, cols=3, depth=10, engine=Engine(0), want_traj=False)
        assert mctx.last_path == "stream_table"
        check_same(got, exp, 500)
    finally:
        mctx.set_option("multi_min_rollouts", 1000)


def test_warmup_pays_the_family_build(ctx):
    """pcu_warmup builds the jump-polynomial family a batch shape needs; the call after it finds it cached."""
    n, depth = 3_000_000 + 12800, 7  # a stride no other test uses
    ctx.warmup(n, depth)
    sums = np.zeros(8, np.uint64)
    ctx.rollout_mold(n, n_actions=2, depth=depth, engine=Engine(0), sums=sums)
    assert ctx.stat("last_family_build_ms") == 0.0
    ctx.rollout_mold(n + 40 * 12800, n_actions=2, depth=depth, engine=Engine(0), sums=sums)  # a new stride: built in the call
    assert ctx.stat("last_family_build_ms") > 0.0


@pytest.mark.parametrize("args", [("80", "7"), ("80", "8", "lane")])
def test_randomised_differential(args):
    """tests/fuzz_parity.py: random shapes, batch kinds and model modes against the oracle, every kernel path."""
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    res = subprocess.run([sys.executable, str(root / "tests" / "fuzz_parity.py"), *args], capture_output=True, text=True,
                         timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "bit-exact" in res.stdout
