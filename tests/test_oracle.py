"""CPU tests: pin the plain-C oracle (oracle/pinyon_oracle.c) against
  (1) tests/golden/reference_vectors.json — produced by the reference's own classes
      (tests/golden/make_golden.py), and
  (2) the compiled reference itself (oracle/_ref/libpinyon_ref.so) when it has been built.
No GPU, no product code on this path except the host-only engine helpers in test_abi.py."""
import json
from pathlib import Path

import numpy as np
import pytest

from oracle import loader

GOLDEN = json.loads((Path(__file__).parent / "golden" / "reference_vectors.json").read_text())


def _case_kwargs(case):
    kw = dict(case)
    kw["thr"] = tuple(kw["thr"])
    return kw


@pytest.mark.parametrize("entry", GOLDEN["tree"], ids=lambda e: "x".join(str(e["case"][k]) for k in ("rows", "cols", "transitions", "depth")))
def test_tree_golden(oracle, entry):
    kw = _case_kwargs(entry["case"])
    n = entry["n"]
    got = loader.random_tree_rollouts(oracle, n, **kw)
    assert got["payoff"].tolist() == entry["payoff_x2"]
    assert got["traj"].reshape(n, -1).tolist() == entry["traj"]
    assert int(got["steps"]) == entry["steps"]
    strat = np.zeros(kw["rows"] * kw["cols"] * kw["transitions"])
    oracle.orc_random_tree_root_strategies(kw.get("state_seed", 0), kw.get("state_pre_discard", 0), kw["depth"],
                                           kw["rows"], kw["cols"], kw["transitions"], *kw["thr"], strat.ctypes.data)
    assert [float(x).hex() for x in strat[:27]] == entry["root_strategies_hex"]
    if "counts" in entry:
        c = entry["counts"]
        big = loader.random_tree_rollouts(oracle, c["n"], want_traj=False, **kw)
        assert big["sums"][:3].tolist() == [c["wins"], c["draws"], c["losses"]]
        assert int(big["sums"][3]) == c["n"] * kw["depth"]


def test_appendix_b_known_answers(oracle):
    """SURVEY.md Appendix B1/B2 literals (reference MonteCarloModel, g++ 13 / libstdc++ 13)."""
    b1 = [1, 0, 1, .5, 0, 1, 0, 1, 0, 1, 1, 1, 0, 1, .5, 1, 0, .5, 1, .5, 1, 1, 1, 1, 1, 1, .5, .5, 0, 1, 0, 0]
    got = loader.random_tree_rollouts(oracle, 32, rows=3, cols=3, depth=10)
    assert (got["payoff"] / 2.0).tolist() == b1
    assert got["traj"][1].tolist() == [[2, 2, 0], [0, 2, 0], [0, 0, 0], [0, 1, 0], [0, 1, 0], [2, 0, 0], [0, 2, 0],
                                       [0, 1, 0], [1, 2, 0], [1, 0, 0]]
    b2 = [0, 1, .5, .5, 1, 0, .5, 1, 1, 1, 0, 0, 0, 0, .5, 0, 1, 0, 0, 1, 1, 1, 1, 1, 1, .5, 0, 1, 0, 1, 0, 0]
    got = loader.random_tree_rollouts(oracle, 32, rows=9, cols=9, transitions=3, depth=20)
    assert (got["payoff"] / 2.0).tolist() == b2
    assert got["traj"][0][-3:].tolist() == [[0, 7, 1], [4, 5, 1], [7, 3, 1]]
    strat = np.zeros(243)
    oracle.orc_random_tree_root_strategies(0, 0, 20, 9, 9, 3, 0, 1, strat.ctypes.data)
    assert strat[:3].tolist() == [0.19047619047619047, 0.42857142857142855, 0.38095238095238093]
    u = np.zeros(2, np.uint64)
    oracle.orc_prng_uniform_64(0, 0, 2, u.ctypes.data)  # prng{0}: the survey's literal is the SECOND draw
    assert [int(x) for x in u] == [10123822223749065263, 13192915183495924928]
    d = np.zeros(1)
    oracle.orc_prng_uniform(0, 0, 1, d.ctypes.data)
    assert d[0] == 0.59284461651668263
    w = np.zeros(1, np.uint32)
    oracle.orc_mt_words(5489, 9999, 1, w.ctypes.data)  # the C++ standard's mt19937 check value
    assert int(w[0]) == 4123659995


@pytest.mark.parametrize("entry", GOLDEN["private"], ids=lambda e: f"{e['rows']}x{e['cols']}x{e['transitions']}-{e['mode']}")
def test_tree_private_roots_golden(oracle, entry):
    n = entry["n"]
    arrays = entry["arrays"]
    got = loader.random_tree_rollouts(
        oracle, n, rows=entry["rows"], cols=entry["cols"], transitions=entry["transitions"], thr=tuple(entry["thr"]),
        state_seed=np.array(arrays["state_seed"], np.uint64), state_pre_discard=np.array(arrays["state_pre_discard"], np.uint32),
        transition_seed=np.array(arrays["transition_seed"], np.uint64), depth=np.array(arrays["depth"], np.uint32),
        payoff_bias=np.array(arrays["payoff_bias"], np.int32), model_seed=entry["model_seed"],
        model_pre_discard=entry["model_pre_discard"],
        model_seeds=np.array(entry["model_seeds"], np.uint64) if entry["mode"] == "P" else None)
    assert got["payoff"].tolist() == entry["payoff_x2"]
    assert got["traj"].reshape(n, -1).tolist() == entry["traj"]
    assert int(got["steps"]) == entry["steps"] == sum(arrays["depth"])


@pytest.mark.parametrize("entry", GOLDEN["mold"], ids=lambda e: f"{e['case']['n_actions']}-{e['case']['depth']}")
def test_mold_golden(oracle, entry):
    n = entry["n"]
    got = loader.mold_rollouts(oracle, n, **entry["case"])
    assert got["traj"].reshape(n, -1).tolist() == entry["traj"]
    assert int(got["steps"]) == entry["steps"] == n * entry["case"]["depth"]
    # MoldState never writes its payoff (test-states.hh:43-49): the reference returns (0, 0)
    assert not any(entry["payoff_row"]) and not any(entry["payoff_col"]) and not got["payoff"].any()
    assert int(got["sums"][1]) == n


def test_prng_golden(oracle):
    for key, g in GOLDEN["prng"].items():
        seed, disc = (int(x) for x in key.split(":"))
        u64 = np.zeros(16, np.uint64)
        ud = np.zeros(16)
        oracle.orc_prng_uniform_64(seed, disc, 16, u64.ctypes.data)
        oracle.orc_prng_uniform(seed, disc, 16, ud.ctypes.data)
        assert [int(x) for x in u64] == g["uniform_64"]
        assert [float(x).hex() for x in ud] == g["uniform_hex"]
        for m, exp in g["random_int"].items():
            a = np.zeros(16, np.int32)
            oracle.orc_prng_random_int(seed, disc, int(m), 16, a.ctypes.data)
            assert a.tolist() == exp


def test_golden_layout_facts():
    """sizeof(prng) and sizeof(RandomTree<>::State) under libstdc++/x86-64."""
    assert GOLDEN["sizeof"] == {"prng": 5040, "random_tree_state": 5232}
    assert GOLDEN["tree_bandit_exp3_b4"]["stats"][:2] == [4096, 4095]


# ---- against the compiled reference itself (only where oracle/_ref/libpinyon_ref.so exists)


def test_oracle_vs_reference_random(oracle, reference):
    rng = np.random.default_rng(17)
    for trial in range(12):
        rows, cols, T = int(rng.integers(1, 10)), int(rng.integers(1, 10)), int(rng.integers(1, 5))
        thr = [(0, 1), (1, 2), (3, 10), (1, 10), (1, 1)][trial % 5]
        kw = dict(rows=rows, cols=cols, transitions=T, thr=thr, depth=int(rng.integers(1, 21)),
                  state_seed=int(rng.integers(0, 2**63)), state_pre_discard=int(rng.integers(0, 2000)),
                  transition_seed=int(rng.integers(0, 2**63)), payoff_bias=int(rng.integers(-3, 4)),
                  model_seed=int(rng.integers(0, 2**63)), model_pre_discard=int(rng.integers(0, 1300)))
        a = loader.random_tree_rollouts(oracle, 150, **kw)
        b = loader.random_tree_rollouts(reference, 150, prefix="ref", **kw)
        assert np.array_equal(a["payoff"], b["payoff"]), kw
        assert np.array_equal(a["traj"], b["traj"]), kw
        assert a["steps"] == b["steps"]


def test_oracle_vs_reference_private_and_mold(oracle, reference):
    rng = np.random.default_rng(5)
    n = 200
    kw = dict(rows=6, cols=5, transitions=3, thr=(1, 5), state_seed=rng.integers(0, 2**63, n).astype(np.uint64),
              state_pre_discard=rng.integers(0, 3000, n).astype(np.uint32),
              transition_seed=rng.integers(0, 2**63, n).astype(np.uint64), depth=rng.integers(1, 25, n).astype(np.uint32),
              payoff_bias=rng.integers(-2, 3, n).astype(np.int32), model_seeds=rng.integers(0, 2**63, n).astype(np.uint64))
    a = loader.random_tree_rollouts(oracle, n, **kw)
    b = loader.random_tree_rollouts(reference, n, prefix="ref", **kw)
    assert np.array_equal(a["payoff"], b["payoff"]) and np.array_equal(a["traj"], b["traj"])
    depth = rng.integers(0, 40, n).astype(np.uint32)
    a = loader.mold_rollouts(oracle, n, n_actions=5, depth=depth, model_seed=3, model_pre_discard=77)
    b = loader.mold_rollouts(reference, n, n_actions=5, depth=depth, model_seed=3, model_pre_discard=77, prefix="ref")
    assert np.array_equal(a["traj"], b["traj"]) and a["steps"] == b["steps"]


def test_oracle_vs_reference_prng(oracle, reference):
    for seed, disc in [(0, 0), (1, 623), (2**40 + 3, 624), (987654321, 10**6)]:
        a, b = np.zeros(4096, np.uint64), np.zeros(4096, np.uint64)
        oracle.orc_prng_uniform_64(seed, disc, 4096, a.ctypes.data)
        reference.ref_prng_uniform_64(seed, disc, 4096, b.ctypes.data)
        assert np.array_equal(a, b)
        x, y = np.zeros(4096), np.zeros(4096)
        oracle.orc_prng_uniform(seed, disc, 4096, x.ctypes.data)
        reference.ref_prng_uniform(seed, disc, 4096, y.ctypes.data)
        assert np.array_equal(x.view(np.uint64), y.view(np.uint64))


def _variant_arrays(e):
    dt = dict(state_seed=np.uint64, state_pre_discard=np.uint32, transition_seed=np.uint64, depth=np.uint32,
              payoff_bias=np.int32)
    return {k: np.array(v, dtype=dt[k]) for k, v in e["arrays"].items()}


@pytest.mark.parametrize("entry", GOLDEN.get("variants", []), ids=lambda e: f"{e['engine']}-g{e['growth']}-{e['rows']}x{e['cols']}x{e['transitions']}")
def test_variant_golden(oracle, entry):
    """SURVEY 8(f)4 variants: the restatement of xor_shift (random.hh:66-121) and of GrowLib::alternating_constant_actions
    (grow-lib.hh:16-24) against vectors the compiled reference produced (make_golden.py)."""
    e = entry
    res = loader.random_tree_rollouts(oracle, e["n"], rows=e["rows"], cols=e["cols"], transitions=e["transitions"],
                                      thr=tuple(e["thr"]), model_seed=e["model_seed"], model_pre_discard=e["model_pre_discard"],
                                      engine=e["engine"], growth=e["growth"], **_variant_arrays(e))
    assert res["payoff"].tolist() == e["payoff_x2"]
    assert res["traj"].reshape(e["n"], -1).tolist() == e["traj"]
    assert int(res["steps"]) == e["steps"]
    if e["engine"] == "xs":
        assert [int(x) for x in res["state_after_ctor"]] == e["state_after_ctor"]


def test_variant_mold_xor_shift_golden(oracle):
    e = GOLDEN["variants_mold_xs"]
    res = loader.mold_rollouts(oracle, e["n"], n_actions=3, depth=np.array(e["depth"], np.uint32), model_seed=e["model_seed"],
                               engine="xs")
    assert res["traj"].reshape(e["n"], -1).tolist() == e["traj"]


def test_variants_oracle_vs_reference(oracle, reference):
    rng = np.random.default_rng(5)
    n = 200
    for engine, growth in (("xs", 0), ("xs", 1), ("mt", 1)):
        for rows, cols, T, thr in ((2, 2, 1, (0, 1)), (1, 4, 3, (3, 10)), (5, 1, 2, (1, 2)), (3, 4, 8, (1, 1))):
            kw = dict(rows=rows, cols=cols, transitions=T, thr=thr, depth=rng.integers(1, 10, n).astype(np.uint32),
                      state_seed=rng.integers(1, 2**63, n).astype(np.uint64),
                      state_pre_discard=rng.integers(0, 30, n).astype(np.uint32),
                      transition_seed=rng.integers(0, 2**63, n).astype(np.uint64),
                      payoff_bias=rng.integers(-2, 3, n).astype(np.int32), engine=engine, growth=growth)
            for ms in (None, rng.integers(1, 2**63, n).astype(np.uint64)):
                a = loader.random_tree_rollouts(oracle, n, model_seed=3, model_seeds=ms, **kw)
                b = loader.random_tree_rollouts(reference, n, model_seed=3, model_seeds=ms, prefix="ref", **kw)
                assert np.array_equal(a["payoff"], b["payoff"]) and np.array_equal(a["traj"], b["traj"])
                if engine == "xs":
                    assert np.array_equal(a["state_after_ctor"], b["state_after_ctor"])


def test_golden_file_is_current(reference):
    """The committed vectors are what the reference produces today (spot check, first case)."""
    e = GOLDEN["tree"][1]
    b = loader.random_tree_rollouts(reference, e["n"], prefix="ref", **_case_kwargs(e["case"]))
    assert b["payoff"].tolist() == e["payoff_x2"]


def test_nine_transitions_overflow_int_in_the_reference(oracle):
    """random-tree.hh:165-184 accumulates n_k/10 as Rational<int> (rational.hh:82-87: p = p*10 + n*q, q *= 10) before
    reducing: at T = 9 the numerator 10^8 * sum(n_k) passes 2^31, so the reference's chance strategies stop being
    probability vectors (signed overflow: undefined behaviour that wraps here).  T = 8 is the last exact case —
    SURVEY.md App. C6 put the limit at 9 — and libpinyon_cuda refuses T > 8 (tests/test_gpu_parity.py)."""
    def cells(T, seed):
        strat = np.zeros(2 * 2 * T)
        oracle.orc_random_tree_root_strategies(seed, 0, 5, 2, 2, T, 0, 1, strat.ctypes.data)
        return strat.reshape(4, T)

    def is_distribution(cell):
        return abs(cell.sum() - 1) < 1e-12 and (cell >= 0).all() and (cell <= 1).all()

    assert all(is_distribution(c) for seed in range(20) for c in cells(8, seed))
    assert not any(is_distribution(c) for seed in range(20) for c in cells(9, seed))
