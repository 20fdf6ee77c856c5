"""pcu_leaf_batch_noalloc (include/pinyon_cuda.h): the leaf-batch path with its scratch in static device memory.

No counterpart in the reference — the entry point exists for callers that cannot call cudaMalloc.  The check is the
existing entry point (pcu_rollout_random_tree, mode W, path "leaf_table_W"), bit-exact like every other parity test
here (integer work: tolerance 0), on two calls in a row that share the static array.  Collected last on purpose
(file name), and self-contained: it passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

from pinyon_b200 import Engine, PcuError

pytestmark = pytest.mark.gpu

NOALLOC_SCRATCH_BYTES = 64 * 1024


def _words(n, depth, seed0):
    return np.concatenate([Engine(seed0 + i).take_raw(4 * depth) for i in range(n)])


def _both(ctx, n, *, rows, cols, T, thr, depth, state_seed, pre, tseed, bias, seed0):
    import torch

    words = _words(n, depth, seed0)
    pos = 2 * rows * cols * T + pre
    # existing entry point, host buffers
    exp_pay = np.full(n, 255, np.uint8)
    exp_sums = np.zeros(8, np.uint64)
    ctx.rollout_random_tree(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, state_seed=state_seed,
                            state_pos=pos, transition_seed=tseed, payoff_bias=bias, model_words=words, payoff=exp_pay,
                            sums=exp_sums)
    assert ctx.last_path == "leaf_table_W"
    # new entry point: the caller owns every device buffer (torch here), the library allocates nothing
    d_words = torch.from_numpy(words.view(np.int32)).cuda()
    d_pay = torch.full((n,), 255, dtype=torch.uint8, device="cuda")
    d_sums = torch.full((8,), -1, dtype=torch.int64, device="cuda")  # garbage: the call clears it
    torch.cuda.synchronize()
    ctx.leaf_batch_noalloc(n, rows=rows, cols=cols, transitions=T, thr=thr, depth=depth, state_seed=state_seed,
                           state_pos=pos, transition_seed=tseed, payoff_bias=bias, words=d_words, payoff=d_pay, sums=d_sums)
    assert ctx.last_path == "leaf_table_noalloc"
    got_pay = d_pay.cpu().numpy()
    got_sums = d_sums.cpu().numpy().view(np.uint64)
    assert np.array_equal(got_pay, exp_pay)
    assert np.array_equal(got_sums, exp_sums)
    assert int(got_sums[0] + got_sums[1] + got_sums[2]) == n
    return got_pay


def test_noalloc_matches_existing_entry_point_twice_in_a_row(ctx):
    # call 1: the largest table of the two (9x9x3, depth 20: ~30 KB with the chance-action half)
    a = _both(ctx, 4096 + 17, rows=9, cols=9, T=3, thr=(0, 1), depth=20, state_seed=21, pre=37, tseed=4, bias=0, seed0=1000)
    # call 2 right behind it on the same static array: another stream, a SHORTER table (stale bytes behind it), T = 1
    b = _both(ctx, 2048 + 5, rows=3, cols=3, T=1, thr=(0, 1), depth=10, state_seed=5, pre=0, tseed=0, bias=1, seed0=7000)
    # and the first shape again: the result must not depend on what the previous call left behind
    c = _both(ctx, 4096 + 17, rows=9, cols=9, T=3, thr=(0, 1), depth=20, state_seed=21, pre=37, tseed=4, bias=0, seed0=1000)
    assert np.array_equal(a, c)
    assert len(set(a.tolist())) > 1 and len(set(b.tolist())) > 1


def test_noalloc_rejects_sizes_beyond_the_static_scratch(ctx):
    import torch

    d_words = torch.zeros(16, dtype=torch.int32, device="cuda")
    d_pay = torch.zeros(4, dtype=torch.uint8, device="cuda")
    # 9x9x3 at depth 60: (486 + 731 * 60 + 8) * 2 bytes > 64 KiB
    assert (486 + 731 * 60 + 8) * 2 > NOALLOC_SCRATCH_BYTES
    with pytest.raises(PcuError) as e:
        ctx.leaf_batch_noalloc(1, rows=9, cols=9, transitions=3, depth=60, words=d_words, payoff=d_pay)
    assert "static" in str(e.value) and "scratch" in str(e.value)
    with pytest.raises(PcuError):
        ctx.leaf_batch_noalloc((1 << 24) + 1, rows=3, cols=3, depth=1, words=d_words, payoff=d_pay)
