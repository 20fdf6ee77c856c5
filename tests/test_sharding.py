"""Multi-GPU host logic on CPU: world_size-2 gloo process group.

Each rank takes its contiguous slice of ONE model stream (pinyon_b200.sharding.plan), jumps its
copy of the model engine there with the library's host-side pcu_mt_discard, and the win/draw/
loss/step/hash counters are summed with one all-reduce — exactly what bench.py does over NCCL.
The rollouts themselves are played by the oracle here (no GPU in this container), which is what
makes the test meaningful: the sharded result must equal the unsharded reference semantics
(monte-carlo-model.hh:58-67, one engine consumed in index order)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import loader
from pinyon_b200 import Engine, sharding

N_TOTAL = 3001  # odd on purpose: ranks get 1501 and 1500 rollouts
CFG = dict(rows=3, cols=3, transitions=1, depth=10)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = loader.load_oracle()
        sh = sharding.plan(N_TOTAL, world, rank, CFG["depth"])
        # the engine this rank would hand to pcu_rollout_random_tree: Model{prng{0}} jumped to the slice
        eng = Engine(0).discard(sh.engine_offset)
        first_words = [eng.copy().next() for _ in range(1)]
        exp_word = np.zeros(1, np.uint32)
        orc.orc_mt_words(0, sh.engine_offset, 1, exp_word.ctypes.data)
        assert first_words[0] == int(exp_word[0])
        part = loader.random_tree_rollouts(orc, sh.count, model_seed=0, model_pre_discard=sh.engine_offset,
                                           want_traj=False, **CFG)
        sums = torch.from_numpy(part["sums"].astype(np.int64))
        sharding.reduce_sums(sums, dist)
        gathered = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(gathered, torch.tensor([sh.first, sh.count]))
        if rank == 0:
            np.save(out_path, np.concatenate([sums.numpy(), torch.stack(gathered).numpy().ravel()]))
    finally:
        dist.destroy_process_group()


def test_plan_covers_batch_exactly():
    for n, world in ((0, 1), (1, 4), (7, 2), (2**24, 8), (1000003, 8)):
        shards = [sharding.plan(n, world, r, 10) for r in range(world)]
        assert shards[0].first == 0 and sum(s.count for s in shards) == n
        for a, b in zip(shards, shards[1:]):
            assert a.first + a.count == b.first
            assert b.engine_offset == 4 * 10 * b.first
    depths = [3, 1, 4, 1, 5, 9, 2, 6]
    s1 = sharding.plan_ragged(depths, 2, 1)
    assert (s1.first, s1.count, s1.engine_offset) == (4, 4, 4 * (3 + 1 + 4 + 1))
    with pytest.raises(ValueError):
        sharding.plan(10, 2, 2, 10)


@pytest.mark.timeout(300)
def test_two_rank_gloo_matches_unsharded(tmp_path):
    out = tmp_path / "sums.npy"
    mp.spawn(_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    got = np.load(out)
    whole = loader.random_tree_rollouts(loader.load_oracle(), N_TOTAL, want_traj=False, **CFG)
    assert np.array_equal(got[:8].astype(np.uint64), whole["sums"])
    assert got[8:].tolist() == [0, 1501, 1501, 1500]
