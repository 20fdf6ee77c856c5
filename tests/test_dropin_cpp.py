"""The header-only C++ drop-in model (include/pinyon_cuda/*.hh) against the reference's own
MonteCarloModel, TreeBandit<Exp3> and OffPolicy<Exp3> — tests/cpp/test_dropin.cc."""
import subprocess
import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent / "cpp"))
import build as cpp_build  # noqa: E402


def _binary():
    from pinyon_b200 import build_library

    build_library()
    if cpp_build.reference_available():
        return cpp_build.build()
    if cpp_build.BIN.exists():
        return cpp_build.BIN
    pytest.skip("test_dropin was not prebuilt and the reference headers are absent")


def _run(*flags):
    res = subprocess.run([str(_binary()), *flags], capture_output=True, text=True, timeout=900)
    print(res.stdout)
    print(res.stderr, file=sys.stderr)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-2000:]
    assert "ALL OK" in res.stdout
    return res.stdout


def test_dropin_host_logic_and_concepts():
    """static_asserts of IsValue/Single/Batch/PolicyModelTypes and IsSearchTypes<TreeBandit<Exp3<Gpu>>> hold at
    compile time; engine access, the stream locator and the no-GPU failure are checked at run time."""
    out = _run("--cpu")
    assert "engine access ok" in out and "stream locator ok" in out and "fibers ok" in out
    import torch

    if not torch.cuda.is_available():
        assert "no CPU fallback" in out


@pytest.mark.gpu
def test_dropin_matches_reference_on_gpu():
    out = _run("--gpu")
    for part in ("single inference ok", "batch inference ok", "TreeBandit<Exp3>", "OffPolicy<Exp3>", "threaded copies ok",
                 "batching adapter ok", "every tree == reference", "variants ok"):
        assert part in out


@pytest.mark.gpu
def test_dropin_multi_gpu_batch_equals_single():
    """One model.inference(batch_in, batch_out) on a multi-GPU context (pcu_create_multi) == the same call on one GPU,
    byte for byte, including model.device afterwards: 2^24 leaves when the box has several GPUs, 2^22 shards sharing the
    one GPU otherwise."""
    import torch

    n = 1 << (24 if torch.cuda.device_count() >= 2 else 22)
    out = _run("--multi", str(n))
    assert "multi-GPU batch ok" in out


@pytest.mark.gpu
def test_search_bench_binary_ref_equals_gpu():
    """tests/cpp/bench_search.cc (bench.py --workload search_offpolicy, BASELINE configs[4]): the reference's
    OffPolicy<Exp3> with 4096-leaf batches builds the same tree over the GPU model as over the reference model."""
    import json

    _binary()
    if cpp_build.reference_available():
        binary = cpp_build.build_bench()
    elif cpp_build.BENCH_BIN.exists():
        binary = cpp_build.BENCH_BIN
    else:
        pytest.skip("bench_search was not prebuilt and the reference headers are absent")
    outs = []
    for impl in ("ref", "gpu"):
        res = subprocess.run([str(binary), impl, "0", "5", "4096", "2", "1"], capture_output=True, text=True, timeout=600)
        assert res.returncode == 0, res.stderr[-2000:]
        outs.append(json.loads(res.stdout.strip().splitlines()[-1]))
    for key in ("nodes", "visits", "value_total_row", "row_visits", "col_visits"):
        assert outs[0][key] == outs[1][key], key
    assert outs[0]["nodes"] == 3 * 4096


def _callers_binary():
    from pinyon_b200 import build_library

    build_library()
    if cpp_build.reference_available():
        return cpp_build.build_callers()
    if cpp_build.CALLERS_BIN.exists():
        return cpp_build.CALLERS_BIN
    pytest.skip("test_callers was not prebuilt and the reference headers are absent")


def _run_callers(flag):
    res = subprocess.run([str(_callers_binary()), flag], capture_output=True, text=True, timeout=900)
    print(res.stdout)
    print(res.stderr, file=sys.stderr)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-2000:]
    assert "ALL OK" in res.stdout
    return res.stdout


def test_callers_concepts_and_stand_ins():
    """tests/cpp/test_callers.cc compiles: IsBanditAlgorithmTypes<MatrixUCBFixed<Gpu>>, IsSearchTypes<TreeBandit<MatrixUCB
    <Gpu>>> and FullTraversal's constraints hold over the GPU model; the lrslib stand-in and the MatrixUCB fix-up run the
    reference's FullTraversal / MatrixUCB / TreeBanditFlat over the reference model on the host."""
    out = _run_callers("--cpu")
    assert "FullTraversal value" in out


@pytest.mark.gpu
def test_matrix_ucb_full_traversal_and_flat_over_gpu_model():
    """SURVEY 8(f)3 + tests/flat.cc: FullTraversal (model.inference at the depth cut-off, LRSNash::solve per node),
    TreeBandit<MatrixUCB> and TreeBandit vs TreeBanditFlat give identical results over the GPU model and the reference's."""
    out = _run_callers("--gpu")
    assert "TreeBandit<MatrixUCB>" in out and "flat equivalence" in out and out.count("FullTraversal seed") == 3
