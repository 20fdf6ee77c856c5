"""CPU tests of the drop-in boundary: libpinyon_cuda.so loads, exports every symbol
include/pinyon_cuda.h declares, its host-only engine helpers (seed / next / discard with
polynomial jump-ahead) agree with the oracle, and without a GPU the product path fails
loudly instead of falling back to anything."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from pinyon_b200 import Engine, PcuError, capi

ROOT = Path(__file__).resolve().parent.parent


def header_symbols():
    text = (ROOT / "include" / "pinyon_cuda.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pcu_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    declared = header_symbols()
    assert len(declared) >= 16
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing
    assert sorted(capi.EXPORTS) == declared  # the ctypes binding tracks the header
    assert lib.pcu_version() == 200


def test_struct_layouts_match_header():
    """ctypes mirrors vs the C structs, field by field (sizes/offsets from a tiny C probe)."""
    import subprocess
    import tempfile

    probe = r"""
    #include <stdio.h>
    #include <stddef.h>
    #include "pinyon_cuda.h"
    int main(void) {
      printf("%zu %zu %zu %zu\n", sizeof(pcu_model_rng), sizeof(pcu_rt_batch), sizeof(pcu_mold_batch), sizeof(pcu_out));
      printf("%zu %zu %zu %zu\n", offsetof(pcu_rt_batch, state_pos_u), offsetof(pcu_rt_batch, depth_u),
             offsetof(pcu_rt_batch, rows), offsetof(pcu_rt_batch, thr_q));
      printf("%zu %zu %zu %zu %zu %zu %zu\n", offsetof(pcu_out, traj_stride), offsetof(pcu_out, sums), offsetof(pcu_mold_batch, n_actions),
             offsetof(pcu_out, payoff_packed2), offsetof(pcu_rt_batch, growth), offsetof(pcu_model_rng, engine),
             offsetof(pcu_model_rng, xs_state));
      return 0; }
    """
    with tempfile.TemporaryDirectory() as d:
        src = Path(d) / "p.c"
        src.write_text(probe)
        subprocess.run(["gcc", "-I", str(ROOT / "include"), str(src), "-o", str(Path(d) / "p")], check=True)
        out = subprocess.run([str(Path(d) / "p")], check=True, capture_output=True, text=True).stdout.split()
    vals = [int(x) for x in out]
    assert vals[:4] == [C.sizeof(capi.ModelRng), C.sizeof(capi.RtBatch), C.sizeof(capi.MoldBatch), C.sizeof(capi.Out)]
    assert vals[4:8] == [capi.RtBatch.state_pos_u.offset, capi.RtBatch.depth_u.offset, capi.RtBatch.rows.offset,
                         capi.RtBatch.thr_q.offset]
    assert vals[8:] == [capi.Out.traj_stride.offset, capi.Out.sums.offset, capi.MoldBatch.n_actions.offset,
                        capi.Out.payoff_packed2.offset, capi.RtBatch.growth.offset, capi.ModelRng.engine.offset,
                        capi.ModelRng.xs_state.offset]


def test_host_engine_matches_oracle(oracle):
    for seed in (0, 1, 5489, 2**32 + 5, 12345678901234):
        e = Engine(seed)
        exp = np.zeros(2000, np.uint32)
        oracle.orc_mt_words(seed, 0, 2000, exp.ctypes.data)
        assert [e.next() for _ in range(2000)] == exp.tolist()
    x = np.zeros(624, np.uint32)
    p = np.zeros(1, np.uint32)
    oracle.orc_mt_state(77, 1000, x.ctypes.data, p.ctypes.data)
    e = Engine(77).discard(1000)
    assert np.array_equal(e.x, x) and int(e.p[0]) == int(p[0])


@pytest.mark.parametrize("n", [1 << 22, (1 << 22) + 12345, 40 * (1 << 24), 3 * 10**10 + 7])
def test_host_jump_ahead_matches_stepping(oracle, n):
    """pcu_mt_discard switches to polynomial jump-ahead at 2^22 draws; the stream that follows
    must be the one plain stepping reaches (oracle, for sizes it can step in seconds) and must
    compose: discard(a) then discard(b) == discard(a + b)."""
    e = Engine(3)
    e.next()  # leave the block boundary so the window reconstruction is exercised
    e.discard(n)
    got = [e.next() for _ in range(1300)]
    if n <= (1 << 30):
        exp = np.zeros(1300, np.uint32)
        oracle.orc_mt_words(3, 1 + n, 1300, exp.ctypes.data)
        assert got == exp.tolist()
    a = n // 3
    e2 = Engine(3)
    e2.next()
    e2.discard(a).discard(n - a)
    assert got == [e2.next() for _ in range(1300)]


def test_jump_family_matches_stepping(oracle):
    """The polynomial family x^(c * stride) mod phi the device jump-ahead uploads (squarings,
    carry-less multiplies, threaded level fill) must land every checkpoint where plain stepping
    does.  stride = 4 words * depth 10 * 224 rollouts, the unit of BASELINE configs[1]."""
    from pinyon_b200.capi import load_library

    lib = load_library()
    stride = 4 * 10 * 224
    for c in (0, 1, 2, 3, 5, 8, 31, 64, 77, 100):  # grows the cached family across several levels
        e = Engine(11)
        e.next()
        assert lib.pcu_mt_discard_family(e.x.ctypes.data, e.p.ctypes.data, stride, c) == 0
        exp = np.zeros(700, np.uint32)
        oracle.orc_mt_words(11, 1 + c * stride, 700, exp.ctypes.data)
        assert [e.next() for _ in range(700)] == exp.tolist(), c
    # a large stride (2^24 rollouts over 592 CTAs) composes with pcu_mt_discard
    big = 4 * 10 * 224 * 128
    e = Engine(5)
    assert lib.pcu_mt_discard_family(e.x.ctypes.data, e.p.ctypes.data, big, 591) == 0
    f = Engine(5).discard(591 * big)
    assert [e.next() for _ in range(700)] == [f.next() for _ in range(700)]


def test_window_reconstruction_every_position(oracle):
    """The jump path rebuilds the 624 words preceding the next draw from libstdc++'s (x, p)
    image for any p; check positions across a block against stepping."""
    for pre in (0, 1, 2, 226, 227, 228, 396, 397, 398, 622, 623, 624, 625, 1000):
        e = Engine(9).discard(pre)
        e.discard(1 << 22)
        exp = np.zeros(64, np.uint32)
        oracle.orc_mt_words(9, pre + (1 << 22), 64, exp.ctypes.data)
        assert [e.next() for _ in range(64)] == exp.tolist(), pre


def test_jump_family_slices_are_the_family(oracle):
    """A multi-GPU shard applies polynomials x^((first + j) * stride): its slice of ONE global family.  The host
    cache serves any [first, first + count) of a stride and survives eviction of other strides."""
    lib = capi.load_library()
    stride = 4 * 10 * 128
    lib.pcu_set_option  # the byte cap is a context option; exercised on the GPU box
    for first, c in ((0, 3), (37, 0), (37, 5), (500, 1)):
        e = Engine(4)
        e.next()
        assert lib.pcu_mt_discard_family(e.x.ctypes.data, e.p.ctypes.data, stride, first + c) == 0
        exp = np.zeros(100, np.uint32)
        oracle.orc_mt_words(4, 1 + (first + c) * stride, 100, exp.ctypes.data)
        assert [e.next() for _ in range(100)] == exp.tolist(), (first, c)


def test_take_raw_is_the_untempered_stream(oracle):
    """pcu_mt_take_raw hands out the RAW words a mode-W rollout consumes and advances the engine like prng::discard."""
    def temper(y):
        y = y ^ (y >> np.uint32(11))
        y = y ^ ((y << np.uint32(7)) & np.uint32(0x9D2C5680))
        y = y ^ ((y << np.uint32(15)) & np.uint32(0xEFC60000))
        return y ^ (y >> np.uint32(18))

    for pre, n in ((0, 40), (600, 80), (623, 1), (624, 1300), (5, 0)):
        e = Engine(17).discard(pre)
        raw = e.take_raw(n)
        exp = np.zeros(n + 3, np.uint32)
        oracle.orc_mt_words(17, pre, n + 3, exp.ctypes.data)
        assert np.array_equal(temper(raw), exp[:n])
        assert [e.next() for _ in range(3)] == exp[n:].tolist()  # the engine continues behind the words taken


def test_xor_shift_discard_matches_stepping(oracle):
    """pcu_xs_discard (powers of the 64 x 64 step matrix over GF(2)) == xor_shift::discard (random.hh:107-111)."""
    lib = capi.load_library()
    oracle.orc_set_engine_kind(1)
    try:
        for state, n in ((1, 0), (5, 1), (0x9E3779B97F4A7C15, 1000), (12345, 2**20 + 3)):
            out = np.zeros(1, np.uint64)
            oracle.orc_prng_uniform_64(state, n, 1, out.ctypes.data)  # the draw AFTER n discarded steps
            nxt = lib.pcu_xs_discard(state, n + 1)
            assert int(out[0]) == nxt
    finally:
        oracle.orc_set_engine_kind(0)
    assert lib.pcu_xs_discard(77, 2**40) == lib.pcu_xs_discard(lib.pcu_xs_discard(77, 2**39), 2**39)


def test_unpack2_roundtrip():
    rng = np.random.default_rng(0)
    for n in (0, 1, 3, 4, 5, 1001):
        pay = rng.integers(0, 3, n).astype(np.uint8)
        packed = np.zeros((n + 3) // 4, np.uint8)
        for i, v in enumerate(pay):
            packed[i // 4] |= v << (2 * (i % 4))
        assert np.array_equal(capi.unpack2(packed, n), pay)


def test_no_gpu_means_loud_failure():
    """Without an sm_100 device there is no product path: pcu_create fails, nothing computes."""
    try:
        import torch

        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    with pytest.raises(PcuError) as e:
        capi.Context(0)
    assert e.value.code == capi.PCU_ERR_CUDA
    with pytest.raises(PcuError) as e:
        capi.Context([0, 1])  # the multi-GPU parent fails the same way: no device, no context
    assert e.value.code == capi.PCU_ERR_CUDA


def test_product_does_not_reach_into_oracle():
    """Nothing under pinyon_b200/ or include/ imports, links or names the oracle."""
    for path in list((ROOT / "pinyon_b200").rglob("*")) + list((ROOT / "include").rglob("*")):
        if path.suffix in {".py", ".cu", ".cuh", ".h", ".hh", ".cc"}:
            text = path.read_text()
            assert "liboracle" not in text and "from oracle" not in text and "import oracle" not in text, path
            assert "pinyon_oracle" not in text and "libpinyon_ref" not in text, path
