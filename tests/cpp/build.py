"""Build tests/cpp/_bin/test_dropin: the C++23 drop-in test of PinyonCuda::MonteCarloModel.

Needs the reference headers (present in the build container only); the binary is git-ignored
and travels to the GPU box with the snapshot, like libpinyon_cuda.so.
"""
from __future__ import annotations

import os
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
REFERENCE = Path(os.environ.get("PINYON_REFERENCE", "/root/reference"))
BIN = HERE / "_bin" / "test_dropin"
BENCH_BIN = HERE / "_bin" / "bench_search"  # bench.py --workload search_offpolicy (BASELINE configs[4])
CALLERS_BIN = HERE / "_bin" / "test_callers"  # MatrixUCB, FullTraversal, TreeBanditFlat over the GPU model (SURVEY 8(f)1,3)
LRS_SHIM = HERE / "lrs_shim"  # stand-in for the absent extern/lrslib submodule (test infrastructure)
HEADERS = [*sorted((ROOT / "include" / "pinyon_cuda").glob("*.hh")), ROOT / "include" / "pinyon_cuda.h"]
SOURCES = [HERE / "test_dropin.cc", *HEADERS]


def reference_available() -> bool:
    return (REFERENCE / "src" / "model" / "monte-carlo-model.hh").exists()


def is_stale(binary: Path = BIN, source: Path = HERE / "test_dropin.cc") -> bool:
    if not binary.exists():
        return True
    t = binary.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [source, *HEADERS])


def _compile(binary: Path, source: Path, force: bool, extra_sources=(), extra_includes=()) -> Path:
    if not reference_available():
        raise RuntimeError(f"{REFERENCE} is not present: {binary.name} can only be built next to the reference")
    if not force and not is_stale(binary, source):
        return binary
    binary.parent.mkdir(exist_ok=True)
    inc = []
    for d in extra_includes:
        inc += ["-I", str(d)]
    cmd = ["g++", "-std=c++23", "-O2", "-pthread", "-I", str(ROOT / "oracle" / "shim"), *inc, "-I", str(REFERENCE / "src"),
           "-I", str(ROOT / "include"), "-o", str(binary), str(source), *map(str, extra_sources), "-L",
           str(ROOT / "pinyon_b200"), "-lpinyon_cuda", "-Wl,-rpath,$ORIGIN/../../../pinyon_b200"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed:\n" + res.stderr[-4000:])
    return binary


def build(force: bool = False) -> Path:
    return _compile(BIN, HERE / "test_dropin.cc", force)


def build_bench(force: bool = False) -> Path:
    return _compile(BENCH_BIN, HERE / "bench_search.cc", force)


def build_callers(force: bool = False) -> Path:
    # libpinyon/lrslib.hh includes "../../extern/lrslib/include/lib.h": resolved against lrs_shim/anchor/here
    return _compile(CALLERS_BIN, HERE / "test_callers.cc", force, extra_sources=[LRS_SHIM / "lrs_shim.cc"],
                    extra_includes=[LRS_SHIM / "anchor" / "here", LRS_SHIM])


if __name__ == "__main__":
    print(build(force=True))
    print(build_bench(force=True))
    print(build_callers(force=True))
