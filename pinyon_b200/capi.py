"""ctypes binding of libpinyon_cuda's C ABI (include/pinyon_cuda.h).

This is plumbing for tests/ and bench.py; the drop-in a Pinyon user links is the
header-only C++ model in include/pinyon_cuda/monte_carlo_model.hh.  There is no CPU
path here: if the library is missing or no sm_100 device is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import weakref
from pathlib import Path
from typing import Optional

import numpy as np

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libpinyon_cuda.so"

PCU_OK = 0
PCU_ERR_INVALID = -1
PCU_ERR_CUDA = -2
PCU_ERR_UNSUPPORTED = -3
PCU_MEM_HOST = 0
PCU_MEM_DEVICE = 1
PCU_MODEL_STREAM = 0
PCU_MODEL_PER_ROLLOUT = 1
PCU_MODEL_WORDS = 2
PCU_ENGINE_MT19937 = 0
PCU_ENGINE_XOR_SHIFT = 1
PCU_GROWTH_DEFAULT = 0
PCU_GROWTH_ALTERNATING_CONSTANT = 1

HASH_INIT = 2166136261
HASH_MUL = 16777619


def unpack2(packed: np.ndarray, n: int) -> np.ndarray:
    """pcu_out.payoff_packed2 -> one payoff_x2 byte per rollout."""
    sh = np.arange(4, dtype=np.uint8) * 2
    return ((packed[:, None] >> sh[None, :]) & 3).astype(np.uint8).reshape(-1)[:n]

# every symbol include/pinyon_cuda.h declares
EXPORTS = [
    "pcu_create", "pcu_destroy", "pcu_last_error", "pcu_version", "pcu_set_stream", "pcu_kernel_launches", "pcu_last_kernel_ms",
    "pcu_set_option", "pcu_host_alloc", "pcu_host_free", "pcu_mt_seed", "pcu_mt_discard", "pcu_mt_next", "pcu_mt_discard_family",
    "pcu_rollout_random_tree", "pcu_rollout_mold", "pcu_last_path", "pcu_measure_int_peak",
    "pcu_create_multi", "pcu_device_count", "pcu_get_stat", "pcu_warmup", "pcu_visible_devices", "pcu_mt_take_raw", "pcu_xs_discard",
    "pcu_leaf_batch_noalloc",
]


class ModelRng(C.Structure):
    _fields_ = [("mode", C.c_int32), ("mt_x", C.c_void_p), ("mt_p", C.c_void_p), ("seeds", C.c_void_p),
                ("words", C.c_void_p), ("engine", C.c_uint32), ("xs_state", C.c_void_p)]


class RtBatch(C.Structure):
    _fields_ = [
        ("n", C.c_uint64), ("mem", C.c_int32),
        ("state_seed", C.c_void_p), ("state_seed_u", C.c_uint64),
        ("state_pos", C.c_void_p), ("state_pos_u", C.c_uint32),
        ("transition_seed", C.c_void_p), ("transition_seed_u", C.c_uint64),
        ("depth", C.c_void_p), ("depth_u", C.c_uint16),
        ("payoff_bias", C.c_void_p), ("payoff_bias_u", C.c_int32),
        ("rows", C.c_uint32), ("cols", C.c_uint32), ("transitions", C.c_uint32),
        ("thr_p", C.c_int32), ("thr_q", C.c_int32), ("growth", C.c_uint32),
    ]


class MoldBatch(C.Structure):
    _fields_ = [("n", C.c_uint64), ("mem", C.c_int32), ("depth", C.c_void_p), ("depth_u", C.c_uint16),
                ("n_actions", C.c_uint32)]


class Out(C.Structure):
    _fields_ = [("mem", C.c_int32), ("payoff_x2", C.c_void_p), ("traj", C.c_void_p), ("traj_stride", C.c_uint32),
                ("traj_hash", C.c_void_p), ("sums", C.c_void_p), ("payoff_packed2", C.c_void_p)]


_lib = None


def load_library(build_if_missing: bool = True) -> C.CDLL:
    """Load libpinyon_cuda.so (building it in-tree first if nvcc is available)."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing:
        from . import build as _build

        try:
            _build.build_library()
        except RuntimeError:
            if not LIB_PATH.exists():
                raise
    if not LIB_PATH.exists():
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m pinyon_b200.build` (no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    vp, u64, i32, u32 = C.c_void_p, C.c_uint64, C.c_int32, C.c_uint32
    lib.pcu_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.pcu_create.restype = C.c_int
    lib.pcu_destroy.argtypes = [vp]
    lib.pcu_destroy.restype = None
    lib.pcu_last_error.argtypes = [vp]
    lib.pcu_last_error.restype = C.c_char_p
    lib.pcu_last_path.argtypes = [vp]
    lib.pcu_last_path.restype = C.c_char_p
    lib.pcu_version.restype = C.c_int
    lib.pcu_set_stream.argtypes = [vp, vp]
    lib.pcu_kernel_launches.argtypes = [vp]
    lib.pcu_kernel_launches.restype = u64
    lib.pcu_last_kernel_ms.argtypes = [vp]
    lib.pcu_last_kernel_ms.restype = C.c_double
    lib.pcu_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    lib.pcu_host_alloc.argtypes = [vp, C.c_size_t]
    lib.pcu_host_alloc.restype = vp
    lib.pcu_host_free.argtypes = [vp, vp]
    lib.pcu_host_free.restype = None
    lib.pcu_mt_seed.argtypes = [u64, vp, vp]
    lib.pcu_mt_seed.restype = None
    lib.pcu_mt_discard.argtypes = [vp, vp, u64]
    lib.pcu_mt_discard_family.argtypes = [vp, vp, u64, u64]
    lib.pcu_mt_take_raw.argtypes = [vp, vp, u64, vp]
    lib.pcu_xs_discard.argtypes = [u64, u64]
    lib.pcu_xs_discard.restype = u64
    lib.pcu_mt_next.argtypes = [vp, vp]
    lib.pcu_mt_next.restype = u32
    lib.pcu_rollout_random_tree.argtypes = [vp, C.POINTER(RtBatch), C.POINTER(ModelRng), C.POINTER(Out)]
    lib.pcu_rollout_mold.argtypes = [vp, C.POINTER(MoldBatch), C.POINTER(ModelRng), C.POINTER(Out)]
    lib.pcu_measure_int_peak.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.pcu_create_multi.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
    lib.pcu_create_multi.restype = C.c_int
    lib.pcu_device_count.argtypes = [vp]
    lib.pcu_visible_devices.restype = C.c_int
    lib.pcu_get_stat.argtypes = [vp, C.c_char_p, C.POINTER(C.c_double)]
    lib.pcu_warmup.argtypes = [vp, u64, u32]
    lib.pcu_leaf_batch_noalloc.argtypes = [vp, C.POINTER(RtBatch), vp, vp, vp]
    lib.pcu_leaf_batch_noalloc.restype = C.c_int
    _lib = lib
    return lib


class PcuError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libpinyon_cuda error {code}: {msg}")
        self.code = code


class Engine:
    """Host image of a model PRNG: std::mt19937 in libstdc++'s (x[624], p) form."""

    def __init__(self, seed: int = 0):
        self.x = np.zeros(624, np.uint32)
        self.p = np.zeros(1, np.uint32)
        load_library().pcu_mt_seed(seed & 0xFFFFFFFFFFFFFFFF, self.x.ctypes.data, self.p.ctypes.data)

    def discard(self, n: int) -> "Engine":
        rc = load_library().pcu_mt_discard(self.x.ctypes.data, self.p.ctypes.data, n)
        if rc:
            raise PcuError(rc, "pcu_mt_discard")
        return self

    def next(self) -> int:
        return load_library().pcu_mt_next(self.x.ctypes.data, self.p.ctypes.data)

    def take_raw(self, n: int) -> np.ndarray:
        """The next n RAW state words (pcu_mt_take_raw): what a mode-W rollout of n / 4 steps consumes."""
        out = np.zeros(n, np.uint32)
        rc = load_library().pcu_mt_take_raw(self.x.ctypes.data, self.p.ctypes.data, n, out.ctypes.data)
        if rc:
            raise PcuError(rc, "pcu_mt_take_raw")
        return out

    def copy(self) -> "Engine":
        e = Engine.__new__(Engine)
        e.x = self.x.copy()
        e.p = self.p.copy()
        return e


def _ptr(a) -> Optional[int]:
    """Address of a numpy array or torch tensor (None passes through)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def _is_device(*arrays) -> bool:
    dev = [a for a in arrays if a is not None and not isinstance(a, np.ndarray) and a.is_cuda]
    host = [a for a in arrays if a is not None and (isinstance(a, np.ndarray) or not a.is_cuda)]
    if dev and host:
        raise ValueError("mixing host and device arrays in one call")
    return bool(dev)


class Context:
    """One libpinyon_cuda context (one GPU)."""

    def __init__(self, device=0):
        """device: one CUDA device index (pcu_create) or a sequence of them (pcu_create_multi: one context over
        several GPUs, batches sharded by rollout index inside the library)."""
        self.lib = load_library()
        h = C.c_void_p()
        if isinstance(device, (list, tuple)):
            ids = (C.c_int * len(device))(*device)
            rc = self.lib.pcu_create_multi(ids, len(device), C.byref(h))
        else:
            rc = self.lib.pcu_create(device, C.byref(h))
        if rc != PCU_OK:
            raise PcuError(rc, f"pcu_create(device={device}) failed: no sm_100 GPU? (there is no CPU fallback)")
        self.h = h
        self._pinned = []

    def close(self):
        if getattr(self, "h", None):
            self.lib.pcu_destroy(self.h)
            self.h = None

    @property
    def device_count(self) -> int:
        return self.lib.pcu_device_count(self.h)

    def stat(self, key: str) -> float:
        v = C.c_double()
        self._check(self.lib.pcu_get_stat(self.h, key.encode(), C.byref(v)))
        return v.value

    def warmup(self, n: int, depth: int):
        self._check(self.lib.pcu_warmup(self.h, n, depth))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc != PCU_OK:
            raise PcuError(rc, self.lib.pcu_last_error(self.h).decode())

    def set_option(self, key: str, value: int):
        self._check(self.lib.pcu_set_option(self.h, key.encode(), value))

    def set_stream(self, cuda_stream: int):
        self._check(self.lib.pcu_set_stream(self.h, cuda_stream))

    @property
    def launches(self) -> int:
        return self.lib.pcu_kernel_launches(self.h)

    @property
    def last_kernel_ms(self) -> float:
        return self.lib.pcu_last_kernel_ms(self.h)

    @property
    def last_path(self) -> str:
        return self.lib.pcu_last_path(self.h).decode()

    def int_peak(self, variant: int = 1, reps: int = 10) -> float:
        v = C.c_double()
        self._check(self.lib.pcu_measure_int_peak(self.h, variant, reps, C.byref(v)))
        return v.value

    def pinned(self, shape, dtype) -> np.ndarray:
        """A numpy array backed by pinned host memory from pcu_host_alloc."""
        dtype = np.dtype(dtype)
        count = int(np.prod(shape))
        nbytes = max(1, count * dtype.itemsize)
        p = self.lib.pcu_host_alloc(self.h, nbytes)
        if not p:
            raise MemoryError("pcu_host_alloc failed")
        buf = (C.c_uint8 * nbytes).from_address(p)
        arr = np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)
        arr.flags.writeable = True
        # cudaFreeHost once the last view of the buffer is gone (the ctypes object is the base of every view)
        weakref.finalize(buf, self.lib.pcu_host_free, None, p)
        return arr

    # -- the hot path ---------------------------------------------------------------
    def _rng(self, engine, model_seeds, model_words=None, xs_state=None, xor_shift=False):
        r = ModelRng()
        if xs_state is not None or xor_shift:  # Pinyon's xor_shift PRNG: xs_state = np.uint64[1] (mode S) or seeds (mode P)
            r.engine = PCU_ENGINE_XOR_SHIFT
            if xs_state is not None:
                r.mode = PCU_MODEL_STREAM
                r.xs_state = xs_state.ctypes.data
            else:
                r.mode = PCU_MODEL_PER_ROLLOUT
                r.seeds = _ptr(model_seeds)
            return r
        if model_words is not None:
            r.mode = PCU_MODEL_WORDS
            r.words = _ptr(model_words)
        elif model_seeds is not None:
            r.mode = PCU_MODEL_PER_ROLLOUT
            r.seeds = _ptr(model_seeds)
        else:
            if engine is None:
                raise ValueError("mode S needs an Engine")
            r.mode = PCU_MODEL_STREAM
            r.mt_x = engine.x.ctypes.data
            r.mt_p = engine.p.ctypes.data
        return r

    @staticmethod
    def _out(device: bool, payoff, traj, traj_stride, traj_hash, sums, packed=None):
        o = Out()
        o.payoff_packed2 = _ptr(packed)
        o.mem = PCU_MEM_DEVICE if device else PCU_MEM_HOST
        o.payoff_x2 = _ptr(payoff)
        o.traj = _ptr(traj)
        o.traj_stride = traj_stride
        o.traj_hash = _ptr(traj_hash)
        o.sums = _ptr(sums)
        return o

    def rollout_random_tree(self, n, *, rows, cols, transitions=1, thr=(0, 1), depth=None, state_seed=0,
                            state_pos=None, transition_seed=0, payoff_bias=0, engine: Engine = None,
                            model_seeds=None, payoff=None, traj=None, traj_stride=0, traj_hash=None, sums=None,
                            payoff_packed=None, model_words=None, growth=0, xs_state=None, xor_shift=False):
        """pcu_rollout_random_tree.  Scalar arguments are batch-uniform; numpy arrays
        (host) or torch CUDA tensors (device) are per-rollout.  state_pos defaults to a
        freshly constructed root: 2*rows*cols*transitions draws."""
        if state_pos is None:
            state_pos = 2 * rows * cols * transitions
        b = RtBatch()
        b.n = n
        arrays = []

        def put(name, val, dtype):
            if np.isscalar(val) or isinstance(val, int):
                setattr(b, name + "_u", int(val))
                setattr(b, name, None)
            else:
                if isinstance(val, np.ndarray):
                    assert val.dtype == dtype, (name, val.dtype, dtype)
                arrays.append(val)
                setattr(b, name, _ptr(val))

        put("state_seed", state_seed, np.uint64)
        put("state_pos", state_pos, np.uint32)
        put("transition_seed", transition_seed, np.uint64)
        put("depth", depth, np.uint16)
        put("payoff_bias", payoff_bias, np.int32)
        b.rows, b.cols, b.transitions = rows, cols, transitions
        b.thr_p, b.thr_q = thr
        b.growth = growth
        device = _is_device(*arrays, model_seeds, model_words, payoff, traj, traj_hash, sums, payoff_packed)
        b.mem = PCU_MEM_DEVICE if device else PCU_MEM_HOST
        r = self._rng(engine, model_seeds, model_words, xs_state, xor_shift)
        o = self._out(device, payoff, traj, traj_stride, traj_hash, sums, payoff_packed)
        self._check(self.lib.pcu_rollout_random_tree(self.h, C.byref(b), C.byref(r), C.byref(o)))

    def leaf_batch_noalloc(self, n, *, rows, cols, transitions=1, thr=(0, 1), depth, state_seed=0, state_pos=None,
                           transition_seed=0, payoff_bias=0, words, payoff=None, sums=None):
        """pcu_leaf_batch_noalloc: leaves of ONE state, mode W, no device allocation inside the library.  `words`,
        `payoff`, `sums` are torch CUDA tensors (uint32-as-int32 / uint8 / int64 storage) owned by the caller."""
        b = RtBatch()
        b.n = n
        b.state_seed_u, b.transition_seed_u = int(state_seed), int(transition_seed)
        b.state_pos_u = 2 * rows * cols * transitions if state_pos is None else int(state_pos)
        b.depth_u, b.payoff_bias_u = int(depth), int(payoff_bias)
        b.rows, b.cols, b.transitions = rows, cols, transitions
        b.thr_p, b.thr_q = thr
        b.mem = PCU_MEM_DEVICE
        self._check(self.lib.pcu_leaf_batch_noalloc(self.h, C.byref(b), _ptr(words), _ptr(payoff), _ptr(sums)))

    def rollout_mold(self, n, *, n_actions, depth, engine: Engine = None, model_seeds=None, payoff=None, traj=None,
                     traj_stride=0, traj_hash=None, sums=None, payoff_packed=None, model_words=None, xs_state=None,
                     xor_shift=False):
        b = MoldBatch()
        b.n = n
        arrays = []
        if np.isscalar(depth) or isinstance(depth, int):
            b.depth_u = int(depth)
            b.depth = None
        else:
            arrays.append(depth)
            b.depth = _ptr(depth)
        b.n_actions = n_actions
        device = _is_device(*arrays, model_seeds, model_words, payoff, traj, traj_hash, sums, payoff_packed)
        b.mem = PCU_MEM_DEVICE if device else PCU_MEM_HOST
        r = self._rng(engine, model_seeds, model_words, xs_state, xor_shift)
        o = self._out(device, payoff, traj, traj_stride, traj_hash, sums, payoff_packed)
        self._check(self.lib.pcu_rollout_mold(self.h, C.byref(b), C.byref(r), C.byref(o)))
