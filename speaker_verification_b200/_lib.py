"""ctypes binding of libspeechpy_b200.so (include/speechpy_b200.h).

There is no CPU fallback: if the shared library is missing, or no B200 is
visible, the first call raises ``RuntimeError`` -- it never routes to numpy.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SPB_LIBRARY", os.path.join(_HERE, "libspeechpy_b200.so"))

SPB_OK, SPB_ERR_INVALID, SPB_ERR_NEGATIVE_DIM, SPB_ERR_CUDA, SPB_ERR_UNSUPPORTED, SPB_ERR_NOMEM = range(6)
F32, F64, I16, I32 = range(4)

c_i32, c_i64, c_u64 = C.c_int32, C.c_int64, C.c_uint64
c_vp = C.c_void_p


class FeatureParams(C.Structure):
    _fields_ = [("frame_len", c_i32), ("frame_stride", c_i32), ("nfft", c_i32),
                ("num_filters", c_i32), ("num_cepstral", c_i32), ("dc_elimination", c_i32),
                ("cmvn_window", c_i32), ("cmvn_variance", c_i32)]


# name -> (restype, argtypes); mirrors include/speechpy_b200.h one to one
PROTOTYPES = {
    "spb_version": (C.c_int, []),
    "spb_init": (C.c_int, [C.c_int, C.POINTER(c_vp)]),
    "spb_destroy": (C.c_int, [c_vp]),
    "spb_last_error": (C.c_char_p, [c_vp]),
    "spb_sync": (C.c_int, [c_vp]),
    "spb_stream": (C.c_int, [c_vp, C.POINTER(c_u64)]),
    "spb_device_info": (C.c_int, [c_vp, C.POINTER(c_i32), C.POINTER(c_i32), C.POINTER(c_i32), C.POINTER(c_u64)]),
    "spb_launch_count": (C.c_int, [c_vp, C.POINTER(c_u64)]),
    "spb_timing_enable": (C.c_int, [c_vp, c_i32]),
    "spb_timing_read": (C.c_int, [c_vp, c_i32, C.POINTER(C.c_double), C.POINTER(c_u64)]),
    "spb_set_option": (C.c_int, [C.c_char_p, C.c_double]),
    "spb_get_option": (C.c_int, [C.c_char_p, C.POINTER(C.c_double)]),
    "spb_host_alloc": (C.c_int, [C.POINTER(c_vp), C.c_size_t]),
    "spb_host_free": (C.c_int, [c_vp]),
    "spb_num_frames": (C.c_int, [c_i64, c_i32, c_i32, c_i32, C.POINTER(c_i64)]),
    "spb_stack_frames": (C.c_int, [c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_i64, c_vp]),
    "spb_power_spectrum": (C.c_int, [c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp]),
    "spb_plan_create": (C.c_int, [c_vp, C.POINTER(FeatureParams), c_vp, C.POINTER(c_vp)]),
    "spb_plan_destroy": (C.c_int, [c_vp]),
    "spb_plan_set_preemphasis": (C.c_int, [c_vp, C.c_double]),
    "spb_mfe": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_i64, c_vp, c_vp]),
    "spb_lmfe": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_i64, c_vp]),
    "spb_mfcc": (C.c_int, [c_vp, c_vp, c_vp, c_i32, c_i64, c_vp]),
    "spb_delta_cube": (C.c_int, [c_vp, c_vp, c_i64, c_i32, c_vp]),
    "spb_derivative": (C.c_int, [c_vp, c_vp, c_i64, c_i32, c_i32, c_vp]),
    "spb_cmvn": (C.c_int, [c_vp, c_vp, c_i64, c_i32, c_i32, c_vp]),
    "spb_cmvnw": (C.c_int, [c_vp, c_vp, c_i64, c_i32, c_i32, c_i32, c_vp]),
    "spb_preemphasis": (C.c_int, [c_vp, c_vp, c_i32, c_i64, c_i64, C.c_double, c_vp]),
    "spb_batch_create": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, C.POINTER(c_vp)]),
    "spb_batch_destroy": (C.c_int, [c_vp]),
    "spb_batch_total_frames": (C.c_int, [c_vp, C.POINTER(c_i64)]),
    "spb_batch_frame_offsets": (C.c_int, [c_vp, c_vp]),
    "spb_lmfe_cmvnw_batch_dev": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp]),
    "spb_lmfe_batch_dev": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp]),
    "spb_mfcc_delta_batch_dev": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp]),
    "spb_lmfe_cmvnw_batch": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp]),
    "spb_mfcc_delta_batch": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp]),
    "spb_lmfe_cmvnw_batch_dtype": (C.c_int, [c_vp, c_vp, c_vp, c_i32, C.c_double, c_i64, c_vp]),
    "spb_mfcc_delta_batch_dtype": (C.c_int, [c_vp, c_vp, c_vp, c_i32, C.c_double, c_i64, c_vp]),
    "spb_feature_cubes_batch_dev": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_i32, c_i32, c_vp, c_i32, c_i32, c_vp]),
    "spb_feature_cubes_batch": (C.c_int, [c_vp, c_vp, c_vp, c_i32, C.c_double, c_i64, c_i32, c_i32, c_vp, c_i32,
                                          c_i32, c_vp]),
    "spb_cosine_scores": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i32, c_vp]),
    "spb_cosine_scores_dev": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i32, c_vp]),
    "spb_cosine_scores_dev_noalloc": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i32, c_vp]),
    "spb_trial_scores_dev": (C.c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_vp, c_i64, c_vp]),
    "spb_argmax_rows_dev": (C.c_int, [c_vp, c_vp, c_i64, c_i64, c_vp]),
    "spb_eer": (C.c_int, [c_vp, c_vp, c_vp, c_i64, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "spb_eer_dev": (C.c_int, [c_vp, c_vp, c_vp, c_i64, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "spb_roc_curve": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_i64, C.POINTER(c_i64)]),
    "spb_roc_curve_dev": (C.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_i64, C.POINTER(c_i64)]),
}

_lib = None
_lock = threading.RLock()
_contexts: dict = {}


def load():
    """dlopen the library and attach prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m speaker_verification_b200.build` "
                "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)  # AttributeError here == header and library out of sync
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class SpbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def check(rc: int, ctx=None):
    """Translate a status code into the exception type the reference would raise."""
    if rc == SPB_OK:
        return
    msg = load().spb_last_error(ctx)
    msg = msg.decode() if msg else f"libspeechpy_b200 error {rc}"
    if rc == SPB_ERR_NEGATIVE_DIM:
        raise ValueError(msg)  # numpy: "negative dimensions are not allowed"
    if rc == SPB_ERR_INVALID:
        raise AssertionError(msg)
    if rc == SPB_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise SpbError(rc, msg)


def set_option(name: str, value: float):
    """spb_set_option: process-wide kernel selection / guard knobs (see the header)."""
    check(load().spb_set_option(name.encode(), float(value)))


def get_option(name: str) -> float:
    v = C.c_double()
    check(load().spb_get_option(name.encode(), C.byref(v)))
    return v.value


class options:
    """Context manager: set options for the duration of a block, then restore them."""

    def __init__(self, **kw):
        self.kw = kw
        self.old = {}

    def __enter__(self):
        for k, v in self.kw.items():
            self.old[k] = get_option(k)
            set_option(k, v)
        return self

    def __exit__(self, *exc):
        for k, v in self.old.items():
            set_option(k, v)
        return False


class Context:
    """One device + stream; created lazily per (process, device)."""

    def __init__(self, device: int = 0):
        lib = load()
        h = c_vp()
        check(lib.spb_init(device, C.byref(h)))
        self.handle = h
        self.device = device
        self.lib = lib
        self._plans: dict = {}

    def close(self):
        if self.handle:
            for p in self._plans.values():
                self.lib.spb_plan_destroy(p)
            self._plans.clear()
            self.lib.spb_destroy(self.handle)
            self.handle = None

    def plan(self, L, S, nfft, M, C_, dc, bank: np.ndarray | None, cm_win=0, cm_var=0, preemphasis=0.0):
        """Cached feature plan; `bank` is the dense float64 (M, nfft/2+1) filterbank.  preemphasis != 0:
        processing.preemphasis(signal, shift=1, cof=preemphasis) fused into the framing (processing.py:45-58)."""
        key = (L, S, nfft, M, C_, dc, cm_win, cm_var, float(preemphasis),
               None if bank is None else hash(bank.tobytes()))
        p = self._plans.get(key)
        if p is None:
            params = FeatureParams(L, S, nfft, M, C_, dc, cm_win, cm_var)
            h = c_vp()
            bptr = None
            if bank is not None:
                bank = np.ascontiguousarray(bank, dtype=np.float64)
                bptr = bank.ctypes.data
            check(self.lib.spb_plan_create(self.handle, C.byref(params), bptr, C.byref(h)), self.handle)
            if preemphasis:
                check(self.lib.spb_plan_set_preemphasis(h, float(preemphasis)), self.handle)
            self._plans[key] = p = h
        return p

    def launch_count(self) -> int:
        n = c_u64()
        check(self.lib.spb_launch_count(self.handle, C.byref(n)), self.handle)
        return n.value

    def timing_enable(self, on: bool):
        check(self.lib.spb_timing_enable(self.handle, 1 if on else 0), self.handle)

    def timing_read(self, kernel: int):
        ms, n = C.c_double(), c_u64()
        check(self.lib.spb_timing_read(self.handle, kernel, C.byref(ms), C.byref(n)), self.handle)
        return ms.value, n.value

    def stream(self) -> int:
        s = c_u64()
        check(self.lib.spb_stream(self.handle, C.byref(s)), self.handle)
        return s.value

    def sync(self):
        check(self.lib.spb_sync(self.handle), self.handle)

    def device_info(self):
        sm, ma, mi, mem = c_i32(), c_i32(), c_i32(), c_u64()
        check(self.lib.spb_device_info(self.handle, C.byref(sm), C.byref(ma), C.byref(mi), C.byref(mem)),
              self.handle)
        return {"sm_count": sm.value, "cc": (ma.value, mi.value), "total_mem": mem.value}


def default_device() -> int:
    dev = os.environ.get("SPB_DEVICE")
    if dev is not None:
        return int(dev)
    lr = os.environ.get("LOCAL_RANK")
    return int(lr) if lr is not None else 0


def context(device: int | None = None) -> Context:
    """Per-process, per-device context (created after fork: keyed by pid)."""
    device = default_device() if device is None else device
    key = (os.getpid(), device)
    ctx = _contexts.get(key)
    if ctx is None:
        with _lock:
            ctx = _contexts.get(key)
            if ctx is None:
                ctx = _contexts[key] = Context(device)
    return ctx


def ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data
