"""ctypes front-end of the CPU oracle (oracle/sib_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package never does (tests/test_layout.py greps for it).

PARITY UNPINNED: the reference holds no golden vectors for this path; see sib_oracle.c header.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libsiboracle.so")

EJSON, ESHAPE, ECAP = -1, -2, -3


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "sib_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.sibo_u01.restype = C.c_float
        L.sibo_u01.argtypes = [C.c_uint64, C.c_uint64]
        L.sibo_fill_u01.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
        L.sibo_make_preset.restype = C.c_long
        L.sibo_make_preset.argtypes = [C.c_int, C.c_void_p, C.c_size_t]
        L.sibo_model_parse.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.sibo_forward_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_int]
        L.sibo_forward_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p]
        L.sibo_json_decode.restype = C.c_long
        L.sibo_json_decode.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_long, C.c_void_p]
        L.sibo_float_to_string.argtypes = [C.c_float, C.c_char_p]
        L.sibo_json_encode.restype = C.c_long
        L.sibo_json_encode.argtypes = [C.c_void_p, C.c_long, C.c_int, C.c_char_p, C.c_size_t]
        L.sibo_execute.restype = C.c_long
        L.sibo_execute.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        L.sibo_run_tuples.restype = C.c_long
        L.sibo_run_tuples.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_long, C.c_int, C.c_void_p]
        L.sibo_sizeof_model.restype = C.c_size_t
        L.sibo_latency_run.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int, C.c_int,
                                       C.c_double, C.c_long, C.c_int, C.c_void_p]
        L.sibo_conv2d_valid.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.sibo_maxpool2.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.sibo_dense.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.sibo_softmax.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def u01(seed: int, start: int, n: int) -> np.ndarray:
    out = np.empty(n, np.float32)
    lib().sibo_fill_u01(seed, start, n, _p(out))
    return out


def make_preset(preset: int) -> bytes:
    n = lib().sibo_make_preset(preset, None, 0)
    assert n > 0
    buf = C.create_string_buffer(n)
    assert lib().sibo_make_preset(preset, buf, n) == n
    return buf.raw


class Model:
    """Parsed model blob (keeps the blob alive: the C struct points into it)."""

    def __init__(self, blob: bytes):
        self._blob = C.create_string_buffer(blob, len(blob))
        self._m = C.create_string_buffer(lib().sibo_sizeof_model())
        rc = lib().sibo_model_parse(self._blob, len(blob), self._m)
        if rc != 0:
            raise ValueError(f"oracle: bad model blob ({rc})")
        hdr = np.frombuffer(blob, np.uint32, 8)
        self.h, self.w, self.c = int(hdr[2]), int(hdr[3]), int(hdr[4])
        m32 = np.frombuffer(self._m.raw, np.uint32, 5)
        self.n_classes = int(m32[4])

    def forward(self, x: np.ndarray, threads: int = 1):
        """x: [N,H,W,C] fp32 -> (logits, probs) fp32 [N,ncls]."""
        x = np.ascontiguousarray(x, np.float32)
        n = x.shape[0]
        lg = np.empty((n, self.n_classes), np.float32)
        pr = np.empty((n, self.n_classes), np.float32)
        rc = lib().sibo_forward_f32(self._m, _p(x), n, _p(lg), _p(pr), threads)
        assert rc == 0
        return lg, pr

    def forward64(self, x: np.ndarray):
        x = np.ascontiguousarray(x, np.float32)
        n = x.shape[0]
        lg = np.empty((n, self.n_classes), np.float64)
        pr = np.empty((n, self.n_classes), np.float64)
        assert lib().sibo_forward_f64(self._m, _p(x), n, _p(lg), _p(pr)) == 0
        return lg, pr

    def execute(self, json_in: bytes):
        """Mirror of InferenceBolt.execute(): returns the output message, None (emit null), or raises on shape."""
        out = C.create_string_buffer(max(4096, len(json_in)))
        r = lib().sibo_execute(self._m, json_in, len(json_in), out, len(out))
        if r == EJSON:
            return None
        if r == ESHAPE:
            raise ValueError("IllegalArgumentException: instances shape does not match model input")
        assert r > 0, r
        return out.raw[:r]

    def run_tuples(self, tuples: list[bytes], threads: int = 0):
        offs = np.zeros(len(tuples) + 1, np.int64)
        offs[1:] = np.cumsum([len(t) for t in tuples])
        blob = b"".join(tuples)
        ob = C.c_long(0)
        ok = lib().sibo_run_tuples(self._m, blob, _p(offs), len(tuples), threads, C.byref(ob))
        return ok, ob.value


def latency_run(model: "Model", payloads, rate_per_s: float, n_tuples: int, threads: int = 4, json_mode: bool = True):
    """Open-loop latency of the CPU path (see sibo_latency_run). -> int64 ndarray of per-tuple latencies in ns."""
    keep, ptrs, lens = [], (C.c_void_p * len(payloads))(), (C.c_size_t * len(payloads))()
    for i, p in enumerate(payloads):
        if isinstance(p, (bytes, bytearray)):
            b = C.create_string_buffer(bytes(p), len(p)); keep.append(b)
            ptrs[i] = C.cast(b, C.c_void_p); lens[i] = len(p)
        else:
            a = np.ascontiguousarray(p, np.float32); keep.append(a)
            ptrs[i] = a.ctypes.data; lens[i] = a.nbytes
    lat = np.empty(n_tuples, np.int64)
    lib().sibo_latency_run(model._m, ptrs, lens, len(payloads), 0 if json_mode else 1, rate_per_s, n_tuples, threads, _p(lat))
    return lat


def json_decode(text: bytes, cap: int | None = None):
    """-> (array [N,H,W,C] fp32) or negative error code."""
    cap = cap or (len(text) // 2 + 8)
    out = np.empty(cap, np.float32)
    dims = (C.c_long * 4)()
    r = lib().sibo_json_decode(text, len(text), _p(out), cap, dims)
    if r < 0:
        return int(r)
    return out[:r].reshape(tuple(dims)).copy()


def float_to_string(x) -> str:
    buf = C.create_string_buffer(32)
    n = lib().sibo_float_to_string(C.c_float(float(np.float32(x))), buf)
    return buf.raw[:n].decode()


def json_encode(probs: np.ndarray) -> bytes:
    probs = np.ascontiguousarray(probs, np.float32)
    n, k = probs.shape
    cap = 64 + n * (k * 18 + 4)
    buf = C.create_string_buffer(cap)
    r = lib().sibo_json_encode(_p(probs), n, k, buf, cap)
    assert r > 0
    return buf.raw[:r]


def max_threads() -> int:
    return int(lib().sibo_max_threads())


# --- single ops for known-answer tests ---------------------------------------------------------

def conv2d_valid(x: np.ndarray, f: np.ndarray) -> np.ndarray:
    h, w, cin = x.shape
    kh, kw, _, cout = f.shape
    out = np.empty((h - kh + 1, w - kw + 1, cout), np.float32)
    lib().sibo_conv2d_valid(_p(np.ascontiguousarray(x, np.float32)), h, w, cin,
                            _p(np.ascontiguousarray(f, np.float32)), kh, kw, cout, _p(out))
    return out


def maxpool2(x: np.ndarray) -> np.ndarray:
    h, w, c = x.shape
    out = np.empty((h // 2, w // 2, c), np.float32)
    lib().sibo_maxpool2(_p(np.ascontiguousarray(x, np.float32)), h, w, c, _p(out))
    return out


def dense(x: np.ndarray, wt: np.ndarray) -> np.ndarray:
    k, m = wt.shape
    out = np.empty(m, np.float32)
    lib().sibo_dense(_p(np.ascontiguousarray(x, np.float32)), k, _p(np.ascontiguousarray(wt, np.float32)), m, _p(out))
    return out


def softmax(x: np.ndarray) -> np.ndarray:
    out = np.empty(len(x), np.float32)
    lib().sibo_softmax(_p(np.ascontiguousarray(x, np.float32)), len(x), _p(out))
    return out
