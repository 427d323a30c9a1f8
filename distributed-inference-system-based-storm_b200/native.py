"""ctypes binding of the C ABI in include/sib.h.  No compute here and no fallback: if libsib.so is missing the
import fails, and without a CUDA device Engine() raises (SIB_ECUDA)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsib.so")

OK, EJSON, ESHAPE, ECAP, EAGAIN, EINVAL, ECUDA, ECLOSED = 0, -1, -2, -3, -4, -5, -6, -7
PREC_FP16, PREC_BF16, PREC_FP32 = 0, 1, 2
STATIC_MAX_IMAGES = 16384

if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is not built: run `python __graft_entry__.py build` (nvcc, sm_100a). "
                      "There is no Python or CPU fallback for the InferenceBolt hot path.")

lib = C.CDLL(LIB_PATH)


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("task_index", C.c_int32), ("max_batch", C.c_int32),
                ("deadline_us", C.c_int32), ("precision", C.c_int32), ("n_slots", C.c_int32),
                ("guard", C.c_int32), ("guard_tau", C.c_float), ("json_on_gpu", C.c_int32), ("strict_batching", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("tag", C.c_uint64), ("n", C.c_int32), ("n_classes", C.c_int32), ("status", C.c_int32),
                ("reserved", C.c_int32), ("probs", C.POINTER(C.c_float)), ("argmax", C.POINTER(C.c_int32))]


class Stats(C.Structure):
    _fields_ = [("tuples", C.c_uint64), ("images", C.c_uint64), ("batches", C.c_uint64),
                ("flush_by_size", C.c_uint64), ("flush_by_deadline", C.c_uint64), ("flush_explicit", C.c_uint64),
                ("guard_recomputed", C.c_uint64), ("batch_hist", C.c_uint64 * 16),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("json_gpu_decoded", C.c_uint64), ("json_host_redone", C.c_uint64)]


class ModelInfo(C.Structure):
    _fields_ = [("in_h", C.c_int32), ("in_w", C.c_int32), ("in_c", C.c_int32), ("n_layers", C.c_int32),
                ("n_classes", C.c_int32), ("n_weights", C.c_int32), ("flops_per_image", C.c_int64),
                ("bytes_per_image", C.c_int64)]


# every symbol include/sib.h declares (tests/test_abi.py checks the header against this table)
SYMBOLS = {
    "sib_device_count": (C.c_int, []),
    "sib_model_describe": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(ModelInfo)]),
    "sib_model_import_pb": (C.c_long, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "sib_engine_open": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "sib_engine_close": (C.c_int, [C.c_void_p]),
    "sib_engine_info": (C.c_int, [C.c_void_p, C.POINTER(ModelInfo), C.POINTER(Config)]),
    "sib_client_open": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "sib_client_close": (C.c_int, [C.c_void_p]),
    "sib_client_submit_json": (C.c_int, [C.c_void_p, C.c_uint64, C.c_char_p, C.c_size_t]),
    "sib_client_submit_f32": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]),
    "sib_client_poll": (C.c_int, [C.c_void_p, C.POINTER(Result), C.c_int, C.c_int]),
    "sib_client_flush": (C.c_int, [C.c_void_p]),
    "sib_client_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "sib_submit_json": (C.c_int, [C.c_void_p, C.c_uint64, C.c_char_p, C.c_size_t]),
    "sib_submit_f32": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]),
    "sib_submit_pinned": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]),
    "sib_poll": (C.c_int, [C.c_void_p, C.POINTER(Result), C.c_int, C.c_int]),
    "sib_poll_json_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "sib_flush": (C.c_int, [C.c_void_p]),
    "sib_format_json": (C.c_long, [C.POINTER(Result), C.c_char_p, C.c_size_t]),
    "sib_stats_get": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "sib_forward_sync": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sib_forward_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sib_forward_device_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sib_debug_stage_dump": (C.c_long, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_long]),
    "sib_json_decode": (C.c_long, [C.c_char_p, C.c_size_t, C.c_void_p, C.c_long, C.POINTER(C.c_int64)]),
    "sib_json_encode": (C.c_long, [C.c_void_p, C.c_long, C.c_int, C.c_char_p, C.c_size_t]),
    "sib_float_to_string": (C.c_int, [C.c_float, C.c_char_p]),
    "sib_last_error": (C.c_char_p, []),
    "sib_strerror": (C.c_char_p, [C.c_int]),
    "sib_version": (C.c_char_p, []),
}
for _name, (_res, _args) in SYMBOLS.items():
    _f = getattr(lib, _name)
    _f.restype = _res
    _f.argtypes = _args


class SibError(RuntimeError):
    def __init__(self, code: int, where: str):
        self.code = code
        msg = lib.sib_last_error().decode(errors="replace")
        super().__init__(f"{where}: {lib.sib_strerror(code).decode()} [{code}] {msg}")


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def last_error() -> str:
    return lib.sib_last_error().decode(errors="replace")


def device_count() -> int:
    return int(lib.sib_device_count())


def describe(blob: bytes) -> ModelInfo:
    info = ModelInfo()
    rc = lib.sib_model_describe(blob, len(blob), C.byref(info))
    if rc:
        raise SibError(rc, "sib_model_describe")
    return info


def import_pb(saved_model_pb: bytes) -> bytes:
    """TensorFlow SavedModel / frozen GraphDef bytes (the reference's model/saved_model.pb) -> model blob."""
    need = lib.sib_model_import_pb(saved_model_pb, len(saved_model_pb), None, 0)
    if need < 0:
        raise SibError(int(need), "sib_model_import_pb")
    buf = C.create_string_buffer(need)
    got = lib.sib_model_import_pb(saved_model_pb, len(saved_model_pb), buf, need)
    if got != need:
        raise SibError(int(got), "sib_model_import_pb")
    return buf.raw


def json_decode(text: bytes, cap: int | None = None):
    """-> ndarray [N,H,W,C] fp32, or the negative error code."""
    cap = cap if cap is not None else len(text) // 2 + 8
    out = np.empty(max(cap, 1), np.float32)
    dims = (C.c_int64 * 4)()
    r = lib.sib_json_decode(text, len(text), _p(out), cap, dims)
    if r < 0:
        return int(r)
    return out[:r].reshape(tuple(dims)).copy()


def json_encode(probs: np.ndarray) -> bytes:
    probs = np.ascontiguousarray(probs, np.float32)
    n, k = probs.shape
    cap = 64 + n * (k * 18 + 4)
    buf = C.create_string_buffer(cap)
    r = lib.sib_json_encode(_p(probs), n, k, buf, cap)
    if r < 0:
        raise SibError(int(r), "sib_json_encode")
    return buf.raw[:r]


def float_to_string(x) -> str:
    buf = C.create_string_buffer(32)
    n = lib.sib_float_to_string(C.c_float(float(np.float32(x))), buf)
    return buf.raw[:n].decode()


class Client:
    """A bolt's view of the process-wide engine shared by co-located bolts (include/sib.h, sib_client_*)."""

    def __init__(self, blob: bytes, img_elems: int, **cfg_kw):
        d = dict(device=-1, task_index=0, max_batch=4096, deadline_us=200, precision=PREC_FP16, n_slots=4, guard=1,
                 guard_tau=0.0, json_on_gpu=0, strict_batching=0)
        d.update(cfg_kw)
        cfg = Config(*[d[k] for k, _ in Config._fields_])
        self._h = C.c_void_p()
        rc = lib.sib_client_open(blob, len(blob), C.byref(cfg), C.byref(self._h))
        if rc:
            raise SibError(rc, "sib_client_open")
        self.img_elems = img_elems
        self._res = (Result * 1024)()

    def close(self):
        if self._h:
            lib.sib_client_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def submit_f32(self, tag: int, x: np.ndarray) -> int:
        x = np.ascontiguousarray(x, np.float32)
        return int(lib.sib_client_submit_f32(self._h, tag, _p(x), x.size // self.img_elems))

    def submit_json(self, tag: int, text: bytes) -> int:
        return int(lib.sib_client_submit_json(self._h, tag, text, len(text)))

    def flush(self):
        lib.sib_client_flush(self._h)

    def stats(self) -> Stats:
        st = Stats()
        lib.sib_client_stats(self._h, C.byref(st))
        return st

    def poll(self, max_results: int = 1024, timeout_us: int = 0):
        """-> list of (tag, probs[n,ncls] copy, argmax[n] copy, status)."""
        n = lib.sib_client_poll(self._h, self._res, min(max_results, 1024), timeout_us)
        if n < 0:
            raise SibError(n, "sib_client_poll")
        out = []
        for i in range(n):
            r = self._res[i]
            if r.n == 0 or not r.probs:
                out.append((int(r.tag), None, None, int(r.status)))
                continue
            out.append((int(r.tag), np.ctypeslib.as_array(r.probs, (r.n, r.n_classes)).copy(),
                        np.ctypeslib.as_array(r.argmax, (r.n,)).copy(), int(r.status)))
        return out


class Engine:
    """One engine = one executor's model replica on one GPU (InferenceBolt.prepare(), InferenceBolt.java:43-62)."""

    def __init__(self, blob: bytes, device: int = -1, task_index: int = 0, max_batch: int = 4096,
                 deadline_us: int = 200, precision: int = PREC_FP16, n_slots: int = 4, guard: bool = True,
                 guard_tau: float = 0.0, json_on_gpu: bool = False, strict_batching: bool = False):
        self._h = C.c_void_p()
        self._blob = blob
        cfg = Config(device, task_index, max_batch, deadline_us, precision, n_slots, 1 if guard else 0, guard_tau,
                     int(json_on_gpu), 1 if strict_batching else 0)
        rc = lib.sib_engine_open(blob, len(blob), C.byref(cfg), C.byref(self._h))
        if rc:
            raise SibError(rc, "sib_engine_open")
        self.info = ModelInfo()
        self.cfg = Config()
        lib.sib_engine_info(self._h, C.byref(self.info), C.byref(self.cfg))
        self.n_classes = int(self.info.n_classes)
        self.img_elems = int(self.info.in_h * self.info.in_w * self.info.in_c)
        self._res = (Result * 4096)()

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h:
            lib.sib_engine_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def submit_json(self, tag: int, text: bytes) -> int:
        return int(lib.sib_submit_json(self._h, tag, text, len(text)))

    def submit_f32(self, tag: int, x: np.ndarray) -> int:
        x = np.ascontiguousarray(x, np.float32)
        n = x.size // self.img_elems
        return int(lib.sib_submit_f32(self._h, tag, _p(x), n))

    def submit_pinned(self, tag: int, ptr: int, n: int) -> int:
        """ptr: address of n images of fp32 NHWC in page-locked host memory (kept alive by the caller)."""
        return int(lib.sib_submit_pinned(self._h, tag, C.c_void_p(ptr), n))

    def flush(self):
        rc = lib.sib_flush(self._h)
        if rc:
            raise SibError(rc, "sib_flush")

    def poll_raw(self, max_results: int = 4096, timeout_us: int = 0):
        n = lib.sib_poll(self._h, self._res, min(max_results, 4096), timeout_us)
        if n < 0:
            raise SibError(n, "sib_poll")
        return n, self._res

    def poll(self, max_results: int = 4096, timeout_us: int = 0):
        """-> list of (tag, probs[n,ncls] copy, argmax[n] copy, status)."""
        n, res = self.poll_raw(max_results, timeout_us)
        out = []
        for i in range(n):
            r = res[i]
            if r.n == 0 or not r.probs:
                out.append((int(r.tag), None, None, int(r.status)))
                continue
            pr = np.ctypeslib.as_array(r.probs, (r.n, r.n_classes)).copy()
            am = np.ctypeslib.as_array(r.argmax, (r.n,)).copy()
            out.append((int(r.tag), pr, am, int(r.status)))
        return out

    def poll_json_batch(self, max_results: int = 256, timeout_us: int = 0, arena_bytes: int = 1 << 16):
        """sib_poll_json_batch -> list of (tag, message bytes | None, status): one native pass per poll."""
        arena = np.empty(arena_bytes, np.uint8)
        tags = np.empty(max_results, np.uint64)
        offs = np.empty(max_results + 1, np.uint32)
        st = np.empty(max_results, np.int32)
        n = lib.sib_poll_json_batch(self._h, _p(arena), arena.size, _p(tags), _p(offs), _p(st), max_results, timeout_us)
        if n < 0:
            raise SibError(n, "sib_poll_json_batch")
        raw = arena.tobytes()
        return [(int(tags[i]), raw[offs[i]:offs[i + 1]] if st[i] == OK and offs[i + 1] > offs[i] else None, int(st[i])) for i in range(n)]

    def poll_json(self, max_results: int = 4096, timeout_us: int = 0):
        """-> list of (tag, message bytes) formatted by the library (PredObj JSON)."""
        n, res = self.poll_raw(max_results, timeout_us)
        out = []
        buf = C.create_string_buffer(1 << 16)
        for i in range(n):
            if res[i].status != OK or res[i].n == 0:
                out.append((int(res[i].tag), None))   # the bolt emits a null message for these
                continue
            need = 64 + res[i].n * (res[i].n_classes * 18 + 4)
            if need > len(buf):
                buf = C.create_string_buffer(need)
            ln = lib.sib_format_json(C.byref(res[i]), buf, len(buf))
            if ln < 0:
                raise SibError(int(ln), "sib_format_json")
            out.append((int(res[i].tag), buf.raw[:ln]))
        return out

    def stats(self) -> Stats:
        s = Stats()
        lib.sib_stats_get(self._h, C.byref(s))
        return s

    def forward_sync(self, x: np.ndarray, want_logits: bool = True):
        x = np.ascontiguousarray(x, np.float32)
        n = x.size // self.img_elems if x.size else 0
        probs = np.empty((n, self.n_classes), np.float32)
        logits = np.empty((n, self.n_classes), np.float32) if want_logits else None
        argmax = np.empty(n, np.int32)
        rc = lib.sib_forward_sync(self._h, _p(x), n, _p(probs), _p(logits) if want_logits else None, _p(argmax))
        if rc:
            raise SibError(rc, "sib_forward_sync")
        return probs, logits, argmax

    def forward_device_static(self, d_in: int, n: int, d_probs: int, d_logits: int = 0, d_argmax: int = 0, stream: int = 0):
        """sib_forward_device_static: guard scratch from the library's static device array; returns the C return code."""
        return int(lib.sib_forward_device_static(self._h, C.c_void_p(d_in), n, C.c_void_p(d_probs), C.c_void_p(d_logits or None),
                                                 C.c_void_p(d_argmax or None), C.c_void_p(stream or None)))

    def stage_dump(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        out = np.zeros(1 << 16, np.float32)
        r = lib.sib_debug_stage_dump(self._h, _p(x), x.size // self.img_elems, _p(out), out.size)
        if r < 0:
            raise SibError(int(r), "sib_debug_stage_dump")
        return out[:r]

    def forward_device(self, d_in: int, n: int, d_probs: int, d_logits: int = 0, d_argmax: int = 0, stream: int = 0):
        rc = lib.sib_forward_device(self._h, C.c_void_p(d_in), n, C.c_void_p(d_probs), C.c_void_p(d_logits or None),
                                    C.c_void_p(d_argmax or None), C.c_void_p(stream or None))
        if rc:
            raise SibError(rc, "sib_forward_device")


# ---- open-loop load generator (include/sib_loadgen.h) -------------------------------------------------------------

LG_JSON, LG_F32 = 0, 1


class LoadgenCfg(C.Structure):
    _fields_ = [("rate_per_s", C.c_double), ("n_tuples", C.c_int64), ("mode", C.c_int32), ("producers", C.c_int32),
                ("poisson", C.c_int32), ("reserved", C.c_int32), ("seed", C.c_uint64)]


class LoadgenReport(C.Structure):
    _fields_ = [("completed", C.c_int64), ("rejected", C.c_int64), ("images", C.c_int64), ("wall_s", C.c_double),
                ("achieved_tuples_per_s", C.c_double), ("p50_ns", C.c_int64), ("p90_ns", C.c_int64),
                ("p99_ns", C.c_int64), ("p999_ns", C.c_int64), ("max_ns", C.c_int64), ("mean_ns", C.c_int64),
                ("eagain_spins", C.c_int64), ("late_submits", C.c_int64)]


LOADGEN_SYMBOLS = ["sib_loadgen_run"]
lib.sib_loadgen_run.restype = C.c_int
lib.sib_loadgen_run.argtypes = [C.c_void_p, C.POINTER(LoadgenCfg), C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int,
                                C.c_void_p, C.POINTER(LoadgenReport)]


def loadgen(engine: "Engine", payloads, rate_per_s: float, n_tuples: int, mode: int = LG_JSON, producers: int = 1,
            poisson: bool = False, seed: int = 99, want_latencies: bool = False):
    """payloads: list of bytes (JSON mode) or contiguous float32 ndarrays (F32 mode). -> dict report (+ latencies)."""
    keep = []
    ptrs = (C.c_void_p * len(payloads))()
    lens = (C.c_size_t * len(payloads))()
    for i, p in enumerate(payloads):
        if isinstance(p, (bytes, bytearray)):
            b = C.create_string_buffer(bytes(p), len(p))
            keep.append(b)
            ptrs[i] = C.cast(b, C.c_void_p)
            lens[i] = len(p)
        else:
            a = np.ascontiguousarray(p, np.float32)
            keep.append(a)
            ptrs[i] = a.ctypes.data
            lens[i] = a.nbytes
    cfg = LoadgenCfg(rate_per_s, n_tuples, mode, producers, 1 if poisson else 0, 0, seed)
    rep = LoadgenReport()
    lat = np.empty(n_tuples, np.int64) if want_latencies else None
    rc = lib.sib_loadgen_run(engine.handle, C.byref(cfg), ptrs, lens, len(payloads), _p(lat) if want_latencies else None,
                             C.byref(rep))
    if rc:
        raise SibError(rc, "sib_loadgen_run")
    out = {f: getattr(rep, f) for f, _ in LoadgenReport._fields_}
    if want_latencies:
        out["latencies_ns"] = lat
    return out
