"""Python face of the bolt mirror (csrc/sib_bolt.cpp) for tests and the latency bench.

Same names as the reference's bolt API (InferenceBolt.java:43-105): prepare / execute / declareOutputFields /
cleanup.  All logic is native; this class only marshals arguments to the harness entry points in include/sib_bolt.h.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

from . import native

_lib = native.lib
_lib.sib_harness_open.restype = C.c_int
_lib.sib_harness_open.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
_lib.sib_harness_execute.restype = C.c_int
_lib.sib_harness_execute.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p, C.c_size_t]
_lib.sib_harness_cleanup.restype = C.c_int
_lib.sib_harness_cleanup.argtypes = [C.c_void_p]
_lib.sib_harness_close.restype = C.c_int
_lib.sib_harness_close.argtypes = [C.c_void_p]
_lib.sib_harness_event_count.restype = C.c_long
_lib.sib_harness_event_count.argtypes = [C.c_void_p]
_lib.sib_harness_event.restype = C.c_int
_lib.sib_harness_event.argtypes = [C.c_void_p, C.c_long, C.POINTER(C.c_int), C.POINTER(C.c_uint64),
                                   C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_int64)]
_lib.sib_harness_output_fields.restype = C.c_int
_lib.sib_harness_output_fields.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
_lib.sib_harness_wait_events.restype = C.c_long
_lib.sib_harness_wait_events.argtypes = [C.c_void_p, C.c_long, C.c_long]

HARNESS_SYMBOLS = ["sib_harness_open", "sib_harness_execute", "sib_harness_cleanup", "sib_harness_close",
                   "sib_harness_event_count", "sib_harness_event", "sib_harness_output_fields",
                   "sib_harness_wait_events"]

EV_EMIT, EV_ACK = 1, 2


@dataclass
class Tuple:
    """org.apache.storm.tuple.Tuple as execute() sees it: field 0 is the JSON string (InferenceBolt.java:71)."""
    id: int
    value: bytes

    def getString(self, i: int) -> bytes:
        assert i == 0
        return self.value


class OutputCollector:
    """Recording collector: `events` is the ordered list of ("emit", message|None) / ("ack", tuple_id)."""

    def __init__(self):
        self.events = []


class IllegalArgumentException(ValueError):
    pass


class InferenceBolt:
    def __init__(self, model_blob: bytes):
        self._blob = model_blob
        self._h = C.c_void_p()
        self._collector = None
        self._seen = 0

    def prepare(self, conf: dict, task_index: int, collector: OutputCollector):
        text = "".join(f"{k}={v}\n" for k, v in conf.items()).encode()
        rc = _lib.sib_harness_open(self._blob, len(self._blob), text, task_index, C.byref(self._h))
        if rc:
            raise native.SibError(rc, "InferenceBolt.prepare")
        self._collector = collector

    def execute(self, tup: Tuple):
        s = tup.getString(0)
        rc = _lib.sib_harness_execute(self._h, tup.id, s, len(s))
        if rc == native.ESHAPE:
            raise IllegalArgumentException(native.last_error())
        if rc:
            raise native.SibError(rc, "InferenceBolt.execute")

    def declareOutputFields(self) -> list[str]:
        buf = C.create_string_buffer(256)
        n = _lib.sib_harness_output_fields(self._h, buf, len(buf))
        return buf.raw[:n].decode().split(",")

    def sync_events(self, wait_for: int = 0, timeout_us: int = 5_000_000, with_time: bool = False):
        """Copies newly recorded collector events into collector.events."""
        if wait_for:
            _lib.sib_harness_wait_events(self._h, wait_for, timeout_us)
        n = _lib.sib_harness_event_count(self._h)
        kind, tid, msg, ln, t = C.c_int(), C.c_uint64(), C.c_char_p(), C.c_size_t(), C.c_int64()
        for i in range(self._seen, n):
            _lib.sib_harness_event(self._h, i, C.byref(kind), C.byref(tid), C.byref(msg), C.byref(ln), C.byref(t))
            if kind.value == EV_EMIT:
                m = C.string_at(msg, ln.value) if msg.value is not None else None
                ev = ("emit", m)
            else:
                ev = ("ack", int(tid.value))
            self._collector.events.append(ev + (int(t.value),) if with_time else ev)
        self._seen = n
        return self._collector.events

    def cleanup(self):
        if self._h:
            _lib.sib_harness_cleanup(self._h)
            self.sync_events()

    def close(self):
        if self._h:
            _lib.sib_harness_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
