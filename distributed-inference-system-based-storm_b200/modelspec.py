"""Model presets, the model-blob writer and the deterministic synthetic generators.

Product-side tooling (numpy only): benchmarks and examples build their weights and inputs here.  The oracle has its
own independent C implementation of the same generators (oracle/sib_oracle.c); tests check the two agree bit for bit.

Blob layout -- the artefact that stands where the reference keeps model/saved_model.pb (InferenceBolt.java:49):
    "SIBM" | u32 version=1 | u32 in_h,in_w,in_c | u32 n_layers | u32 n_weights | u32 rsvd
    n_layers x { u32 kind,a,b,c,d,flags,w_off,b_off }     conv: a=kh b=kw c=cin d=cout ; dense: a=in b=out, flags&1=ReLU
    n_weights x f32                                       conv filter HWIO, dense kernel [in,out] (TensorFlow layouts)
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np

KIND_CONV, KIND_DENSE = 1, 2

PRESETS = {
    # name: (id, H, W, C, conv channels, dense widths)
    "mnist-lenet5": (0, 28, 28, 1, (6, 16), (120, 84, 10)),
    "cifar-lenet5": (1, 32, 32, 3, (6, 16), (120, 84, 10)),
    "mnist-wide": (2, 28, 28, 1, (16, 32), (64, 10)),
    "cifar-wide": (3, 32, 32, 3, (16, 32), (64, 10)),
}

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def u01(seed: int, start: int, n: int) -> np.ndarray:
    """Counter-based U[0,1): 24 random bits * 2^-24 -- identical bits in C, numpy and CUDA (D2)."""
    with np.errstate(over="ignore"):
        base = _mix64(np.array([seed], np.uint64))[0]
        idx = np.arange(start, start + n, dtype=np.uint64)
        r = _mix64(base + idx)
    return (r >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)


@dataclass
class Layer:
    kind: int
    a: int
    b: int
    c: int
    d: int
    flags: int
    w_off: int
    b_off: int


@dataclass
class Model:
    name: str
    h: int
    w: int
    c: int
    layers: list
    weights: np.ndarray

    @property
    def n_classes(self) -> int:
        return self.layers[-1].b

    @property
    def img_elems(self) -> int:
        return self.h * self.w * self.c

    def to_blob(self) -> bytes:
        out = [b"SIBM", struct.pack("<7I", 1, self.h, self.w, self.c, len(self.layers), len(self.weights), 0)]
        for L in self.layers:
            out.append(struct.pack("<8I", L.kind, L.a, L.b, L.c, L.d, L.flags, L.w_off, L.b_off))
        out.append(np.ascontiguousarray(self.weights, "<f4").tobytes())
        return b"".join(out)

    def flops_per_image(self) -> int:
        h, w, f = self.h, self.w, 0
        for L in self.layers:
            if L.kind == KIND_CONV:
                oh, ow = h - L.a + 1, w - L.b + 1
                f += 2 * oh * ow * L.d * L.a * L.b * L.c
                h, w = oh // 2, ow // 2
            else:
                f += 2 * L.a * L.b
        return f

    def bytes_per_image(self) -> int:
        return self.img_elems * 4 + self.n_classes * 4


def layout(h: int, w: int, c: int, convs, dense) -> tuple[list, int]:
    layers, off = [], 0
    for cout in convs:
        layers.append(Layer(KIND_CONV, 5, 5, c, cout, 1, off, off + 25 * c * cout))
        off += 25 * c * cout + cout
        h, w, c = (h - 4) // 2, (w - 4) // 2, cout
    feat = h * w * c
    for i, width in enumerate(dense):
        relu = 1 if i + 1 < len(dense) else 0
        layers.append(Layer(KIND_DENSE, feat, width, 0, 0, relu, off, off + feat * width))
        off += feat * width + width
        feat = width
    return layers, off


def make_preset(name: str) -> Model:
    """Synthetic weights per D2: U(-a,a), a = sqrt(6/fan_in) (x4 on the last layer); biases U(-0.1,0.1).
    Seed 0xC0FFEE00 + preset id; the RNG index of a weight is its position in the weight section."""
    pid, h, w, c, convs, dense = PRESETS[name]
    layers, nw = layout(h, w, c, convs, dense)
    u = u01(0xC0FFEE00 + pid, 0, nw)
    wt = np.empty(nw, np.float32)
    for i, L in enumerate(layers):
        fan_in = L.a * L.b * L.c if L.kind == KIND_CONV else L.a
        n_w = fan_in * L.d if L.kind == KIND_CONV else L.a * L.b
        n_b = L.d if L.kind == KIND_CONV else L.b
        a = np.float32(np.sqrt(6.0 / fan_in))
        if i + 1 == len(layers):
            a = np.float32(a * np.float32(4.0))
        ws = slice(L.w_off, L.w_off + n_w)
        bs = slice(L.b_off, L.b_off + n_b)
        wt[ws] = a * (np.float32(2.0) * u[ws] - np.float32(1.0))
        wt[bs] = np.float32(0.1) * (np.float32(2.0) * u[bs] - np.float32(1.0))
    return Model(name, h, w, c, layers, wt)


def custom_model(name: str, h: int, w: int, c: int, convs, dense, seed: int = 7) -> Model:
    """Same family, arbitrary sizes (edge-case tests of the generic fp32 path)."""
    layers, nw = layout(h, w, c, convs, dense)
    u = u01(seed, 0, nw)
    wt = np.empty(nw, np.float32)
    for i, L in enumerate(layers):
        fan_in = L.a * L.b * L.c if L.kind == KIND_CONV else L.a
        n_w = fan_in * L.d if L.kind == KIND_CONV else L.a * L.b
        n_b = L.d if L.kind == KIND_CONV else L.b
        a = np.float32(np.sqrt(6.0 / fan_in))
        wt[L.w_off:L.w_off + n_w] = a * (np.float32(2.0) * u[L.w_off:L.w_off + n_w] - np.float32(1.0))
        wt[L.b_off:L.b_off + n_b] = np.float32(0.1) * (np.float32(2.0) * u[L.b_off:L.b_off + n_b] - np.float32(1.0))
    return Model(name, h, w, c, layers, wt)


def synthetic_images(model: Model, batch_index: int, n: int, kind: str = "uniform") -> np.ndarray:
    """[n,H,W,C] fp32 synthetic images, seed 1234 + batch_index (D2).
    kind "uniform": pixels ~ U[0,1).  kind "patches": the same noise modulated by a per-image 8x8-cell random
    brightness map, which spreads the logits over more classes (used by parity tests as a second distribution)."""
    x = u01(1234 + batch_index, 0, n * model.img_elems).reshape(n, model.h, model.w, model.c)
    if kind == "uniform":
        return x
    assert kind == "patches"
    cell = 8
    gh, gw = (model.h + cell - 1) // cell, (model.w + cell - 1) // cell
    g = u01(0x9A7C4E5 + batch_index, 0, n * gh * gw * model.c).reshape(n, gh, gw, model.c)
    g = np.repeat(np.repeat(g, cell, 1), cell, 2)[:, :model.h, :model.w, :]
    return (g * (np.float32(0.5) + np.float32(0.5) * x)).astype(np.float32)


def _f32_text(v) -> str:
    # shortest decimal that round-trips the fp32 value (what a producer serialising fp32 pixels would send)
    s = np.format_float_positional(np.float32(v), unique=True, trim="0")
    return s


def instances_json(x: np.ndarray) -> bytes:
    """{"instances": [N][H][W][C]} for a batch of images (the KafkaSpout payload, README.md:22-34 of the reference)."""
    x = np.asarray(x, np.float32)
    if x.ndim == 3:
        x = x[None]
    n, h, w, c = x.shape
    flat = [_f32_text(v) for v in x.reshape(-1)]
    it = iter(flat)
    parts = []
    for _ in range(n):
        rows = []
        for _ in range(h):
            px = []
            for _ in range(w):
                px.append("[" + ",".join(next(it) for _ in range(c)) + "]")
            rows.append("[" + ",".join(px) + "]")
        parts.append("[" + ",".join(rows) + "]")
    return ('{"instances":[' + ",".join(parts) + "]}").encode()
