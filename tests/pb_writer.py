"""Minimal protobuf *encoder* for TensorFlow GraphDef / SavedModel messages -- test infrastructure only.

Independent of the C++ reader in csrc/sib_pbimport.cpp (that one only decodes). Field numbers are the public ones of
tensorflow/core/framework/{graph,node_def,attr_value,tensor,tensor_shape}.proto and
tensorflow/core/protobuf/{saved_model,meta_graph}.proto. No TensorFlow exists on this box, so these bytes are what a
TF 1.13 `simple_save` of a frozen LeNet would contain *as far as the schema goes*; parity with a real TF export stays
unpinned.
"""
from __future__ import annotations

import struct

import numpy as np

DT_FLOAT, DT_INT32 = 1, 3


def varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def f_varint(field: int, v: int) -> bytes:
    return varint(field << 3 | 0) + varint(v)


def f_bytes(field: int, b: bytes) -> bytes:
    return varint(field << 3 | 2) + varint(len(b)) + b


def f_f32(field: int, x: float) -> bytes:
    return varint(field << 3 | 5) + struct.pack("<f", x)


def shape(dims) -> bytes:  # TensorShapeProto
    return b"".join(f_bytes(2, f_varint(1, d)) for d in dims)


def tensor(a: np.ndarray, encoding: str = "content") -> bytes:  # TensorProto
    a = np.asarray(a)
    dt = DT_FLOAT if a.dtype == np.float32 else DT_INT32
    out = f_varint(1, dt) + f_bytes(2, shape(a.shape))
    if encoding == "content":
        out += f_bytes(4, a.astype("<f4" if dt == DT_FLOAT else "<i4").tobytes())
    elif encoding == "packed":      # float_val, packed
        out += f_bytes(5, a.astype("<f4").tobytes())
    elif encoding == "repeated":    # float_val, one tag per element (legal for a repeated scalar)
        out += b"".join(f_f32(5, float(x)) for x in a.ravel())
    else:
        raise ValueError(encoding)
    return out


def attr_s(s: str) -> bytes:
    return f_bytes(2, s.encode())


def attr_b(v: bool) -> bytes:
    return f_varint(5, int(v))


def attr_type(t: int) -> bytes:
    return f_varint(6, t)


def attr_shape(dims) -> bytes:
    return f_bytes(7, shape(dims))


def attr_tensor(a, encoding="content") -> bytes:
    return f_bytes(8, tensor(a, encoding))


def attr_ints(vals, packed=True) -> bytes:  # AttrValue.list.i
    body = f_bytes(3, b"".join(varint(v) for v in vals)) if packed else b"".join(f_varint(3, v) for v in vals)
    return f_bytes(1, body)


def node(name: str, op: str, inputs=(), **attrs) -> bytes:  # NodeDef
    out = f_bytes(1, name.encode()) + f_bytes(2, op.encode())
    for i in inputs:
        out += f_bytes(3, i.encode())
    for k, v in attrs.items():
        out += f_bytes(5, f_bytes(1, k.encode()) + f_bytes(2, v))
    return out


def graph(nodes) -> bytes:  # GraphDef: node = 1, versions = 4
    return b"".join(f_bytes(1, n) for n in nodes) + f_bytes(4, f_varint(1, 27))


def saved_model(graph_def: bytes) -> bytes:  # SavedModel{1: schema version, 2: MetaGraphDef{1: meta info, 2: graph_def}}
    meta_info = f_bytes(4, b"serve") + f_bytes(5, b"1.13.1")  # tags, tensorflow_version
    return f_varint(1, 1) + f_bytes(2, f_bytes(1, meta_info) + f_bytes(2, graph_def))


def lenet_graph(model, *, identity=True, encoding="content", placeholder_shape=True, relu_first=True, add_op="BiasAdd",
                packed_lists=True, conv_padding="VALID", frozen=True, softmax_name="output/Softmax") -> bytes:
    """GraphDef of `model` (a modelspec.Model) the way TF 1.x lays a frozen LeNet out."""
    W = model.weights
    nodes = []
    ph_attrs = {"dtype": attr_type(DT_FLOAT)}
    if placeholder_shape:
        ph_attrs["shape"] = attr_shape([-1, model.h, model.w, model.c])
    nodes.append(node("input", "Placeholder", **ph_attrs))
    cur = "input"

    def const(name, arr):
        if not frozen and name.endswith("kernel") and "conv1" in name:
            nodes.append(node(name, "VariableV2", dtype=attr_type(DT_FLOAT), shape=attr_shape(arr.shape)))
        else:
            nodes.append(node(name, "Const", dtype=attr_type(DT_FLOAT), value=attr_tensor(arr, encoding)))
        if identity:
            nodes.append(node(name + "/read", "Identity", [name], T=attr_type(DT_FLOAT)))
            return name + "/read"
        return name

    n_dense = 0
    for li, L in enumerate(model.layers):
        if L.kind == 1:
            nm = f"conv{li + 1}"
            k = const(nm + "/kernel", W[L.w_off:L.w_off + L.a * L.b * L.c * L.d].reshape(L.a, L.b, L.c, L.d))
            b = const(nm + "/bias", W[L.b_off:L.b_off + L.d])
            nodes.append(node(nm + "/Conv2D", "Conv2D", [cur, k], T=attr_type(DT_FLOAT), strides=attr_ints([1, 1, 1, 1], packed_lists),
                              padding=attr_s(conv_padding), data_format=attr_s("NHWC"), dilations=attr_ints([1, 1, 1, 1], packed_lists),
                              use_cudnn_on_gpu=attr_b(True)))
            nodes.append(node(nm + "/BiasAdd", add_op, [nm + "/Conv2D", b], T=attr_type(DT_FLOAT)))
            pool_attrs = dict(T=attr_type(DT_FLOAT), ksize=attr_ints([1, 2, 2, 1], packed_lists), strides=attr_ints([1, 2, 2, 1], packed_lists),
                              padding=attr_s("VALID"), data_format=attr_s("NHWC"))
            if relu_first:
                nodes.append(node(nm + "/Relu", "Relu", [nm + "/BiasAdd"], T=attr_type(DT_FLOAT)))
                nodes.append(node(nm + "/MaxPool", "MaxPool", [nm + "/Relu:0"], **pool_attrs))
                cur = nm + "/MaxPool"
            else:
                nodes.append(node(nm + "/MaxPool", "MaxPool", [nm + "/BiasAdd"], **pool_attrs))
                nodes.append(node(nm + "/Relu", "Relu", [nm + "/MaxPool"], T=attr_type(DT_FLOAT)))
                cur = nm + "/Relu"
        else:
            if n_dense == 0:
                nodes.append(node("flatten/shape", "Const", dtype=attr_type(DT_INT32), value=attr_tensor(np.array([-1, L.a], np.int32))))
                nodes.append(node("flatten/Reshape", "Reshape", [cur, "flatten/shape"], T=attr_type(DT_FLOAT), Tshape=attr_type(DT_INT32)))
                cur = "flatten/Reshape"
            n_dense += 1
            last = li == len(model.layers) - 1
            nm = "output" if last else f"fc{n_dense}"
            k = const(nm + "/kernel", W[L.w_off:L.w_off + L.a * L.b].reshape(L.a, L.b))
            b = const(nm + "/bias", W[L.b_off:L.b_off + L.b])
            nodes.append(node(nm + "/MatMul", "MatMul", [cur, k], T=attr_type(DT_FLOAT), transpose_a=attr_b(False), transpose_b=attr_b(False)))
            nodes.append(node(nm + "/BiasAdd", add_op, [nm + "/MatMul", b, "^input"] if last else [nm + "/MatMul", b], T=attr_type(DT_FLOAT)))
            cur = nm + "/BiasAdd"
            if L.flags & 1:
                nodes.append(node(nm + "/Relu", "Relu", [cur], T=attr_type(DT_FLOAT)))
                cur = nm + "/Relu"
    nodes.append(node(softmax_name, "Softmax", [cur], T=attr_type(DT_FLOAT)))
    nodes.append(node("init", "NoOp"))
    return graph(nodes)
