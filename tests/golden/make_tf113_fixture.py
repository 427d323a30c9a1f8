#!/usr/bin/env python
"""Writes tests/golden/lenet_tf113_saved_model.pb: a saved_model.pb in the shape TensorFlow 1.13 gives a frozen
tf.layers LeNet-5 (what the reference keeps at model/saved_model.pb, InferenceBolt.java:49-58).

No TensorFlow exists on this box, so the file is assembled by hand, field by field, from the public .proto files of
TF 1.13 (field numbers quoted beside every write). It deliberately does NOT use tests/pb_writer.py, the encoder the
other importer tests share: this script has its own ten-line wire writer, writes the messages in the order and with
the extras a real export carries, and is the second, independent witness the importer is checked against.

What a real TF-1.13 export has that tests/pb_writer.py never emits, all present here:
  * SavedModel.saved_model_schema_version, MetaGraphDef.meta_info_def (stripped_op_list, tags, versions),
    saver_def, signature_def["serving_default"] with TensorInfo for input:0 / output/Softmax:0;
  * GraphDef.versions and an (empty) function library; nodes carry T / dtype / _class / _output_shapes attrs;
  * frozen variables as Const -> Identity ("<var>/read") pairs, Const tensors with dtype and tensor_shape BEFORE
    tensor_content;
  * tf.layers.flatten's shape subgraph: Shape -> StridedSlice (three int32 Const stacks) -> Pack with a -1 Const ->
    Reshape;
  * tf.layers.dense as MatMul + BiasAdd and a hand-written last layer as MatMul + Add;
  * a Placeholder whose batch dimension is unknown (-1), a NoOp and a control input ("^group_deps").
The weights are modelspec's mnist-lenet5 preset, so the imported blob must equal modelspec's blob byte for byte.

Run from the repo root:  python tests/golden/make_tf113_fixture.py
"""
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import sib_b200 as sib  # noqa: E402

# ---- protobuf wire format (developers.google.com/protocol-buffers/docs/encoding) --------------------------------------


def varint(v: int) -> bytes:
    v &= (1 << 64) - 1  # negative int64 -> ten bytes, two's complement
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def f_varint(field: int, v: int) -> bytes:  # wire type 0
    return varint(field << 3 | 0) + varint(v)


def f_bytes(field: int, payload: bytes) -> bytes:  # wire type 2: strings, bytes, embedded messages, packed lists
    return varint(field << 3 | 2) + varint(len(payload)) + payload


def f_str(field: int, s: str) -> bytes:
    return f_bytes(field, s.encode())


# ---- tensorflow/core/framework/types.proto -----------------------------------------------------------------------------
DT_FLOAT, DT_INT32 = 1, 3

# ---- tensorflow/core/framework/tensor_shape.proto ----------------------------------------------------------------------


def tensor_shape(dims) -> bytes:
    # TensorShapeProto { repeated Dim dim = 2; }   Dim { int64 size = 1; }
    return b"".join(f_bytes(2, f_varint(1, d)) for d in dims)


# ---- tensorflow/core/framework/tensor.proto ----------------------------------------------------------------------------


def tensor_proto(dtype: int, dims, content: bytes) -> bytes:
    # TensorProto { DataType dtype = 1; TensorShapeProto tensor_shape = 2; bytes tensor_content = 4; }
    return f_varint(1, dtype) + f_bytes(2, tensor_shape(dims)) + f_bytes(4, content)


def tensor_proto_int_scalar(v: int) -> bytes:
    # a scalar int32 Const as TF writes it: empty shape, the value in int_val (field 7, packed)
    return f_varint(1, DT_INT32) + f_bytes(2, b"") + f_bytes(7, varint(v))


# ---- tensorflow/core/framework/attr_value.proto ------------------------------------------------------------------------
# AttrValue { oneof: bytes s = 2; int64 i = 3; float f = 4; bool b = 5; DataType type = 6; TensorShapeProto shape = 7;
#             TensorProto tensor = 8; ListValue list = 1; }
# ListValue { repeated bytes s = 2; repeated int64 i = 3 [packed]; repeated TensorShapeProto shape = 7; }


def a_type(t): return f_varint(6, t)
def a_bool(b): return f_varint(5, 1 if b else 0)
def a_int(i): return f_varint(3, i)
def a_str(s): return f_bytes(2, s.encode())
def a_shape(dims): return f_bytes(7, tensor_shape(dims))
def a_tensor(tp): return f_bytes(8, tp)
def a_ints(v): return f_bytes(1, f_bytes(3, b"".join(varint(x) for x in v)))  # packed, as proto3 writes them
def a_strs(v): return f_bytes(1, b"".join(f_bytes(2, s.encode()) for s in v))
def a_shapes(v): return f_bytes(1, b"".join(f_bytes(7, tensor_shape(d)) for d in v))


# ---- tensorflow/core/framework/node_def.proto --------------------------------------------------------------------------


def node(name: str, op: str, inputs=(), **attrs) -> bytes:
    # NodeDef { string name = 1; string op = 2; repeated string input = 3; string device = 4;
    #           map<string, AttrValue> attr = 5; }      map entry { key = 1; value = 2; }
    out = f_str(1, name) + f_str(2, op) + b"".join(f_str(3, i) for i in inputs)
    for k in sorted(attrs):  # TF serialises the attr map in key order
        out += f_bytes(5, f_str(1, k) + f_bytes(2, attrs[k]))
    return out


def frozen_variable(name: str, arr: np.ndarray):
    """convert_variables_to_constants: VariableV2 -> Const, its '/read' Identity stays."""
    dims = list(arr.shape)
    c = node(name, "Const", dtype=a_type(DT_FLOAT), value=a_tensor(tensor_proto(DT_FLOAT, dims, arr.astype("<f4").tobytes())),
             _output_shapes=a_shapes([dims]))
    r = node(name + "/read", "Identity", [name], T=a_type(DT_FLOAT), _class=a_strs(["loc:@" + name]), _output_shapes=a_shapes([dims]))
    return [c, r]


def int_const(name: str, values) -> bytes:
    content = b"".join(struct.pack("<i", v) for v in values)
    return node(name, "Const", dtype=a_type(DT_INT32), value=a_tensor(tensor_proto(DT_INT32, [len(values)], content)))


def build_graph(m) -> bytes:
    W, L = m.weights, m.layers
    nodes = []
    # Placeholder: dtype, shape with an unknown batch dimension
    nodes.append(node("input", "Placeholder", dtype=a_type(DT_FLOAT), shape=a_shape([-1, m.h, m.w, m.c]),
                      _output_shapes=a_shapes([[-1, m.h, m.w, m.c]])))
    cur = "input"
    conv_names = ["conv2d", "conv2d_1"]
    pool_names = ["max_pooling2d/MaxPool", "max_pooling2d_1/MaxPool"]
    for li in range(2):  # tf.layers.conv2d(activation=relu) + tf.layers.max_pooling2d
        lay, nm = L[li], conv_names[li]
        k = W[lay.w_off:lay.w_off + lay.a * lay.b * lay.c * lay.d].reshape(lay.a, lay.b, lay.c, lay.d)  # HWIO
        b = W[lay.b_off:lay.b_off + lay.d]
        nodes += frozen_variable(nm + "/kernel", k) + frozen_variable(nm + "/bias", b)
        nodes.append(node(nm + "/Conv2D", "Conv2D", [cur, nm + "/kernel/read"], T=a_type(DT_FLOAT), data_format=a_str("NHWC"),
                          dilations=a_ints([1, 1, 1, 1]), padding=a_str("VALID"), strides=a_ints([1, 1, 1, 1]),
                          use_cudnn_on_gpu=a_bool(True)))
        nodes.append(node(nm + "/BiasAdd", "BiasAdd", [nm + "/Conv2D", nm + "/bias/read"], T=a_type(DT_FLOAT), data_format=a_str("NHWC")))
        nodes.append(node(nm + "/Relu", "Relu", [nm + "/BiasAdd"], T=a_type(DT_FLOAT)))
        nodes.append(node(pool_names[li], "MaxPool", [nm + "/Relu"], T=a_type(DT_FLOAT), data_format=a_str("NHWC"),
                          ksize=a_ints([1, 2, 2, 1]), padding=a_str("VALID"), strides=a_ints([1, 2, 2, 1])))
        cur = pool_names[li]
    # tf.layers.flatten: reshape(x, [shape(x)[0], -1])
    nodes.append(node("flatten/Shape", "Shape", [cur], T=a_type(DT_FLOAT), out_type=a_type(DT_INT32)))
    nodes += [int_const("flatten/strided_slice/stack", [0]), int_const("flatten/strided_slice/stack_1", [1]),
              int_const("flatten/strided_slice/stack_2", [1])]
    nodes.append(node("flatten/strided_slice", "StridedSlice",
                      ["flatten/Shape", "flatten/strided_slice/stack", "flatten/strided_slice/stack_1", "flatten/strided_slice/stack_2"],
                      Index=a_type(DT_INT32), T=a_type(DT_INT32), begin_mask=a_int(0), ellipsis_mask=a_int(0), end_mask=a_int(0),
                      new_axis_mask=a_int(0), shrink_axis_mask=a_int(1)))
    nodes.append(node("flatten/Reshape/shape/1", "Const", dtype=a_type(DT_INT32), value=a_tensor(tensor_proto_int_scalar(-1))))
    nodes.append(node("flatten/Reshape/shape", "Pack", ["flatten/strided_slice", "flatten/Reshape/shape/1"], N=a_int(2),
                      T=a_type(DT_INT32), axis=a_int(0)))
    nodes.append(node("flatten/Reshape", "Reshape", [cur, "flatten/Reshape/shape"], T=a_type(DT_FLOAT), Tshape=a_type(DT_INT32)))
    cur = "flatten/Reshape"
    for li, nm in ((2, "dense"), (3, "dense_1")):  # tf.layers.dense(activation=relu): MatMul + BiasAdd + Relu
        lay = L[li]
        nodes += frozen_variable(nm + "/kernel", W[lay.w_off:lay.w_off + lay.a * lay.b].reshape(lay.a, lay.b))
        nodes += frozen_variable(nm + "/bias", W[lay.b_off:lay.b_off + lay.b])
        nodes.append(node(nm + "/MatMul", "MatMul", [cur, nm + "/kernel/read"], T=a_type(DT_FLOAT), transpose_a=a_bool(False),
                          transpose_b=a_bool(False)))
        nodes.append(node(nm + "/BiasAdd", "BiasAdd", [nm + "/MatMul", nm + "/bias/read"], T=a_type(DT_FLOAT), data_format=a_str("NHWC")))
        nodes.append(node(nm + "/Relu", "Relu", [nm + "/BiasAdd"], T=a_type(DT_FLOAT)))
        cur = nm + "/Relu"
    # the last layer written by hand inside tf.name_scope("output"): tf.matmul(h, w) + b, then tf.nn.softmax
    lay = L[4]
    nodes += frozen_variable("output/kernel", W[lay.w_off:lay.w_off + lay.a * lay.b].reshape(lay.a, lay.b))
    nodes += frozen_variable("output/bias", W[lay.b_off:lay.b_off + lay.b])
    nodes.append(node("group_deps", "NoOp", ["^input"]))  # a control-only node; nothing data-flows through it
    nodes.append(node("output/MatMul", "MatMul", [cur, "output/kernel/read"], T=a_type(DT_FLOAT), transpose_a=a_bool(False),
                      transpose_b=a_bool(False)))
    nodes.append(node("output/add", "Add", ["output/MatMul", "output/bias/read", "^group_deps"], T=a_type(DT_FLOAT)))  # control input last
    nodes.append(node("output/Softmax", "Softmax", ["output/add"], T=a_type(DT_FLOAT), _output_shapes=a_shapes([[-1, lay.b]])))
    # tensorflow/core/framework/graph.proto: GraphDef { repeated NodeDef node = 1; FunctionDefLibrary library = 2;
    #                                                    VersionDef versions = 4; }   VersionDef { int32 producer = 1; }
    return b"".join(f_bytes(1, n) for n in nodes) + f_bytes(4, f_varint(1, 27)) + f_bytes(2, b"")


def tensor_info(name: str, dims) -> bytes:
    # tensorflow/core/protobuf/meta_graph.proto: TensorInfo { string name = 1; DataType dtype = 2; TensorShapeProto tensor_shape = 3; }
    return f_str(1, name) + f_varint(2, DT_FLOAT) + f_bytes(3, tensor_shape(dims))


def build_saved_model(m) -> bytes:
    graph = build_graph(m)
    ops = ["Add", "BiasAdd", "Const", "Conv2D", "Identity", "MatMul", "MaxPool", "NoOp", "Pack", "Placeholder", "Relu", "Reshape",
           "Shape", "Softmax", "StridedSlice"]
    # MetaInfoDef { string meta_graph_version = 1; OpList stripped_op_list = 2; repeated string tags = 4;
    #               string tensorflow_version = 5; string tensorflow_git_version = 6; }    OpList { repeated OpDef op = 1; } OpDef { string name = 1; }
    meta_info = (f_bytes(2, b"".join(f_bytes(1, f_str(1, o)) for o in ops)) + f_str(4, "serve") + f_str(5, "1.13.1") +
                 f_str(6, "b'v1.13.1-0-g6612da8951'"))
    # SaverDef { string filename_tensor_name = 1; string save_tensor_name = 2; string restore_op_name = 3; int32 max_to_keep = 4;
    #            bool sharded = 5; float keep_checkpoint_every_n_hours = 6; CheckpointFormatVersion version = 7; }
    saver = f_str(1, "save/Const:0") + f_str(2, "save/Identity:0") + f_str(3, "save/restore_all") + f_varint(4, 5) + f_varint(5, 1) + f_varint(7, 2)
    # SignatureDef { map<string, TensorInfo> inputs = 1; map<string, TensorInfo> outputs = 2; string method_name = 3; }
    sig = (f_bytes(1, f_str(1, "input") + f_bytes(2, tensor_info("input:0", [-1, m.h, m.w, m.c]))) +
           f_bytes(2, f_str(1, "output") + f_bytes(2, tensor_info("output/Softmax:0", [-1, m.n_classes]))) +
           f_str(3, "tensorflow/serving/predict"))
    # MetaGraphDef { MetaInfoDef meta_info_def = 1; GraphDef graph_def = 2; SaverDef saver_def = 3;
    #                map<string, SignatureDef> signature_def = 5; }
    meta_graph = f_bytes(1, meta_info) + f_bytes(2, graph) + f_bytes(3, saver) + f_bytes(5, f_str(1, "serving_default") + f_bytes(2, sig))
    # tensorflow/core/protobuf/saved_model.proto: SavedModel { int64 saved_model_schema_version = 1; repeated MetaGraphDef meta_graphs = 2; }
    return f_varint(1, 1) + f_bytes(2, meta_graph)


if __name__ == "__main__":
    model = sib.modelspec.make_preset("mnist-lenet5")
    pb = build_saved_model(model)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lenet_tf113_saved_model.pb")
    with open(out, "wb") as f:
        f.write(pb)
    print(out, len(pb), "bytes")
