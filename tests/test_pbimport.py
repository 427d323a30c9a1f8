"""SavedModel / GraphDef importer (csrc/sib_pbimport.cpp; replaces SavedModelBundle.load,
InferenceBolt.java:49-58). CPU-only: builds protobuf bytes with tests/pb_writer.py (an independent encoder) and checks
that the imported blob is byte-identical to the blob modelspec writes for the same weights."""
import numpy as np
import pytest

import pb_writer as pw
from sib_b200 import modelspec, native

PRESETS = ["mnist-lenet5", "cifar-lenet5", "mnist-wide", "cifar-wide"]


@pytest.mark.parametrize("preset", PRESETS)
@pytest.mark.parametrize("container", ["saved_model", "graph_def"])
def test_import_equals_blob(preset, container):
    m = modelspec.make_preset(preset)
    g = pw.lenet_graph(m)
    pb = pw.saved_model(g) if container == "saved_model" else g
    assert native.import_pb(pb) == m.to_blob()


@pytest.mark.parametrize("kw", [
    dict(identity=False),
    dict(encoding="packed"),
    dict(encoding="repeated"),
    dict(relu_first=False),            # pool(relu(x)) == relu(pool(x)): both orders are the same layer
    dict(add_op="Add"),
    dict(add_op="AddV2"),
    dict(packed_lists=False),
    dict(placeholder_shape=False),      # geometry inferred from the first dense layer (square input)
    dict(softmax_name="probabilities"),  # only one Softmax in the graph
])
def test_import_variants(kw):
    m = modelspec.make_preset("mnist-lenet5")
    assert native.import_pb(pw.saved_model(pw.lenet_graph(m, **kw))) == m.to_blob()


def test_custom_shape_roundtrip():
    m = modelspec.custom_model("odd", 20, 20, 2, convs=[4, 8], dense=[24, 7], seed=11)
    blob = native.import_pb(pw.saved_model(pw.lenet_graph(m)))
    assert blob == m.to_blob()
    info = native.describe(blob)
    assert (info.in_h, info.in_w, info.in_c, info.n_classes) == (20, 20, 2, 7)


def test_describe_and_codec_take_the_pb_directly():
    """Every entry point that takes a model takes saved_model.pb bytes as they are."""
    m = modelspec.make_preset("cifar-lenet5")
    pb = pw.saved_model(pw.lenet_graph(m))
    a, b = native.describe(pb), native.describe(m.to_blob())
    for f in ("in_h", "in_w", "in_c", "n_layers", "n_classes", "n_weights", "flops_per_image", "bytes_per_image"):
        assert getattr(a, f) == getattr(b, f)


def test_transpose_b_is_undone():
    # the MatMul kernel is stored [out,in] with transpose_b = true; the blob wants TF's usual [in,out]
    nodes = [
        pw.node("input", "Placeholder", dtype=pw.attr_type(pw.DT_FLOAT), shape=pw.attr_shape([-1, 4, 4, 1])),
    ]
    # a tiny 1-conv net is enough: conv 3x3x1x2 on 4x4 -> 2x2x2 -> pool 1x1x2 -> dense 2 -> 3
    rng = np.random.default_rng(0)
    k = rng.standard_normal((3, 3, 1, 2)).astype(np.float32)
    kb = rng.standard_normal(2).astype(np.float32)
    d = rng.standard_normal((2, 3)).astype(np.float32)
    db = rng.standard_normal(3).astype(np.float32)
    f = pw.attr_type(pw.DT_FLOAT)
    nodes += [
        pw.node("k", "Const", dtype=f, value=pw.attr_tensor(k)), pw.node("kb", "Const", dtype=f, value=pw.attr_tensor(kb)),
        pw.node("d", "Const", dtype=f, value=pw.attr_tensor(np.ascontiguousarray(d.T))), pw.node("db", "Const", dtype=f, value=pw.attr_tensor(db)),
        pw.node("c", "Conv2D", ["input", "k"], strides=pw.attr_ints([1, 1, 1, 1]), padding=pw.attr_s("VALID")),
        pw.node("cb", "BiasAdd", ["c", "kb"]), pw.node("cr", "Relu", ["cb"]),
        pw.node("cp", "MaxPool", ["cr"], ksize=pw.attr_ints([1, 2, 2, 1]), strides=pw.attr_ints([1, 2, 2, 1]), padding=pw.attr_s("VALID")),
        pw.node("fl", "Reshape", ["cp", "shape"]),
        pw.node("mm", "MatMul", ["fl", "d"], transpose_b=pw.attr_b(True)),
        pw.node("out", "BiasAdd", ["mm", "db"]), pw.node("output/Softmax", "Softmax", ["out"]),
    ]
    blob = native.import_pb(pw.graph(nodes))
    want = modelspec.Model("t", 4, 4, 1, [modelspec.Layer(1, 3, 3, 1, 2, 1, 0, 18), modelspec.Layer(2, 2, 3, 0, 0, 0, 20, 26)],
                           np.concatenate([k.ravel(), kb, d.ravel(), db]).astype(np.float32))
    assert blob == want.to_blob()


@pytest.mark.parametrize("kw,needle", [
    (dict(frozen=False), "not a Const"),
    (dict(conv_padding="SAME"), "VALID"),
])
def test_refusals(kw, needle):
    m = modelspec.make_preset("mnist-lenet5")
    with pytest.raises(native.SibError) as e:
        native.import_pb(pw.saved_model(pw.lenet_graph(m, **kw)))
    assert e.value.code == native.EINVAL and needle in str(e.value)


def test_refuses_garbage_and_truncation():
    m = modelspec.make_preset("mnist-lenet5")
    pb = pw.saved_model(pw.lenet_graph(m))
    for bad in (b"", b"\x00" * 64, b"not a protobuf at all, just text....", pb[: len(pb) // 2], pb[:-3]):
        with pytest.raises(native.SibError):
            native.import_pb(bad)
    with pytest.raises(native.SibError):  # no Softmax
        native.import_pb(pw.graph([pw.node("input", "Placeholder")]))


def test_mutated_bytes_never_crash():
    """Bit flips, truncations and splices of a valid file: the reader returns a blob or an error, nothing else."""
    m = modelspec.make_preset("mnist-lenet5")
    pb = bytearray(pw.saved_model(pw.lenet_graph(m, encoding="repeated")))  # the encoding with most tags per byte
    rng = np.random.default_rng(5)
    head = len(pb) // 50  # node structure lives at both ends of every message; weights fill the middle
    outcomes = {"ok": 0, "err": 0}
    for trial in range(400):
        b = bytearray(pb)
        for _ in range(int(rng.integers(1, 6))):
            pos = int(rng.integers(0, head)) if trial % 2 else int(rng.integers(0, len(b)))
            b[pos] = int(rng.integers(0, 256))
        if trial % 7 == 0:
            b = b[: int(rng.integers(1, len(b)))]
        try:
            native.import_pb(bytes(b))
            outcomes["ok"] += 1
        except native.SibError:
            outcomes["err"] += 1
    assert outcomes["err"] > 0 and outcomes["ok"] + outcomes["err"] == 400


def test_tf113_shaped_fixture_imports_to_the_preset_blob():
    """tests/golden/lenet_tf113_saved_model.pb is assembled field by field from TF 1.13's public .proto files by
    tests/golden/make_tf113_fixture.py -- a second encoder that shares nothing with pb_writer.py -- in the shape a real
    frozen tf.layers LeNet export has: meta_info_def, saver_def, signature_def, Const -> Identity reads, the
    Shape/StridedSlice/Pack/Reshape flatten, MatMul + Add, an unknown batch dimension, a control input."""
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    pb = open(os.path.join(here, "golden", "lenet_tf113_saved_model.pb"), "rb").read()
    m = modelspec.make_preset("mnist-lenet5")
    assert native.import_pb(pb) == m.to_blob()
    info = native.describe(pb)
    assert (info.in_h, info.in_w, info.in_c, info.n_classes) == (28, 28, 1, 10)
    # the committed bytes are what the committed generator writes
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_tf113_fixture", os.path.join(here, "golden", "make_tf113_fixture.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    assert gen.build_saved_model(m) == pb


def test_cyclic_graphs_are_refused_not_followed():
    """A GraphDef can feed a node its own output; the walk must stop with an error, not loop or grow without bound."""
    f = pw.attr_type(pw.DT_FLOAT)
    w = np.zeros((4, 3), np.float32)
    b = np.zeros(3, np.float32)
    base = [pw.node("input", "Placeholder", dtype=f, shape=pw.attr_shape([-1, 4, 4, 1])),
            pw.node("w", "Const", dtype=f, value=pw.attr_tensor(w)), pw.node("b", "Const", dtype=f, value=pw.attr_tensor(b))]
    dense_cycle = base + [pw.node("mm", "MatMul", ["add", "w"]), pw.node("add", "Add", ["mm", "b"]),   # add -> mm -> add ...
                          pw.node("output/Softmax", "Softmax", ["add"])]
    reshape_cycle = base + [pw.node("fl", "Reshape", ["fl", "shape"]), pw.node("mm", "MatMul", ["fl", "w"]),
                            pw.node("add", "Add", ["mm", "b"]), pw.node("output/Softmax", "Softmax", ["add"])]
    relu_cycle = base + [pw.node("r", "Relu", ["p"]), pw.node("p", "MaxPool", ["r"], ksize=pw.attr_ints([1, 2, 2, 1]),
                                                               strides=pw.attr_ints([1, 2, 2, 1]), padding=pw.attr_s("VALID")),
                         pw.node("fl", "Reshape", ["r", "shape"]), pw.node("mm", "MatMul", ["fl", "w"]),
                         pw.node("add", "Add", ["mm", "b"]), pw.node("output/Softmax", "Softmax", ["add"])]
    for nodes in (dense_cycle, reshape_cycle, relu_cycle):
        with pytest.raises(native.SibError) as e:
            native.import_pb(pw.graph(nodes))
        assert e.value.code == native.EINVAL


@pytest.mark.gpu
def test_engine_runs_from_saved_model_pb():
    """prepare() can hand the extracted saved_model.pb straight to sib_engine_open."""
    m = modelspec.make_preset("mnist-lenet5")
    x = modelspec.synthetic_images(m, 0, 64)
    with native.Engine(pw.saved_model(pw.lenet_graph(m)), precision=native.PREC_FP32) as e_pb, native.Engine(m.to_blob(), precision=native.PREC_FP32) as e_blob:
        p1 = e_pb.forward_sync(x)[0]
        p2 = e_blob.forward_sync(x)[0]
    assert np.array_equal(p1, p2)
