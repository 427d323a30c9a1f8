"""The C-ABI library loads on a box without a GPU and exports every symbol include/*.h declares.
No compute call is made here (those are the -m gpu tests)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sib_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(sib):
    lib = ctypes.CDLL(sib.native.LIB_PATH)
    names = _declared("sib.h") + _declared("sib_bolt.h") + _declared("sib_loadgen.h")
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ but not exported by libsib.so"


def test_binding_table_covers_header(sib):
    from importlib import import_module
    bolt = import_module("sib_b200.bolt")
    assert sorted(sib.native.SYMBOLS) == _declared("sib.h")
    assert sorted(bolt.HARNESS_SYMBOLS) == _declared("sib_bolt.h")
    assert sorted(sib.native.LOADGEN_SYMBOLS) == [n for n in _declared("sib_loadgen.h") if n.startswith("sib_loadgen")]


def test_struct_sizes_match_header(sib):
    n = sib.native
    assert ctypes.sizeof(n.Config) == 40
    assert ctypes.sizeof(n.Result) == 40
    assert ctypes.sizeof(n.ModelInfo) == 40
    assert ctypes.sizeof(n.Stats) == 8 * (7 + 16 + 5)


def test_model_describe_and_errors_without_gpu(sib, presets):
    n = sib.native
    info = n.describe(presets["cifar-lenet5"].to_blob())
    assert (info.in_h, info.in_w, info.in_c, info.n_classes) == (32, 32, 3, 10)
    assert info.flops_per_image == 1303440 and info.bytes_per_image == 12328
    info = n.describe(presets["mnist-lenet5"].to_blob())
    assert info.flops_per_image == 563280 and info.bytes_per_image == 3176
    assert n.describe(presets["cifar-wide"].to_blob()).flops_per_image == 4545280
    bad = bytearray(presets["mnist-lenet5"].to_blob())
    bad[0] = ord("X")
    try:
        n.describe(bytes(bad))
        raise AssertionError("bad magic accepted")
    except n.SibError as e:
        assert e.code == n.EINVAL and "magic" in str(e)
    assert n.lib.sib_strerror(n.ECUDA).decode().startswith("CUDA failure")
    assert b"sm_100a" in n.lib.sib_version()


def test_no_cpu_fallback(sib, presets):
    """Without a GPU the engine must refuse to open (there is no CPU path in the product)."""
    n = sib.native
    if n.device_count() > 0:
        return
    try:
        n.Engine(presets["mnist-lenet5"].to_blob(), precision=n.PREC_FP32)
        raise AssertionError("engine opened without a GPU")
    except n.SibError as e:
        assert e.code == n.ECUDA
