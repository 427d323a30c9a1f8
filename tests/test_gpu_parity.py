"""Parity proper: the CUDA path through the C ABI (sib_forward_sync) against the CPU oracle on the same seeded
inputs and weights.  Bars: fp32 path logits/probabilities max-abs <= 1e-5 (x logit scale), argmax
exact unless the fp64 shadow oracle says the top two are a genuine tie."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _argmax_agrees(got_argmax, lg32, lg64, tie=1e-5):
    want = lg32.argmax(1)
    bad = np.nonzero(got_argmax != want)[0]
    for i in bad:
        s = np.sort(lg64[i])
        assert s[-1] - s[-2] < tie, f"image {i}: argmax {got_argmax[i]} vs oracle {want[i]}, fp64 margin {s[-1]-s[-2]}"
    return len(bad)


@pytest.mark.parametrize("name", ["mnist-lenet5", "cifar-lenet5", "mnist-wide", "cifar-wide"])
@pytest.mark.parametrize("kind", ["uniform", "patches"])
def test_fp32_path_matches_oracle(sib, oracle, presets, name, kind):
    m = presets[name]
    n = 259  # not a multiple of the kernel's images-per-CTA
    x = sib.modelspec.synthetic_images(m, 11, n, kind=kind)
    om = oracle.Model(m.to_blob())
    lg, pr = om.forward(x, threads=0)
    lg64, _ = om.forward64(x[:64])
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=128) as e:  # 259 > 128: chunked
        gp, gl, ga = e.forward_sync(x)
    scale = max(1.0, float(np.abs(lg).max()))
    assert np.abs(gl - lg).max() <= 1e-5 * scale
    assert np.abs(gp - pr).max() <= 1e-5
    np.testing.assert_allclose(gp.sum(1), 1.0, atol=1e-5)
    assert (ga[:64] == gl[:64].argmax(1)).all()
    _argmax_agrees(ga[:64], lg[:64], lg64)
    assert (ga == lg.argmax(1)).mean() > 0.999


@pytest.mark.parametrize("name", ["mnist-lenet5", "cifar-lenet5", "mnist-wide", "cifar-wide"])
def test_fp32_path_matches_golden_fixtures(sib, presets, name):
    m = presets[name]
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    x = sib.modelspec.synthetic_images(m, 0, 16)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32) as e:
        gp, gl, ga = e.forward_sync(x)
    assert np.abs(gl - g["logits64"]).max() <= 1e-5 * max(1.0, np.abs(g["logits64"]).max())
    assert np.abs(gp - g["probs64"]).max() <= 1e-5
    assert (ga == g["logits64"].argmax(1)).all()


@pytest.mark.parametrize("shape", [
    (14, 14, 1, (3, 5), (7, 10)),     # smallest map that survives two conv5+pool stages... 14->10->5->1->x: invalid
    (16, 16, 2, (2, 3), (4,)),         # 16->12->6->2->1
    (17, 21, 3, (5, 7), (33, 10)),     # odd conv outputs: pool drops the last row/col (13->6, 17->8)
    (40, 24, 4, (9, 12), (50, 20, 10)),
])
def test_fp32_generic_shapes(sib, oracle, shape):
    h, w, c, convs, dense = shape
    ms = sib.modelspec
    if min(h, w) < 16:
        m = ms.custom_model("bad", 16, 16, 1, (2, 2), (3,))  # placeholder so the param list stays honest
        blob = bytearray(m.to_blob())
        blob[8] = 14  # in_h: 14 -> conv2 no longer fits; the engine must reject the blob, not crash
        with pytest.raises(sib.native.SibError):
            sib.native.Engine(bytes(blob), precision=sib.native.PREC_FP32)
        return
    m = ms.custom_model("custom", h, w, c, convs, dense, seed=h * 100 + w)
    x = ms.u01(77, 0, 9 * m.img_elems).reshape(9, h, w, c)
    lg, pr = oracle.Model(m.to_blob()).forward(x)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32) as e:
        gp, gl, ga = e.forward_sync(x)
    assert np.abs(gl - lg).max() <= 1e-5 * max(1.0, float(np.abs(lg).max()))
    assert np.abs(gp - pr).max() <= 1e-5
    assert (ga == lg.argmax(1)).all()


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 4097])
def test_batch_edges(sib, oracle, presets, n):
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 2, n)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=4096) as e:
        gp, gl, ga = e.forward_sync(x)
    assert gp.shape == (n, 10)
    if n:
        lg, pr = oracle.Model(m.to_blob()).forward(x, threads=0)
        assert np.abs(gp - pr).max() <= 1e-5


def test_full_size_properties(sib, presets):
    """BASELINE-size batch (4096 CIFAR images): size-independent properties instead of a slow CPU diff."""
    m = presets["cifar-lenet5"]
    n = 4096
    x = sib.modelspec.synthetic_images(m, 5, n)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=n) as e:
        p1, l1, a1 = e.forward_sync(x)
        p2, l2, a2 = e.forward_sync(x)
        assert p1.tobytes() == p2.tobytes() and l1.tobytes() == l2.tobytes()  # deterministic
        perm = np.random.default_rng(0).permutation(n)
        p3, l3, a3 = e.forward_sync(x[perm])
        assert l3.tobytes() == l1[perm].tobytes()                                # images are independent
        np.testing.assert_allclose(p1.sum(1), 1.0, atol=1e-5)
        assert (a1 == l1.argmax(1)).all()
        assert np.isfinite(l1).all()
