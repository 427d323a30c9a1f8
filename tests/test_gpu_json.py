"""GPU-side decode of {"instances": ...} tuples (sib_config.json_on_gpu; csrc/sib_kernels_json.cu) against the host
decoder, which is itself checked bit-for-bit against the oracle in tests/test_codec.py.  The bar is bit-exactness:
text -> double -> float on the device must give the very floats Jackson's rule gives, so the forward pass sees the
same input and returns the same bits (InferenceBolt.java:76-80)."""
import json
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _payload(vals, h, w, c, ws="", sep_ws=""):
    it = iter(vals)
    rows = []
    for _ in range(h):
        px = []
        for _ in range(w):
            px.append("[" + ("," + sep_ws).join(next(it) for _ in range(c)) + "]")
        rows.append("[" + ("," + sep_ws).join(px) + "]")
    return ("%s{%s\"instances\"%s:%s[[%s]]%s}" % (ws, ws, ws, ws, ("," + sep_ws).join(rows), ws)).encode()


def _run(e, tuples, timeout_s=20.0):
    for i, t in enumerate(tuples):
        while True:
            rc = e.submit_json(i, t)
            if rc != -4:
                break
            time.sleep(0.0005)
        assert rc == 1, (i, rc)
    e.flush()
    got, t0 = [], time.time()
    while len(got) < len(tuples) and time.time() - t0 < timeout_s:
        got += e.poll(timeout_us=100000)
    assert [g[0] for g in got] == list(range(len(tuples)))
    return got


def test_gpu_decode_is_bit_exact_with_host_decode(sib, presets):
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    rng = np.random.default_rng(5)
    n_el = m.img_elems
    tuples, want_x = [], []
    styles = [
        lambda v: ms._f32_text(v),                                   # shortest fp32 text (the producer's format)
        lambda v: repr(float(np.float32(v))),                        # double repr of the float (up to 17 digits: host redo)
        lambda v: "%.6e" % float(v),                                 # exponent form with sign and padding
        lambda v: ("%.4f" % float(v)),
        lambda v: str(int(v * 255)),                                 # integers
        lambda v: "%.3E" % float(v * 1e-3 - 5e-4),                   # negatives, E notation
    ]
    for k in range(48):
        v = rng.random(n_el).astype(np.float32)
        txt = [styles[k % len(styles)](x) for x in v]
        ws, sep = [("", ""), (" ", " "), ("\n\t", "\n ")][k % 3]
        tuples.append(_payload(txt, m.h, m.w, m.c, ws, sep))
        want_x.append(np.array([np.float32(float(s)) for s in txt], np.float32).reshape(1, m.h, m.w, m.c))
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=32, deadline_us=10**6, json_on_gpu=True, strict_batching=True) as e:
        got = _run(e, tuples)
        st = e.stats()
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64) as h:
        ref, _, ref_am = h.forward_sync(np.concatenate(want_x))
    for i, (tag, p, a, status) in enumerate(got):
        assert status == 0 and p.shape == (1, 10)
        assert p.tobytes() == ref[i:i + 1].tobytes(), f"tuple {i}: decoded input differs from Jackson's text->double->float"
        assert a[0] == ref_am[i]
    assert st.tuples == 48 and st.images == 48
    # the kernel itself took the tuples it is meant for; only the 17-digit style (1 in 6) went to the host decoder
    assert st.json_gpu_decoded >= 38 and st.json_gpu_decoded + st.json_host_redone == 48


def test_gpu_decode_hands_the_unusual_to_the_host_decoder(sib, oracle, presets):
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 81, 4)
    om = oracle.Model(m.to_blob())
    _, pr = om.forward(x)
    good = [ms.instances_json(x[i:i + 1]) for i in range(4)]
    two = ms.instances_json(x[:2])                                            # N = 2 rows in one tuple
    dup = good[0][:-1] + b',"instances":' + good[1][len(b'{"instances":'):]    # duplicate key: the last one wins
    nul = good[2].replace(b"[[[[", b"[[[[null,", 1)                           # 785 values: ragged -> ESHAPE
    trail = good[3] + b"   trailing bytes are ignored"
    lead0 = good[0].replace(b"[[[[0.", b"[[[[00.", 1) if b"[[[[0." in good[0] else None
    trunc = good[1][:-7]
    unk = good[1].replace(b'"instances"', b'"inputs"')
    small = b'{"instances":[[[[1.0,2.0]]]]}'                                  # not 28x28x1; shorter than a real tuple
    tuples = [good[0], two, dup, nul, trail, trunc, unk, small, good[3]]
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=16, deadline_us=10**6, json_on_gpu=True, strict_batching=True) as e:
        got = _run(e, tuples)
    EJ, ES = sib.native.EJSON, sib.native.ESHAPE
    assert got[0][3] == 0 and np.abs(got[0][1] - pr[0:1]).max() <= 1e-5
    assert got[1][3] == 0 and got[1][1].shape == (2, 10) and np.abs(got[1][1] - pr[:2]).max() <= 1e-5
    assert got[2][3] == 0 and np.abs(got[2][1] - pr[1:2]).max() <= 1e-5      # second "instances" value
    assert got[3][3] == ES and got[3][1] is None
    assert got[4][3] == 0 and np.abs(got[4][1] - pr[3:4]).max() <= 1e-5
    assert got[5][3] == EJ and got[6][3] == EJ
    assert got[7][3] == ES
    assert got[8][3] == 0 and np.abs(got[8][1] - pr[3:4]).max() <= 1e-5


def test_bolt_contract_holds_with_gpu_decode(sib, oracle, presets):
    from importlib import import_module
    B = import_module("sib_b200.bolt")
    m = presets["cifar-lenet5"]
    ms = sib.modelspec
    om = oracle.Model(m.to_blob())
    bolt, col = B.InferenceBolt(m.to_blob()), B.OutputCollector()
    bolt.prepare({"sib.precision": "fp32", "sib.max.batch": 16, "sib.deadline.us": 300, "sib.json.gpu": 1}, 0, col)
    x = ms.synthetic_images(m, 82, 40)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(40)]
    payloads[7] = payloads[7][:100]
    payloads[33] = b'{"instances": 5}'
    for i, p in enumerate(payloads):
        bolt.execute(B.Tuple(i, p))
    bolt.cleanup()
    ev = col.events
    assert [e[1] for e in ev if e[0] == "ack"] == list(range(40))
    emits = [e[1] for e in ev if e[0] == "emit"]
    assert len(emits) == 40 and emits[7] is None and emits[33] is None
    for i in (0, 8, 39):
        got = np.array(json.loads(emits[i])["predictions"], np.float32)
        ref = np.array(json.loads(om.execute(payloads[i]))["predictions"], np.float32)
        assert np.abs(got - ref).max() <= 1e-5
    bolt.close()


def test_gpu_decode_full_rate(sib, presets):
    """A few thousand CIFAR tuples through the load generator with device-side decode: all come back, tensor path."""
    m = presets["cifar-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 83, 16)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(16)]
    with sib.native.Engine(m.to_blob(), max_batch=1024, deadline_us=300, json_on_gpu=True) as e:
        r = sib.native.loadgen(e, payloads, 50000.0, 20000, mode=sib.native.LG_JSON, producers=2)
        assert r["completed"] == 20000 and r["rejected"] == 0
        st = e.stats()
        assert st.images == 20000 and st.json_gpu_decoded == 20000 and st.json_host_redone == 0


def test_auto_decode_uses_both_decoders_and_gives_the_same_bits(sib, presets):
    """json_on_gpu = 2: a tuple that finds the engine idle is decoded on the submitting thread, one that finds it
    busy is shipped as text; predictions do not depend on which (both decoders are bit-exact with each other)."""
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 84, 300)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(300)]
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64, deadline_us=500) as h:
        want, _, _ = h.forward_sync(x)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64, deadline_us=500, json_on_gpu=2) as e:
        got = []
        for i, p in enumerate(payloads):                  # a burst: the first is alone, the rest arrive on a busy engine
            while e.submit_json(i, p) == sib.native.EAGAIN:
                got += e.poll(timeout_us=1000)
        e.flush()
        while len(got) < 300:
            got += e.poll(timeout_us=100000)
        st = e.stats()
    assert [g[0] for g in got] == list(range(300))
    assert np.concatenate([g[1] for g in got]).tobytes() == want.tobytes()
    assert 0 < st.json_gpu_decoded < 300                  # some went each way


def test_gpu_decode_rejects_strays_between_numbers(sib, presets):
    """A run that holds no token start but ends in a digit ("7,e5,8", "7,.5,8", "7,+5,8", "7,E3,8", "7 e5]") and a
    leading zero are malformed for Jackson and for the host decoder; the device decoder must not let them through
    either: every one of them comes back an error (the bolt emits a null message), never a prediction."""
    m = presets["cifar-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 85, 2)
    good = ms.instances_json(x[:1])
    assert good.startswith(b'{"instances":[[[[') and b"," in good
    bad = [good.replace(b",", b"," + stray + b",", 1) for stray in (b"e5", b".5", b"+5", b"E3")]   # [a,e5,b,c]
    bad.append(good.replace(b"]", b" e5]", 1))                                                       # [a,b,c e5]
    bad.append(good.replace(b"[[[[", b"[[[[0", 1))                                                   # leading zero
    assert all(t != good for t in bad)
    tuples = [good] + bad + [ms.instances_json(x[1:2])]
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=16, deadline_us=10**6, json_on_gpu=True, strict_batching=True) as e:
        got = _run(e, tuples)
    assert got[0][3] == 0 and got[-1][3] == 0
    for g in got[1:-1]:
        assert g[3] == sib.native.EJSON and g[1] is None, g[3]


def test_long_tuple_with_tiny_text_area_goes_to_the_host_decoder(sib, presets):
    """max_batch = 1 gives a slot a text area smaller than one 17-digit MNIST tuple (~16 KB): the tuple must take the
    host decoder instead of being copied past the end of the pinned text buffer."""
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 86, 3)
    payloads = [json.dumps({"instances": x[i:i + 1].astype(np.float64).tolist()}).encode() for i in range(3)]  # 17 digits
    assert min(len(p) for p in payloads) > 12000
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=1, deadline_us=200, json_on_gpu=True) as e:
        got = _run(e, payloads)
        st = e.stats()
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=8) as h:
        ref, _, _ = h.forward_sync(x)
    assert [g[3] for g in got] == [0, 0, 0]
    assert np.concatenate([g[1] for g in got]).tobytes() == ref.tobytes()
    assert st.json_gpu_decoded == 0
