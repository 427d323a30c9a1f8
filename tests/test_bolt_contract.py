"""Bolt contract, read the way a test of the reference's InferenceBolt would read (InferenceBolt.java:43-105):
prepare -> execute per tuple -> exactly one emit of field "message" and one ack per tuple, in arrival order;
malformed JSON emits a null message and still acks; a wrong-shaped payload throws IllegalArgumentException."""
import json

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bolt(sib, blob, **conf):
    from importlib import import_module
    B = import_module("sib_b200.bolt")
    bolt = B.InferenceBolt(blob)
    col = B.OutputCollector()
    c = {"sib.precision": "fp32", "sib.max.batch": 32, "sib.deadline.us": 300}
    c.update({k.replace("_", "."): v for k, v in conf.items()})
    bolt.prepare(c, 0, col)
    return B, bolt, col


def test_declares_message_field(sib, presets):
    B, bolt, col = _bolt(sib, presets["mnist-lenet5"].to_blob())
    assert bolt.declareOutputFields() == ["message"]          # InferenceBolt.java:104; KafkaBolt reads "message"
    bolt.close()


def test_exactly_once_in_order_with_nulls(sib, oracle, presets):
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    om = oracle.Model(m.to_blob())
    B, bolt, col = _bolt(sib, m.to_blob())
    x = ms.synthetic_images(m, 31, 100)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(100)]
    payloads[5] = b'{"instances": [[[[0.5'                      # truncated JSON
    payloads[50] = b'{"inputs": [[[[0.5]]]]}'                   # unknown property
    payloads[99] = b'not json'
    for i, p in enumerate(payloads):
        bolt.execute(B.Tuple(i, p))
    bolt.cleanup()
    ev = col.events
    assert len(ev) == 200
    emits = [e for e in ev if e[0] == "emit"]
    acks = [e[1] for e in ev if e[0] == "ack"]
    assert acks == list(range(100))                             # one ack per tuple, arrival order
    for k in range(100):                                        # emit precedes its ack, pairwise
        assert ev[2 * k][0] == "emit" and ev[2 * k + 1] == ("ack", k)
    for i, (_, msg) in enumerate(emits):
        if i in (5, 50, 99):
            assert msg is None                                  # outputJson stays null (InferenceBolt.java:72,92-98)
            continue
        want = om.execute(payloads[i])                          # oracle's mirror of execute(): JSON in -> JSON out
        got = np.array(json.loads(msg)["predictions"], np.float32)
        ref = np.array(json.loads(want)["predictions"], np.float32)
        assert got.shape == (1, 10)
        assert np.abs(got - ref).max() <= 1e-5
        assert set(json.loads(msg)) == {"predictions"}
    bolt.close()


def test_wrong_shape_throws_like_tensorflow(sib, presets):
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    B, bolt, col = _bolt(sib, m.to_blob())
    good = ms.instances_json(ms.synthetic_images(m, 32, 1))
    bolt.execute(B.Tuple(0, good))
    with pytest.raises(B.IllegalArgumentException):
        bolt.execute(B.Tuple(1, b'{"instances": [[[[1.0, 2.0]]]]}'))      # 1x1x2 is not 28x28x1
    with pytest.raises(B.IllegalArgumentException):
        bolt.execute(B.Tuple(2, b'{"instances": [[[[1.0],[2.0, 3.0]]]]}'))  # ragged
    with pytest.raises(B.IllegalArgumentException):
        bolt.execute(B.Tuple(3, b'{}'))                                     # instances == null -> NPE in the reference
    bolt.execute(B.Tuple(4, good))
    bolt.cleanup()
    acks = [e[1] for e in col.events if e[0] == "ack"]
    assert acks == [0, 4]                                       # thrown tuples are neither emitted nor acked
    bolt.close()


def test_deadline_emits_without_more_input(sib, presets):
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    B, bolt, col = _bolt(sib, m.to_blob(), sib_max_batch=4096, sib_deadline_us=1000)
    for i in range(3):
        bolt.execute(B.Tuple(i, ms.instances_json(ms.synthetic_images(m, 40 + i, 1))))
    ev = bolt.sync_events(wait_for=6, timeout_us=5_000_000)    # no cleanup(), no further tuples: deadline must fire
    assert [e[1] for e in ev if e[0] == "ack"] == [0, 1, 2]
    bolt.close()


def test_multi_row_tuple_is_a_documented_superset(sib, presets):
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    B, bolt, col = _bolt(sib, m.to_blob())
    bolt.execute(B.Tuple(0, ms.instances_json(ms.synthetic_images(m, 50, 3))))  # reference: copyTo(new float[1][10]) throws
    bolt.cleanup()
    msg = [e[1] for e in col.events if e[0] == "emit"][0]
    assert np.array(json.loads(msg)["predictions"]).shape == (3, 10)
    bolt.close()


def test_two_bolts_sharing_one_engine_keep_the_contract(sib, oracle, presets):
    """sib.shared = 1: two executors of one worker process, one engine underneath. Each bolt still emits and acks
    exactly its own tuples, once, in its own arrival order, nulls included (InferenceBolt.java:92-99)."""
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    om = oracle.Model(m.to_blob())
    bolts = [_bolt(sib, m.to_blob(), sib_shared=1) for _ in range(2)]
    x = ms.synthetic_images(m, 33, 80)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(80)]
    payloads[7] = b'{"instances": [[[[0.5'
    for i, p in enumerate(payloads):                            # tuples alternate between the two executors
        B, bolt, col = bolts[i % 2]
        bolt.execute(B.Tuple(i, p))
    for B, bolt, col in bolts:
        bolt.cleanup()
    for k, (B, bolt, col) in enumerate(bolts):
        mine = list(range(k, 80, 2))
        assert [e[1] for e in col.events if e[0] == "ack"] == mine
        emits = [e[1] for e in col.events if e[0] == "emit"]
        assert len(emits) == len(mine)
        for i, msg in zip(mine, emits):
            if i == 7:
                assert msg is None
                continue
            got = np.array(json.loads(msg)["predictions"], np.float32)
            ref = np.array(json.loads(om.execute(payloads[i]))["predictions"], np.float32)
            assert np.abs(got - ref).max() <= 1e-5
        bolt.close()
