"""The micro-batcher through the C ABI: tags and order, flush-by-size, flush-by-deadline, JSON submit, back-pressure.
Mirrors what InferenceBolt.execute() promises per tuple (InferenceBolt.java:69-100)."""
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _drain(e, want, timeout_s=10.0):
    out = []
    t0 = time.time()
    while len(out) < want and time.time() - t0 < timeout_s:
        out += e.poll(timeout_us=100000)
    return out


def test_submit_poll_order_and_values(sib, oracle, presets):
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 21, 300)
    _, pr = oracle.Model(m.to_blob()).forward(x, threads=0)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64, deadline_us=1000, n_slots=3, strict_batching=True) as e:
        got = []
        for i in range(300):
            while True:
                rc = e.submit_f32(1000 + i, x[i])
                if rc == sib.native.EAGAIN:
                    got += e.poll(timeout_us=1000)
                    continue
                assert rc == 1
                break
        e.flush()
        got += _drain(e, 300 - len(got))
        assert [t for t, *_ in got] == list(range(1000, 1300))       # submission order, exactly once
        for i, (_, p, a, status) in enumerate(got):
            assert status == 0 and p.shape == (1, 10)
            assert np.abs(p[0] - pr[i]).max() <= 1e-5
            assert a[0] == pr[i].argmax()
        s = e.stats()
        assert s.tuples == 300 and s.images == 300
        assert s.flush_by_size == 4 and s.batches >= 5                 # 4 x 64 + remainder
        assert s.h2d_bytes == 300 * m.img_elems * 4


def test_flush_by_deadline(sib, presets):
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 22, 3)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=4096, deadline_us=2000, strict_batching=True) as e:
        t0 = time.time()
        for i in range(3):
            assert e.submit_f32(i, x[i]) == 1
        got = _drain(e, 3, timeout_s=5.0)                                # no flush() call: the deadline thread launches
        dt = time.time() - t0
        assert [t for t, *_ in got] == [0, 1, 2]
        assert dt < 2.0
        s = e.stats()
        assert s.flush_by_deadline == 1 and s.flush_by_size == 0 and s.flush_explicit == 0
        assert s.batch_hist[1] == 1                                      # one batch of 3 images: bucket [2,4)


def test_multi_image_tuples_and_json_submit(sib, oracle, presets):
    m = presets["cifar-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 23, 5)
    _, pr = oracle.Model(m.to_blob()).forward(x)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=8, deadline_us=500, strict_batching=True) as e:
        assert e.submit_json(7, ms.instances_json(x[:2])) == 2           # N=2 rows in one tuple (InstObj.java:8)
        assert e.submit_f32(8, x[2:5]) == 3
        assert e.submit_json(9, b'{"instances": [[[[1.0]]]') == sib.native.EJSON
        assert e.submit_json(9, b'{"instances": [[[[1.0]]]]}') == sib.native.ESHAPE   # 1x1x1 is not 32x32x3
        assert e.submit_json(9, ms.instances_json(np.zeros((9, 32, 32, 3), np.float32))) == sib.native.ECAP
        e.flush()
        got = _drain(e, 2)
        assert [(t, p.shape) for t, p, *_ in got] == [(7, (2, 10)), (8, (3, 10))]
        assert np.abs(got[0][1] - pr[:2]).max() <= 1e-5
        assert np.abs(got[1][1] - pr[2:]).max() <= 1e-5
        # library-side formatting equals the oracle's PredObj text for the same probabilities
        assert e.submit_json(10, ms.instances_json(x[:1])) == 1
        e.flush()
        msgs = []
        t0 = time.time()
        while not msgs and time.time() - t0 < 5:
            msgs = e.poll_json(timeout_us=100000)
        tag, text = msgs[0]
        assert tag == 10 and text.startswith(b'{"predictions":[[') and text.endswith(b"]]}")
        vals = np.array(eval(text.decode().replace("E", "e"))["predictions"], np.float32)
        assert np.abs(vals - pr[:1]).max() <= 1e-5


def test_back_pressure_is_eagain_not_deadlock(sib, presets):
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 24, 1)
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=2, deadline_us=10**6, n_slots=2, strict_batching=True) as e:
        rcs = [e.submit_f32(i, x[0]) for i in range(6)]
        assert rcs[:4] == [1, 1, 1, 1] and sib.native.EAGAIN in rcs[4:]   # 2 slots x 2 images, nobody polled
        got = _drain(e, 4)
        assert [t for t, *_ in got] == [0, 1, 2, 3]
        e.poll()                                                           # hands the slots back
        assert e.submit_f32(99, x[0]) == 1


def test_close_with_pending_work(sib, presets):
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 25, 10)
    e = sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64, deadline_us=10**6, strict_batching=True)
    for i in range(10):
        e.submit_f32(i, x[i])
    e.close()  # must flush, wait and free without hanging


def test_submit_pinned_is_zero_copy_and_mixes_with_staged(sib, oracle, presets):
    torch = pytest.importorskip("torch")
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 26, 96)
    _, pr = oracle.Model(m.to_blob()).forward(x, threads=0)
    pinned = torch.empty((96, 28, 28, 1), dtype=torch.float32, pin_memory=True)
    pinned.numpy()[:] = x
    with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64, deadline_us=10**6, strict_batching=True) as e:
        # tuples alternate between caller-pinned memory (DMA straight from it) and staged copies, inside one batch
        for i in range(0, 96, 8):
            if (i // 8) % 2 == 0:
                assert e.submit_pinned(i, pinned.data_ptr() + i * 784 * 4, 8) == 8
            else:
                assert e.submit_f32(i, x[i:i + 8]) == 8
        e.flush()
        got = _drain(e, 12)
        assert [t for t, *_ in got] == list(range(0, 96, 8))
        P = np.concatenate([p for _, p, _, _ in got])
        assert np.abs(P - pr).max() <= 1e-5
        st = e.stats()
        assert st.h2d_bytes == 96 * 784 * 4 and st.flush_by_size == 1


def test_loadgen_latency_is_bounded_by_deadline_plus_service(sib, presets):
    """Open-loop arrivals far below capacity: every tuple comes back, and p99 stays near the flush deadline."""
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    x = ms.synthetic_images(m, 27, 8)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(8)]
    with sib.native.Engine(m.to_blob(), max_batch=256, deadline_us=300, strict_batching=True) as e:
        sib.native.loadgen(e, payloads, 2000.0, 500, mode=sib.native.LG_JSON)          # warm-up
        r = sib.native.loadgen(e, payloads, 5000.0, 5000, mode=sib.native.LG_JSON, producers=2, want_latencies=True)
        assert r["completed"] == 5000 and r["rejected"] == 0 and r["images"] == 5000
        lat = r["latencies_ns"]
        assert (lat > 0).all()
        assert r["p50_ns"] < 5_000_000 and r["p99_ns"] < 50_000_000                     # ms-scale, generous for a shared box
        st = e.stats()
        assert st.flush_by_deadline > 0                                                 # low rate: batches close on the deadline
    with sib.native.Engine(m.to_blob(), max_batch=64, deadline_us=300) as e:
        r = sib.native.loadgen(e, [x[i] for i in range(8)], 200000.0, 20000, mode=sib.native.LG_F32, producers=2)
        assert r["completed"] == 20000


def test_engines_on_several_threads_with_different_models(sib, presets):
    """Executors of one worker process each own an engine (InferenceBolt.java:57-58 per executor); models and
    precisions differ, so kernels with different shared-memory needs are configured and launched concurrently."""
    import threading

    combos = [("mnist-lenet5", sib.native.PREC_FP32), ("cifar-wide", sib.native.PREC_FP32),
              ("cifar-lenet5", sib.native.PREC_FP16), ("mnist-wide", sib.native.PREC_FP32),
              ("cifar-lenet5", sib.native.PREC_FP32), ("mnist-lenet5", sib.native.PREC_FP16)]
    want = {}
    for name, prec in combos:  # single-threaded answers first
        m = presets[name]
        x = sib.modelspec.synthetic_images(m, 70, 300)
        with sib.native.Engine(m.to_blob(), precision=prec, max_batch=512) as e:
            want[(name, prec)] = e.forward_sync(x)[0].copy()
    errors = []

    def run(name, prec):
        try:
            m = presets[name]
            x = sib.modelspec.synthetic_images(m, 70, 300)
            for _ in range(3):
                with sib.native.Engine(m.to_blob(), precision=prec, max_batch=512) as e:
                    for _ in range(5):
                        got = e.forward_sync(x)[0]
                        if got.tobytes() != want[(name, prec)].tobytes():
                            errors.append((name, prec, float(np.abs(got - want[(name, prec)]).max())))
        except Exception as ex:  # noqa: BLE001
            errors.append((name, prec, repr(ex)))

    ts = [threading.Thread(target=run, args=c) for c in combos]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors


def test_eager_launches_when_idle_and_still_batches_under_load(sib, oracle, presets):
    """Default policy (sib_config.strict_batching = 0): a tuple that finds nothing queued or in flight does not wait for the deadline; a burst that
    arrives while that batch runs is still grouped; order and values are unchanged."""
    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 71, 400)
    _, pr = oracle.Model(m.to_blob()).forward(x, threads=0)
    with sib.native.Engine(m.to_blob(), max_batch=256, deadline_us=2_000_000) as e:  # default: strict_batching = 0
        t0 = time.time()
        assert e.submit_f32(7, x[0:1]) == 1
        out = _drain(e, 1, timeout_s=1.0)
        assert len(out) == 1 and time.time() - t0 < 0.5          # the 2 s deadline was not waited for
        assert out[0][0] == 7 and np.abs(out[0][1] - pr[0:1]).max() <= 1e-2
        for i in range(1, 400):                                    # a burst: the first launches alone, the rest group up
            while e.submit_f32(100 + i, x[i:i + 1]) == sib.native.EAGAIN:
                _drain(e, 1, timeout_s=1.0)
        e.flush()
        got = out + _drain(e, 399)
        st = e.stats()
    assert [g[0] for g in got[1:]] == [100 + i for i in range(1, 400)]
    assert np.abs(np.concatenate([g[1] for g in got[1:]]) - pr[1:]).max() <= 1e-2
    assert st.flush_by_deadline == 0 and st.batches < 200          # grouped, not one launch per tuple


def test_co_located_bolts_share_one_engine(sib, oracle, presets):
    """sib_client_*: three bolts of one worker process open the same model: one engine, common batches, and every
    bolt gets back exactly its own tuples, in its own order, with the right values (threading row)."""
    import threading

    m = presets["mnist-lenet5"]
    x = sib.modelspec.synthetic_images(m, 72, 600)
    _, pr = oracle.Model(m.to_blob()).forward(x, threads=0)
    img = m.h * m.w * m.c
    kw = dict(precision=sib.native.PREC_FP32, max_batch=128, deadline_us=300)
    clients = [sib.native.Client(m.to_blob(), img, task_index=t * sib.native.device_count(), **kw) for t in range(3)]
    other = sib.native.Client(m.to_blob(), img, precision=sib.native.PREC_FP32, max_batch=64, deadline_us=300)  # another cfg: its own engine
    got = [[] for _ in clients]
    errors = []

    def run(ci):
        try:
            c = clients[ci]
            mine = list(range(ci, 600, 3))
            sent = 0
            while len(got[ci]) < len(mine):
                while sent < len(mine):
                    rc = c.submit_f32(1000 * ci + sent, x[mine[sent]:mine[sent] + 1])
                    if rc == sib.native.EAGAIN:
                        break
                    assert rc == 1, rc
                    sent += 1
                got[ci] += c.poll(timeout_us=2000)
        except Exception as ex:  # noqa: BLE001
            errors.append(repr(ex))

    ts = [threading.Thread(target=run, args=(i,)) for i in range(3)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors
    st = clients[0].stats()
    assert st.tuples == 600 and st.batches < 600          # one engine saw all three bolts' tuples and batched them
    assert other.stats().tuples == 0
    for ci in range(3):
        mine = list(range(ci, 600, 3))
        assert [g[0] for g in got[ci]] == [1000 * ci + k for k in range(len(mine))]
        assert np.abs(np.concatenate([g[1] for g in got[ci]]) - pr[mine]).max() <= 1e-5
    for c in clients + [other]:
        c.close()


def test_poll_json_batch_formats_a_whole_poll_like_the_per_tuple_path(sib, oracle, presets):
    """sib_poll_json_batch (batched egress towards KafkaBolt, KafkaBolt.java:116-166): the messages of one poll, written
    back to back by one native pass, are byte for byte what sib_poll + sib_format_json give tuple by tuple, and within
    fp32 noise of the oracle's writeValueAsString; order, a multi-row tuple and a small arena (tuples that do not fit stay
    queued for the next call) included."""
    import json
    import time
    m = presets["mnist-lenet5"]
    ms = sib.modelspec
    om = oracle.Model(m.to_blob())
    x = ms.synthetic_images(m, 91, 40)
    payloads = [ms.instances_json(x[i:i + 1]) for i in range(40)]
    payloads[5] = ms.instances_json(x[5:8])          # three rows in one tuple

    def run(batched):
        with sib.native.Engine(m.to_blob(), precision=sib.native.PREC_FP32, max_batch=64, deadline_us=300, strict_batching=True) as e:
            for i, p in enumerate(payloads):
                assert e.submit_json(i, p) > 0
            e.flush()
            got, t0 = [], time.time()
            while len(got) < 40 and time.time() - t0 < 20:
                if batched:
                    got += [(t, msg) for t, msg, st in e.poll_json_batch(max_results=7, timeout_us=100000, arena_bytes=1500)]
                else:
                    got += e.poll_json(timeout_us=100000)
        return got

    per_tuple, batch = run(False), run(True)
    assert [g[0] for g in batch] == list(range(40))
    assert batch == per_tuple
    for i, (tag, msg) in enumerate(batch):
        got = np.array(json.loads(msg)["predictions"], np.float32)
        ref = np.array(json.loads(om.execute(payloads[i]))["predictions"], np.float32)
        assert got.shape == ((3, 10) if i == 5 else (1, 10)) and np.abs(got - ref).max() <= 1e-5
