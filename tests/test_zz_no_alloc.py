"""sib_forward_device_static: the forward entry point whose margin-guard scratch is a fixed array inside the library
instead of engine-owned cudaMalloc memory, for callers that may not allocate device memory (a GPU already fully
handed out to other executors, a latency-critical loop that cannot take cudaMalloc's device-wide synchronisation, a
no-allocation phase).  It must give what sib_forward_device gives, call after call, although every call reuses the
same array and finds the previous call's contents in it.  Collected last on purpose: a run that stops at the first
failure has then already run every other test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(e, entry, x, torch):
    n = x.shape[0]
    d_in = torch.from_numpy(x).cuda()
    d_probs = torch.full((n, 10), -1.0, dtype=torch.float32, device="cuda")
    d_logits = torch.full((n, 10), -1.0, dtype=torch.float32, device="cuda")
    d_argmax = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = entry(d_in.data_ptr(), n, d_probs.data_ptr(), d_logits.data_ptr(), d_argmax.data_ptr(), st)
    assert rc in (None, 0), rc
    torch.cuda.synchronize()
    return d_probs.cpu().numpy(), d_logits.cpu().numpy(), d_argmax.cpu().numpy()


def test_static_scratch_entry_matches_engine_scratch_entry_twice_in_a_row(sib, oracle, presets):
    torch = pytest.importorskip("torch")
    m = presets["cifar-lenet5"]
    ms = sib.modelspec
    with sib.native.Engine(m.to_blob(), device=0, precision=sib.native.PREC_FP16, max_batch=8192, guard=True) as e:
        # two calls in a row on different batches and sizes: the second finds the first one's margins, list and
        # counter in the shared array and must not be influenced by them
        batches = [ms.synthetic_images(m, 71, 4096), ms.synthetic_images(m, 72, 1000, kind="patches")]
        want = [_run(e, e.forward_device, x, torch) for x in batches]
        got = [_run(e, e.forward_device_static, x, torch) for x in batches]
        for (wp, wl, wa), (gp, gl, ga), x in zip(want, got, batches):
            assert (ga == wa).all()                               # same predictions, image by image
            assert np.abs(gl - wl).max() <= 1e-6                  # same kernels, same order of operations
            assert np.abs(gp - wp).max() <= 1e-6
            np.testing.assert_allclose(gp.sum(1), 1.0, atol=1e-5)
            # the guard really ran through the static list: low-margin rows carry fp32-grade logits, and they
            # agree with an fp32 engine
            s = np.sort(gl, 1)
            assert ((s[:, -1] - s[:, -2]) < 0.5 * float(e.cfg.guard_tau)).any()
        with sib.native.Engine(m.to_blob(), device=0, precision=sib.native.PREC_FP32, max_batch=8192) as f:
            fp, fl, fa = f.forward_sync(batches[1])
        assert (got[1][2] == fa).all()
        # ... and with the CPU oracle itself: argmax wherever the fp32 margin is not inside fp32 noise, logits within the
        # half-precision path's stated tolerance (tests/test_gpu_tc.py)
        om = oracle.Model(m.to_blob())
        for (gp, gl, ga), x in zip(got, batches):
            lg, pr = om.forward(x, threads=0)
            so = np.sort(lg.astype(np.float64), 1)
            clear = so[:, -1] - so[:, -2] > 1e-4
            assert (ga[clear] == lg.argmax(1)[clear]).all()
            assert np.abs(gl - lg).max() <= 3e-2 and np.abs(gp - pr).max() <= 6e-3
        # same result again on a third call (stale contents from call two)
        again = _run(e, e.forward_device_static, batches[0], torch)
        assert (again[2] == want[0][2]).all() and np.abs(again[1] - want[0][1]).max() <= 1e-6


def test_static_scratch_entry_rejects_what_it_cannot_hold(sib, presets):
    torch = pytest.importorskip("torch")
    m = presets["mnist-lenet5"]
    n_max = sib.native.STATIC_MAX_IMAGES
    buf = torch.zeros(16, dtype=torch.float32, device="cuda")
    with sib.native.Engine(m.to_blob(), device=0, precision=sib.native.PREC_FP16, max_batch=64, guard=True) as e:
        rc = e.forward_device_static(buf.data_ptr(), n_max + 1, buf.data_ptr(), 0, buf.data_ptr(), 0)
        assert rc == sib.native.ECAP
        assert str(n_max) in sib.native.last_error() and "static scratch" in sib.native.last_error()
    with sib.native.Engine(m.to_blob(), device=0, precision=sib.native.PREC_FP32, max_batch=64) as e:
        rc = e.forward_device_static(buf.data_ptr(), 4, buf.data_ptr(), 0, buf.data_ptr(), 0)
        assert rc == sib.native.EINVAL          # the fp32 path has no scratch to replace
