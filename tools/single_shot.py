#!/usr/bin/env python
"""Closed-loop single-tuple round trip through the C ABI (submit -> poll until the result is back), one tuple in
flight: the floor of the per-tuple latency, per entry point. Prints the median and the 99th percentile in us."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import sib_b200 as sib

m = sib.modelspec.make_preset(sys.argv[1] if len(sys.argv) > 1 else "cifar-lenet5")
blob = m.to_blob()
x = sib.modelspec.synthetic_images(m, 5, 64)
js = [sib.modelspec.instances_json(x[i:i + 1]) for i in range(64)]
N = 3000
for name, kw, submit in (("T1 sib_submit_f32", {}, lambda e, i: e.submit_f32(i, x[i % 64:i % 64 + 1])),
                         ("T2 sib_submit_json host decode", {}, lambda e, i: e.submit_json(i, js[i % 64])),
                         ("T2 sib_submit_json gpu decode", {"json_on_gpu": 1}, lambda e, i: e.submit_json(i, js[i % 64])),
                         ("T2 sib_submit_json auto decode", {"json_on_gpu": 2}, lambda e, i: e.submit_json(i, js[i % 64]))):
    with sib.native.Engine(blob, max_batch=64, deadline_us=200, **kw) as e:
        lat = []
        for i in range(N + 200):
            t0 = time.perf_counter_ns()
            assert submit(e, i) == 1
            while True:
                n, _ = e.poll_raw(4, 0)
                if n:
                    break
            lat.append(time.perf_counter_ns() - t0)
        lat = np.array(lat[200:]) / 1e3
        st = e.stats()
    print(json.dumps({"entry": name, "p50_us": float(np.percentile(lat, 50)), "p99_us": float(np.percentile(lat, 99)),
                      "json_gpu_decoded": int(st.json_gpu_decoded), "json_host_redone": int(st.json_host_redone)}), flush=True)
