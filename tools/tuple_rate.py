#!/usr/bin/env python
"""Highest sustained rate of single-image tuples through the C ABI (T1 fp32 tuples and T2 JSON tuples), i.e. the rate
a set of Storm executors can push through one engine. Offered load is set above capacity; achieved rate is reported."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import sib_b200 as sib

m = sib.modelspec.make_preset(sys.argv[1] if len(sys.argv) > 1 else "cifar-lenet5")
blob = m.to_blob()
x = sib.modelspec.synthetic_images(m, 5, 256)
f32 = [np.ascontiguousarray(x[i:i + 1]) for i in range(256)]
js = [sib.modelspec.instances_json(x[i:i + 1]) for i in range(64)]
for mode, payloads, name, gpu_json in ((sib.native.LG_F32, f32, "T1 f32 tuple", False), (sib.native.LG_JSON, js, "T2 json tuple (gpu decode)", 1),
                                     (sib.native.LG_JSON, js, "T2 json tuple (auto decode)", 2)):
    for producers in (2, 4, 8):
        with sib.native.Engine(blob, max_batch=4096, deadline_us=200, n_slots=6, json_on_gpu=gpu_json) as e:
            sib.native.loadgen(e, payloads, 50000.0, 5000, mode=mode, producers=producers)
            r = sib.native.loadgen(e, payloads, 5e6, 600000, mode=mode, producers=producers)
        print(json.dumps({"level": name, "producers": producers, "achieved_tuples_per_s": r["achieved_tuples_per_s"],
                          "p50_us": r["p50_ns"] / 1e3, "eagain_spins": r["eagain_spins"]}), flush=True)
