#!/usr/bin/env python
"""How wrong can the half-precision path be about the top-2 MARGIN (what the guard thresholds), as opposed to about
single logits?  Runs the tensor-core path unguarded and the fp32 path over N images and reports, per preset/precision:
max / high quantiles of |margin_tc - margin_fp32| for the fp32 top-2 classes, the largest tensor-core margin among
images whose argmax differs (the smallest tau that would have caught every flip), and how many images each candidate
tau sends to the guard.   usage: python tools/margin_error.py [n_images] > profiles/margin_error_r02.jsonl"""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
import sib_b200 as sib

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
B = 65536
for preset in ("cifar-lenet5", "mnist-lenet5"):
    m = sib.modelspec.make_preset(preset)
    blob = m.to_blob()
    for pname, prec in (("fp16", sib.native.PREC_FP16), ("bf16", sib.native.PREC_BF16)):
        e_m, flip_margin, margins = [], 0.0, []
        flips = 0
        with sib.native.Engine(blob, precision=prec, max_batch=B, guard=False) as et, sib.native.Engine(blob, precision=sib.native.PREC_FP32, max_batch=B) as ef:
            for b in range(N // B):
                x = sib.modelspec.synthetic_images(m, 1000 + b, B)
                _, lt, at = et.forward_sync(x)
                _, lf, af = ef.forward_sync(x)
                o = np.argsort(lf, 1)
                a, c = o[:, -1], o[:, -2]
                r = np.arange(B)
                mf = lf[r, a] - lf[r, c]
                mt = lt[r, a] - lt[r, c]
                e_m.append(np.abs(mt - mf))
                st = np.sort(lt, 1)
                m16 = st[:, -1] - st[:, -2]
                margins.append(m16)
                fl = at != af
                flips += int(fl.sum())
                if fl.any():
                    flip_margin = max(flip_margin, float(m16[fl].max()))
        e_m = np.concatenate(e_m)
        margins = np.concatenate(margins)
        rec = {"preset": preset, "precision": pname, "images": int(e_m.size), "margin_err_max": float(e_m.max()),
               "margin_err_p9999": float(np.quantile(e_m, 0.9999)), "margin_err_p99": float(np.quantile(e_m, 0.99)),
               "unguarded_argmax_flips": flips, "largest_tc_margin_among_flips": flip_margin,
               "listed_fraction_by_tau": {str(t): float((margins < t).mean()) for t in (0.02, 0.03, 0.04, 0.05, 0.1, 0.2, 0.3)}}
        print(json.dumps(rec), flush=True)
