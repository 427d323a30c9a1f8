"""Which images of the 2^20-image sweep get another argmax than the CPU path, and how close their fp32 top-2 logits are.

    python tools/diag_mismatch.py [preset] > profiles/mismatch_<preset>.jsonl      (GPU box)
"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import numpy as np
import oracle, sib_b200 as sib
m = sib.modelspec.make_preset(sys.argv[1] if len(sys.argv) > 1 else "mnist-lenet5")
blob = m.to_blob()
n = 1 << 20
x = np.concatenate([sib.modelspec.synthetic_images(m, bi, 65536) for bi in range(n // 65536)])
om = oracle.Model(blob)
lg32, _ = om.forward(x, threads=oracle.max_threads())
ref = lg32.argmax(1)
with sib.native.Engine(blob, max_batch=65536) as e:
    got = np.concatenate([e.forward_sync(x[i:i + 65536])[2] for i in range(0, n, 65536)])
    lgt = np.concatenate([e.forward_sync(x[i:i + 65536])[1] for i in range(0, n, 65536)])
bad = np.nonzero(got != ref)[0]
print(json.dumps({"preset": m.name, "images": n, "mismatches": [int(i) for i in bad]}))
with sib.native.Engine(blob, max_batch=65536, precision=sib.native.PREC_FP32) as e32, sib.native.Engine(blob, max_batch=65536, guard=False) as e16:
    for i in bad:
        xi = x[i:i + 1]
        l64 = om.forward64(xi)[0] if hasattr(om, 'forward64') else None
        p32, l32g, a32 = e32.forward_sync(xi)
        p16, l16g, a16 = e16.forward_sync(xi)
        s = np.sort(lg32[i])[::-1]
        print(json.dumps({"index": int(i), "oracle_fp32_argmax": int(ref[i]), "guarded_argmax": int(got[i]), "gpu_fp32_argmax": int(a32[0]), "gpu_fp16_argmax": int(a16[0]),
                          "oracle_fp32_top2_margin": float(s[0] - s[1]), "oracle_fp32_logits": [float(v) for v in lg32[i]],
                          "guarded_logits": [float(v) for v in lgt[i]], "gpu_fp32_logits": [float(v) for v in l32g[0]], "gpu_fp16_logits": [float(v) for v in l16g[0]],
                          "oracle_fp64": (np.asarray(l64).ravel().tolist() if l64 is not None else None)}))
