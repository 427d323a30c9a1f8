import os, sys, json
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import numpy as np
import oracle, sib_b200 as sib
m = sib.modelspec.make_preset("mnist-lenet5")
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
print("mismatches", bad)
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
