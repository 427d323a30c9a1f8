#!/usr/bin/env python
"""Per-stage error budget of the half-precision forward pass (CPU emulation, no GPU needed).

Rounds, one source at a time and all together, what the tensor-core kernel rounds to half precision -- the input
pixels, each layer's weights, the activations p1 / p2 / h1 / h2 that pass through shared memory -- and reports the
max / rms logit error each source causes against the all-fp64 forward pass.  Accumulation is fp64 here (fp32 in the
kernel: its contribution is ~1e-6 and not the subject).  Run: python tools/error_budget.py [preset] [n] [fp16|bf16]
"""
import sys
import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
import sib_b200 as sib

preset = sys.argv[1] if len(sys.argv) > 1 else "cifar-lenet5"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
hp = {"fp16": torch.float16, "bf16": torch.bfloat16}[sys.argv[3] if len(sys.argv) > 3 else "fp16"]
m = sib.modelspec.make_preset(preset)
x = torch.from_numpy(sib.modelspec.synthetic_images(m, 62, n)).double().permute(0, 3, 1, 2)
W = m.weights


def rnd(t, on):
    return t.to(hp).double() if on else t


def forward(src):
    """src: set of rounding sources that are switched on."""
    t = rnd(x, "x" in src)
    flat = None
    names = ["c1", "c2", "f1", "f2", "f3"]
    acts = ["p1", "p2", "h1", "h2", None]
    for L, nm, act in zip(m.layers, names, acts):
        if L.kind == 1:
            filt = torch.from_numpy(W[L.w_off:L.w_off + L.a * L.b * L.c * L.d].reshape(L.a, L.b, L.c, L.d).copy()).double()
            bias = torch.from_numpy(W[L.b_off:L.b_off + L.d].copy()).double()
            t = F.conv2d(t, rnd(filt, "w" + nm in src).permute(3, 2, 0, 1).contiguous(), bias)
            t = rnd(F.max_pool2d(F.relu(t), 2, 2), act in src)
        else:
            if flat is None:
                flat = t.permute(0, 2, 3, 1).reshape(t.shape[0], -1)
            kern = torch.from_numpy(W[L.w_off:L.w_off + L.a * L.b].reshape(L.a, L.b).copy()).double()
            bias = torch.from_numpy(W[L.b_off:L.b_off + L.b].copy()).double()
            flat = flat @ rnd(kern, "w" + nm in src) + bias
            if L.flags & 1:
                flat = rnd(F.relu(flat), act in src)
    return flat


ref = forward(set())
allsrc = ["x", "wc1", "p1", "wc2", "p2", "wf1", "h1", "wf2", "h2", "wf3"]
print(f"{preset} n={n} {hp}: |logit| max {ref.abs().max():.2f}")
s = torch.sort(ref, 1).values
print("top1-top2 margin quantiles (1%, 2%, 5%):", [round(float(q), 4) for q in torch.quantile(s[:, -1] - s[:, -2], torch.tensor([0.01, 0.02, 0.05], dtype=torch.float64))])
rows = [(k, forward({k})) for k in allsrc] + [("all", forward(set(allsrc))), ("all but h2,wf3", forward(set(allsrc) - {"h2", "wf3"})),
                                                ("all but fc (wf*,h*)", forward({"x", "wc1", "p1", "wc2", "p2"})),
                                                ("all but conv1 (x,wc1)", forward(set(allsrc) - {"x", "wc1"}))]
for k, out in rows:
    e = (out - ref).abs()
    print(f"  {k:24s} max {float(e.max()):.2e}  rms {float((e ** 2).mean().sqrt()):.2e}")
