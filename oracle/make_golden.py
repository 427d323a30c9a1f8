"""Generates tests/golden/*.npz from the independent torch-CPU implementation (oracle/torch_ref.py).

    python oracle/make_golden.py

The reference itself holds no golden vectors for this path (no tests, no fixtures, no model file), so
these pin the oracle against a second implementation, not against the reference: parity stays "unpinned".
Inputs are the D2 synthetic images (seed 1234 + 0); weights are the D2 presets. Stored: fp64 logits/probs (torch
float64) and fp32 logits/probs (torch float32) for the first 16 images of each preset.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import sib_b200 as sib  # noqa: E402  (modelspec only: numpy generators)
import torch_ref  # noqa: E402

out = os.path.join(ROOT, "tests", "golden")
os.makedirs(out, exist_ok=True)
for name in sib.modelspec.PRESETS:
    m = sib.modelspec.make_preset(name)
    x = sib.modelspec.synthetic_images(m, 0, 16)
    lg32, pr32 = torch_ref.forward(m.to_blob(), x, torch.float32)
    lg64, pr64 = torch_ref.forward(m.to_blob(), x, torch.float64)
    np.savez_compressed(os.path.join(out, f"{name}.npz"), logits32=lg32, probs32=pr32, logits64=lg64, probs64=pr64,
                        x_checksum=np.float64(x.astype(np.float64).sum()), w_checksum=np.float64(m.weights.astype(np.float64).sum()))
    print(name, "argmax", lg64.argmax(1).tolist(), "logit range", float(lg64.min()), float(lg64.max()))
