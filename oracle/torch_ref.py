"""Independent second implementation of the forward pass on torch-CPU (oneDNN), used only to cross-check the C
oracle and to generate tests/golden/ (oracle/make_golden.py).  TEST INFRASTRUCTURE ONLY.

It shares no code with oracle/sib_oracle.c: weights are permuted HWIO->OIHW and the image NHWC->NCHW, the conv runs
through torch.nn.functional.conv2d, and the result is permuted back so that flatten order is TensorFlow's (h,w,c).
"""
from __future__ import annotations

import struct

import numpy as np
import torch
import torch.nn.functional as F


def parse_blob(blob: bytes):
    magic, ver, h, w, c, nl, nw, _ = struct.unpack_from("<4s7I", blob, 0)
    assert magic == b"SIBM" and ver == 1
    layers = [struct.unpack_from("<8I", blob, 32 + 32 * i) for i in range(nl)]
    wt = np.frombuffer(blob, "<f4", nw, 32 + 32 * nl)
    return (h, w, c), layers, wt


def forward(blob: bytes, x: np.ndarray, dtype=torch.float32, threads: int = 1):
    """x [N,H,W,C] -> (logits, probs) as numpy arrays of `dtype`."""
    (h, w, c), layers, wt = parse_blob(blob)
    torch.set_num_threads(threads)
    t = torch.from_numpy(np.ascontiguousarray(x, np.float32)).to(dtype).permute(0, 3, 1, 2)  # NCHW
    flat = None
    for kind, a, b, cin, cout, flags, w_off, b_off in layers:
        if kind == 1:
            filt = torch.from_numpy(wt[w_off:w_off + a * b * cin * cout].reshape(a, b, cin, cout).copy()).to(dtype)
            bias = torch.from_numpy(wt[b_off:b_off + cout].copy()).to(dtype)
            t = F.conv2d(t, filt.permute(3, 2, 0, 1).contiguous(), bias)  # cross-correlation, VALID, stride 1
            t = F.max_pool2d(F.relu(t), 2, 2)                            # floor mode == TF VALID
        else:
            if flat is None:
                flat = t.permute(0, 2, 3, 1).reshape(t.shape[0], -1)      # TF flatten order (h,w,c)
            kern = torch.from_numpy(wt[w_off:w_off + a * b].reshape(a, b).copy()).to(dtype)
            bias = torch.from_numpy(wt[b_off:b_off + b].copy()).to(dtype)
            flat = flat @ kern + bias
            if flags & 1:
                flat = F.relu(flat)
    logits = flat
    probs = torch.softmax(logits, dim=1)
    return logits.numpy(), probs.numpy()
