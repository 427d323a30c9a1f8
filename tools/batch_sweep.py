#!/usr/bin/env python
"""Micro-batch sweep (BASELINE.json config 2): mnist-lenet5 on one GPU, micro-batch in
{1, 2, 4, ..., 4096} (+ 65536), 2^20 images per point, every prediction diffed against the CPU path's (cfg1) argmax.

    python tools/batch_sweep.py [--preset mnist-lenet5] [--images 1048576] > profiles/batch_sweep_r01.jsonl

T0 (device-resident input, sib_forward_device, guard on) is timed with CUDA events on the launching stream; the
python/ctypes call overhead per launch is inside the timed region for the small batches, as it would be for any
host caller. One JSON line per point; `argmax_mismatches` must be 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="mnist-lenet5")
    ap.add_argument("--images", type=int, default=1 << 20)
    ap.add_argument("--max-launches", type=int, default=1 << 17, help="cap on launches per point (small batches)")
    args = ap.parse_args()
    import torch

    import oracle
    import sib_b200 as sib

    m = sib.modelspec.make_preset(args.preset)
    blob = m.to_blob()
    n = args.images
    x = np.concatenate([sib.modelspec.synthetic_images(m, bi, 65536) for bi in range((n + 65535) // 65536)])[:n]
    t0 = time.time()
    ref_logits, _ = oracle.Model(blob).forward(x, threads=oracle.max_threads())
    ref_arg = ref_logits.argmax(axis=1).astype(np.int32)
    cpu_s = time.time() - t0
    print(json.dumps({"arm": "cpu-port", "preset": m.name, "images": n, "threads": oracle.max_threads(),
                      "images_per_s": n / cpu_s}), flush=True)

    dev = torch.device("cuda:0")
    d_in = torch.from_numpy(x).to(dev)
    d_probs = torch.empty((n, m.n_classes), dtype=torch.float32, device=dev)
    d_arg = torch.empty(n, dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream()
    img_b, prob_b = m.img_elems * 4, m.n_classes * 4
    with sib.native.Engine(blob, max_batch=65536) as e:
        for b in [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 65536]:
            launches = min(n // b, args.max_launches)
            done = launches * b
            d_arg.fill_(-1)
            with torch.cuda.stream(stream):
                for i in range(min(launches, 50)):  # warm-up
                    e.forward_device(d_in.data_ptr() + i * b * img_b, b, d_probs.data_ptr() + i * b * prob_b, 0,
                                     d_arg.data_ptr() + i * b * 4, stream.cuda_stream)
                stream.synchronize()
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record(stream)
                for i in range(launches):
                    e.forward_device(d_in.data_ptr() + i * b * img_b, b, d_probs.data_ptr() + i * b * prob_b, 0,
                                     d_arg.data_ptr() + i * b * 4, stream.cuda_stream)
                ev1.record(stream)
                stream.synchronize()
            ms = ev0.elapsed_time(ev1)
            got = d_arg[:done].cpu().numpy()
            mism = int((got != ref_arg[:done]).sum())
            print(json.dumps({"arm": "b200", "level": "T0 sib_forward_device (fp16 + fp32 guard)", "preset": m.name,
                              "micro_batch": b, "launches": launches, "images": done, "images_per_s": done / (ms / 1e3),
                              "us_per_launch": ms * 1e3 / launches, "argmax_mismatches": mism}), flush=True)


if __name__ == "__main__":
    main()
