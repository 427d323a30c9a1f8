#!/usr/bin/env python
"""Latency under fixed arrival rates (BASELINE.json config 5).

    python tools/latency_bench.py [--preset cifar-lenet5] [--seconds 1.0] [--full]

Open-loop arrivals (fixed interval; --poisson for exponential gaps) against (a) the B200 engine through the C ABI
(T2: JSON tuple in -> JSON message out; T1: fp32 image in -> probabilities out) and (b) the CPU path (oracle port of
execute(), `cores` executors). Latency = result available - scheduled arrival. One JSON line per point.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np  # noqa: E402


def gpu_point(sib, blob, payloads, mode, rate, seconds, producers, max_batch, deadline_us, poisson=False, json_on_gpu=False, strict=False):
    n = max(1000, int(rate * seconds))
    with sib.native.Engine(blob, max_batch=max_batch, deadline_us=deadline_us, n_slots=6, json_on_gpu=json_on_gpu, strict_batching=strict) as e:
        sib.native.loadgen(e, payloads, min(rate, 20000.0), 2000, mode=mode, producers=producers)  # warm-up
        r = sib.native.loadgen(e, payloads, rate, n, mode=mode, producers=producers, poisson=poisson)
        st = e.stats()
    return {"arm": "b200", "level": ("T2 json->json" + (" (gpu decode)" if json_on_gpu else " (host decode)")) if mode == sib.native.LG_JSON else "T1 f32->probs",
            "rate_tuples_s": rate, "tuples": n, "producers": producers, "max_batch": max_batch, "deadline_us": deadline_us, "strict_batching": bool(strict),
            "achieved_tuples_s": r["achieved_tuples_per_s"], "p50_us": r["p50_ns"] / 1e3, "p99_us": r["p99_ns"] / 1e3,
            "p999_us": r["p999_ns"] / 1e3, "max_us": r["max_ns"] / 1e3, "late_submits": r["late_submits"],
            "eagain": r["eagain_spins"], "batches": int(st.batches), "by_deadline": int(st.flush_by_deadline),
            "by_size": int(st.flush_by_size)}


def cpu_point(oracle, om, payloads, json_mode, rate, seconds, threads):
    n = max(500, int(min(rate, 50000) * seconds))
    lat = oracle.latency_run(om, payloads, rate, n, threads=threads, json_mode=json_mode)
    return {"arm": "cpu-port", "level": "T2 json->json" if json_mode else "T1 f32->probs", "rate_tuples_s": rate,
            "tuples": n, "executors": threads, "p50_us": float(np.percentile(lat, 50)) / 1e3,
            "p99_us": float(np.percentile(lat, 99)) / 1e3, "max_us": float(lat.max()) / 1e3,
            "saturated": bool(lat[-len(lat) // 10:].mean() > 4 * max(1.0, lat[:len(lat) // 10].mean()))}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="cifar-lenet5")
    ap.add_argument("--seconds", type=float, default=1.0)
    ap.add_argument("--full", action="store_true", help="whole cfg5 grid (max_batch x deadline)")
    ap.add_argument("--poisson", action="store_true")
    args = ap.parse_args()
    import oracle
    import sib_b200 as sib
    ms = sib.modelspec
    m = ms.make_preset(args.preset)
    blob = m.to_blob()
    om = oracle.Model(blob)
    cores = oracle.max_threads()
    x = ms.synthetic_images(m, 9, 64)
    jp = [ms.instances_json(x[i:i + 1]) for i in range(64)]
    fp = [x[i] for i in range(64)]
    prod = max(1, min(8, cores - 2))
    grid = [(64, 50), (512, 200), (4096, 1000)] if args.full else [(512, 200)]
    for mb, dl in grid:
        for rate in (10_000, 100_000):
            print(json.dumps(gpu_point(sib, blob, jp, sib.native.LG_JSON, rate, args.seconds, prod, mb, dl, args.poisson)), flush=True)
        for rate in (10_000, 100_000, 400_000):
            print(json.dumps(gpu_point(sib, blob, jp, sib.native.LG_JSON, rate, args.seconds, prod, mb, dl, args.poisson, json_on_gpu=True)), flush=True)
        for rate in (10_000, 100_000, 1_000_000):
            print(json.dumps(gpu_point(sib, blob, fp, sib.native.LG_F32, rate, args.seconds, prod, mb, dl, args.poisson)), flush=True)
    # sib_config.strict_batching = 1: batches leave by size and deadline only (the default also launches when the GPU is idle)
    for rate in (1_000, 10_000, 100_000, 400_000):
        print(json.dumps(gpu_point(sib, blob, jp, sib.native.LG_JSON, rate, args.seconds, prod, 512, 200, args.poisson, json_on_gpu=True, strict=True)), flush=True)
    for rate in (1_000, 10_000, 100_000, 1_000_000):
        print(json.dumps(gpu_point(sib, blob, fp, sib.native.LG_F32, rate, args.seconds, prod, 512, 200, args.poisson, strict=True)), flush=True)
    for rate in (1_000, 10_000):
        print(json.dumps(cpu_point(oracle, om, jp, True, rate, args.seconds, cores)), flush=True)
    for rate in (10_000, 100_000):
        print(json.dumps(cpu_point(oracle, om, fp, False, rate, args.seconds, cores)), flush=True)


if __name__ == "__main__":
    main()
