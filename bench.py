#!/usr/bin/env python
"""bench.py -- images/sec of the InferenceBolt hot path (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--preset cifar-lenet5]

A "step" is one pass of the hot path over one batch of synthetic images.  Reported on ONE JSON line:
  value      T0: device-resident fp32 NHWC -> probabilities, whole job (all ranks), CUDA events, max over ranks
  e2e        T1: the same metric through the C ABI with HOST buffers (sib_submit_f32 / sib_poll): pinned staging,
             H2D, kernels, D2H inside the timed region
  roofline   dominant kernel vs MEASURED_PEAKS.json (HBM): algorithmic bytes = (H*W*C*4 + n_classes*4) per image
  cpu_baseline  the CPU oracle (a port of the reference's per-tuple execute(), batch 1 per call) on the host cores
Under torchrun (N>1) every rank drives its own GPU with its own engine (executor -> GPU = task index mod N,
MainTopology.java:62); no collective touches the data path, torch.distributed is the barrier and the max.

--impl reference times the reference's CPU implementation of the path (the oracle port: the reference itself is
Java + TF-Java and cannot run here) on all host cores, rank 0 only.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

METRIC = "images/sec"


def _modelspec():
    """modelspec.py alone (numpy only), WITHOUT importing the package: the package's __init__ dlopens libsib.so, and the
    reference arm must not have the product library in its process."""
    import importlib.util
    if "sib_modelspec_only" in sys.modules:
        return sys.modules["sib_modelspec_only"]
    path = os.path.join(ROOT, "distributed-inference-system-based-storm_b200", "modelspec.py")
    spec = importlib.util.spec_from_file_location("sib_modelspec_only", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["sib_modelspec_only"] = mod
    spec.loader.exec_module(mod)
    return mod


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--preset", default="cifar-lenet5")
    ap.add_argument("--batch", type=int, default=65536, help="images per step per GPU")
    ap.add_argument("--precision", default="auto", choices=["auto", "fp32", "fp16", "bf16"])
    ap.add_argument("--no-guard", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline budget")
    ap.add_argument("--submit-threads", type=int, default=0, help="host threads feeding one GPU in the e2e leg")
    ap.add_argument("--e2e-copy", action="store_true", help="e2e leg through sib_submit_f32 (pageable source, staged copy)")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-latency", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------- clocks

class ClockSampler:
    """nvidia-smi sampling during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.lines: list[str] = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------- CPU arm

def cpu_per_tuple(model, seconds: float, threads: int = 0):
    """The reference's shape on the host cores: one forward per image (batch 1 per call, InferenceBolt.java:81-85),
    `threads` executors each owning a contiguous share. Bounded by wall time. -> (img/s, cores, sample text)."""
    import oracle
    om = oracle.Model(model.to_blob())
    cores = oracle.max_threads() if threads <= 0 else threads
    chunk = 256 * cores
    x = _modelspec().synthetic_images(model, 0, chunk)
    om.forward(x[:cores], threads=cores)  # warm
    done, t0 = 0, time.perf_counter()
    while True:
        om.forward(x, threads=cores)
        done += chunk
        dt = time.perf_counter() - t0
        if dt >= seconds:
            break
    return done / dt, cores, f"{done} images of {model.name} (D2 seed 1234), batch 1 per call, {dt:.1f} s wall"


def cpu_generous(model, seconds: float):
    """'Generous' CPU mode: torch-CPU (oneDNN) batched, b=256, all cores, model only. Independent of the oracle's C."""
    try:
        import torch
        import torch_ref
        torch.set_num_threads(os.cpu_count() or 8)
        x = _modelspec().synthetic_images(model, 0, 256)
        blob = model.to_blob()
        torch_ref.forward(blob, x)
        done, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            torch_ref.forward(blob, x, threads=os.cpu_count() or 8)
            done += 256
        return done / (time.perf_counter() - t0)
    except Exception as ex:  # noqa: BLE001
        return f"unavailable: {ex}"


def cpu_t2_json(model, seconds: float):
    """The bolt's real unit of work on the host cores: JSON tuple in -> JSON message out through the oracle's port of
    execute() (InferenceBolt.java:69-100), one executor per core, one image per tuple. -> (tuples/s, cores, sample)."""
    import oracle
    ms = _modelspec()
    om = oracle.Model(model.to_blob())
    cores = oracle.max_threads()
    x = ms.synthetic_images(model, 9, 32)
    jp = [ms.instances_json(x[i:i + 1]) for i in range(32)]
    batch = jp * max(1, (16 * cores) // 32)
    om.run_tuples(batch[:cores], threads=cores)
    done, t0 = 0, time.perf_counter()
    while True:
        om.run_tuples(batch, threads=cores)
        done += len(batch)
        dt = time.perf_counter() - t0
        if dt >= seconds:
            break
    return done / dt, cores, f"{done} one-image JSON tuples of {model.name} ({len(jp[0])} B each) in {dt:.1f} s"


def tuple_rates(sib, model, blob, prec, guard, dev, rank, world, dist):
    """What a Storm executor actually hands the bolt, flat out, at EVERY N (each rank drives its own GPU; whole-job rate =
    tuples of all ranks / slowest rank's time): one-image fp32 tuples from pageable memory (a JVM heap float[]), and
    one-image JSON tuples in -> JSON messages out with the host decoder and with the GPU decoder."""
    from importlib import import_module
    sharding = import_module("sib_b200.sharding")
    native = sib.native
    ms = sib.modelspec
    cores = os.cpu_count() or 8
    prod = max(1, min(6, cores // max(1, world) - 1))
    x = ms.synthetic_images(model, 9 + rank, 32)
    jp = [ms.instances_json(x[i:i + 1]) for i in range(32)]
    fp = [x[i] for i in range(32)]
    out = {"producers_per_gpu": prod, "json_bytes_per_tuple": int(np.mean([len(j) for j in jp])), "images_per_tuple": 1}

    def one(key, payloads, mode, n, **kw):
        with native.Engine(blob, device=dev, task_index=rank, max_batch=4096, deadline_us=200, precision=prec, n_slots=6, guard=guard, **kw) as e:
            native.loadgen(e, payloads, 20000.0, 2000, mode=mode, producers=prod)
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            r = native.loadgen(e, payloads, 5e7, n, mode=mode, producers=prod)
            dt = time.perf_counter() - t0
        u, t = sharding.aggregate(int(r["completed"]), dt, dist)
        out[key] = u / t

    one("t1_single_image_f32_tuples_per_s", fp, native.LG_F32, 400000)
    one("t2_json_tuples_per_s_host_decode", jp, native.LG_JSON, 40000)
    one("t2_json_tuples_per_s_gpu_decode", jp, native.LG_JSON, 200000, json_on_gpu=True)
    return out


def t2_and_latency(sib, model, blob, prec, guard):
    """Bolt-level numbers through the C ABI's load generator (include/sib_loadgen.h), N=1 GPU only.
    T2 saturation: JSON tuples in -> JSON messages out, producers running flat out.
    Latency: open-loop fixed-interval arrivals; result-available minus scheduled arrival; CPU path beside it."""
    import oracle
    native = sib.native
    ms = sib.modelspec
    cores = os.cpu_count() or 8
    prod = max(1, min(8, cores - 2))
    x = ms.synthetic_images(model, 9, 32)
    jp = [ms.instances_json(x[i:i + 1]) for i in range(32)]
    fp = [x[i] for i in range(32)]
    out = {"producers": prod, "json_bytes_per_tuple": int(np.mean([len(j) for j in jp]))}
    with native.Engine(blob, max_batch=4096, deadline_us=200, precision=prec, n_slots=6, guard=guard) as e:
        native.loadgen(e, jp, 20000.0, 2000, mode=native.LG_JSON, producers=prod)
        r = native.loadgen(e, jp, 5e6, 40000, mode=native.LG_JSON, producers=prod)
        out["t2_json_tuples_per_s_saturated_host_decode"] = r["achieved_tuples_per_s"]
    with native.Engine(blob, max_batch=4096, deadline_us=200, precision=prec, n_slots=6, guard=guard, json_on_gpu=True) as e:
        native.loadgen(e, jp, 20000.0, 2000, mode=native.LG_JSON, producers=prod)
        r = native.loadgen(e, jp, 5e6, 200000, mode=native.LG_JSON, producers=prod)
        out["t2_json_tuples_per_s_saturated_gpu_decode"] = r["achieved_tuples_per_s"]
    lat = {}
    with native.Engine(blob, max_batch=512, deadline_us=200, precision=prec, n_slots=6, guard=guard) as e:
        native.loadgen(e, jp, 10000.0, 2000, mode=native.LG_JSON, producers=prod)
        r = native.loadgen(e, jp, 10000.0, 10000, mode=native.LG_JSON, producers=prod)
        lat["b200_T2_json_10k_per_s"] = {"p50_us": r["p50_ns"] / 1e3, "p99_us": r["p99_ns"] / 1e3, "deadline_us": 200}
        r = native.loadgen(e, fp, 100000.0, 100000, mode=native.LG_F32, producers=prod)
        lat["b200_T1_f32_100k_per_s"] = {"p50_us": r["p50_ns"] / 1e3, "p99_us": r["p99_ns"] / 1e3, "deadline_us": 200}
    with native.Engine(blob, max_batch=512, deadline_us=200, precision=prec, n_slots=6, guard=guard, json_on_gpu=True) as e:
        native.loadgen(e, jp, 10000.0, 2000, mode=native.LG_JSON, producers=prod)
        r = native.loadgen(e, jp, 100000.0, 100000, mode=native.LG_JSON, producers=prod)
        lat["b200_T2_json_100k_per_s_gpu_decode"] = {"p50_us": r["p50_ns"] / 1e3, "p99_us": r["p99_ns"] / 1e3, "deadline_us": 200,
                                                     "late_submits": r["late_submits"]}
    with native.Engine(blob, max_batch=64, deadline_us=50, precision=prec, n_slots=6, guard=guard) as e:
        native.loadgen(e, fp, 10000.0, 2000, mode=native.LG_F32, producers=prod)
        r = native.loadgen(e, fp, 100000.0, 100000, mode=native.LG_F32, producers=prod)
        lat["b200_T1_f32_100k_per_s_deadline50"] = {"p50_us": r["p50_ns"] / 1e3, "p99_us": r["p99_ns"] / 1e3, "deadline_us": 50}
    om = oracle.Model(blob)
    l = oracle.latency_run(om, jp, 10000.0, 10000, threads=cores, json_mode=True)
    lat["cpu_T2_json_10k_per_s"] = {"p50_us": float(np.percentile(l, 50)) / 1e3, "p99_us": float(np.percentile(l, 99)) / 1e3, "executors": cores}
    l = oracle.latency_run(om, fp, 10000.0, 10000, threads=cores, json_mode=False)
    lat["cpu_T1_f32_10k_per_s"] = {"p50_us": float(np.percentile(l, 50)) / 1e3, "p99_us": float(np.percentile(l, 99)) / 1e3, "executors": cores}
    out["latency"] = lat
    return out


def run_reference(args):
    """--impl reference: the CPU port of execute()'s model call on all host cores; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    model = _modelspec().make_preset(args.preset)  # numpy only: libsib.so is never loaded in this process
    per_step = max(1.0, min(8.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_per_tuple(model, 0.2)
    vals = []
    t0 = time.perf_counter()
    sample = ""
    for _ in range(args.steps):
        v, cores, sample = cpu_per_tuple(model, per_step)
        vals.append(v)
    wall = time.perf_counter() - t0
    value = statistics.mean(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, model),
        "cpu_baseline": {"value": value, "unit": "images/s", "cores": cores, "kind": "port",
                         "sample": f"per step: {sample}; oracle/sib_oracle.c (TF-Java/Jackson cannot run: no JVM)"},
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, model) -> dict:
    return {"workload": f"{model.name} ({model.h}x{model.w}x{model.c} -> softmax {model.n_classes}), "
                        f"{args.batch} images per step per GPU, SURVEY.md D2 synthetic pixels and weights",
            "preset": model.name, "images_per_step_per_gpu": args.batch,
            "parallelism": f"executor-per-gpu x{args.gpus}, shuffle grouping, no collective",
            "l2": "4 rotating input buffers per GPU, each step's batch is read from HBM (4 x batch x image bytes > 126 MB L2)"}


# ------------------------------------------------------------------------------------------- GPU arm

def run_ours(args):
    import torch
    import torch.distributed as dist

    import sib_b200 as sib
    from importlib import import_module
    sharding = import_module("sib_b200.sharding")
    native = sib.native

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ngpu = torch.cuda.device_count()
    dev = sharding.device_for_task(local, ngpu)
    torch.cuda.set_device(dev)

    model = sib.modelspec.make_preset(args.preset)
    blob = model.to_blob()
    prec_name = args.precision
    if prec_name == "auto":
        prec_name = sib.default_bench_precision()
    prec = {"fp32": native.PREC_FP32, "bf16": native.PREC_BF16, "fp16": native.PREC_FP16}[prec_name]
    B = args.batch
    eng = native.Engine(blob, device=dev, task_index=rank, max_batch=B, deadline_us=200, precision=prec,
                        n_slots=4, guard=not args.no_guard)
    nc, img = model.n_classes, model.img_elems

    # ---- T0: device-resident. 4 rotating batches so every step streams its input from HBM, not L2.
    R = 4
    host_batches = []
    for r in range(R):  # page-locked host memory: what a bolt's direct buffers are
        t = torch.empty((B, model.h, model.w, model.c), dtype=torch.float32, pin_memory=True)
        t.numpy()[:] = sib.modelspec.synthetic_images(model, 100 + rank * R + r, B)
        host_batches.append(t)
    d_in = [h.cuda() for h in host_batches]
    d_probs = torch.empty((B, nc), dtype=torch.float32, device="cuda")
    d_argmax = torch.empty((B,), dtype=torch.int32, device="cuda")
    stream = torch.cuda.Stream()  # a real (non-null) stream: kernels and timing events share it
    stream.wait_stream(torch.cuda.current_stream())

    def step(i):
        eng.forward_device(d_in[i % R].data_ptr(), B, d_probs.data_ptr(), 0, d_argmax.data_ptr(), stream.cuda_stream)

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches0 = eng.stats().kernel_launches
    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    torch.cuda.synchronize()
    evs[0].record(stream)
    for i in range(args.steps):
        step(i)
        evs[i + 1].record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.stats().kernel_launches - launches0
    t0_local = evs[0].elapsed_time(evs[-1]) * 1e-3
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    units, secs = sharding.aggregate(B * args.steps, t0_local, dist if world > 1 else None)
    value = units / secs
    # ---- the same step, sustained: blocks of `steps` launches until >= 2 s have run, its own clock record. The 20-step
    # region above is ~7 ms; this is what the kernel does once the chip is warm and power-managed.
    sustained = None
    if not os.environ.get("SIB_BENCH_NO_SUSTAINED"):
        samp2 = ClockSampler(dev)
        if rank == 0:
            samp2.start()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        n_sus, t_w = 0, time.perf_counter()
        while time.perf_counter() - t_w < 2.0:
            for i in range(args.steps):
                step(n_sus + i)
            n_sus += args.steps
            stream.synchronize()
        e1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        c2 = samp2.stop() if rank == 0 else None
        u_s, t_s = sharding.aggregate(B * n_sus, e0.elapsed_time(e1) * 1e-3, dist if world > 1 else None)
        sustained = {"value": u_s / t_s, "unit": "images/s", "steps": n_sus, "seconds": t_s, "clocks": c2}
    torch.cuda.current_stream().wait_stream(stream)
    checksum = float(d_probs.sum().item())  # every row sums to 1: B
    assert os.environ.get("SIB_BENCH_NOCHECK") or abs(checksum - B) < 1e-2 * B, checksum  # (diagnostic kernel builds skip the check)

    # ---- dominant kernel alone (forward_tc_kernel): same batches through a guard-off engine, CUDA events on the
    # launching stream. The step time above also holds the guard's recompute kernel; the roofline is the kernel's.
    tc_ms = None
    if prec != native.PREC_FP32 and not args.no_guard:
        eng0 = native.Engine(blob, device=dev, task_index=rank, max_batch=B, precision=prec, n_slots=2, guard=False)
        with torch.cuda.stream(stream):
            for i in range(args.warmup):
                eng0.forward_device(d_in[i % R].data_ptr(), B, d_probs.data_ptr(), 0, d_argmax.data_ptr(), stream.cuda_stream)
            ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
            torch.cuda.synchronize()
            ev2[0].record(stream)
            for i in range(args.steps):
                eng0.forward_device(d_in[i % R].data_ptr(), B, d_probs.data_ptr(), 0, d_argmax.data_ptr(), stream.cuda_stream)
                ev2[i + 1].record(stream)
            torch.cuda.synchronize()
        tc_ms = statistics.median(ev2[i].elapsed_time(ev2[i + 1]) for i in range(args.steps))
        eng0.close()

    # ---- the same kernel at R x B images per launch (N = 1 only): a launch costs ~17 us of set-up, fill and drain
    # whatever its size (profiles/fixed_cost_r02_*.jsonl), so this is what the kernel does once that is amortised. The
    # R rotating batches become one contiguous input (R x B images >> L2), read from HBM by every launch.
    large = None
    if tc_ms is not None and world == 1 and not os.environ.get("SIB_BENCH_NO_LARGE"):
        try:
            BL = R * B
            x_big = torch.cat(d_in, dim=0)
            p_big = torch.empty((BL, model.n_classes), dtype=torch.float32, device=dev)
            a_big = torch.empty(BL, dtype=torch.int32, device=dev)
            engL = native.Engine(blob, device=dev, task_index=rank, max_batch=BL, precision=prec, n_slots=1, guard=False)
            with torch.cuda.stream(stream):
                for _ in range(3):
                    engL.forward_device(x_big.data_ptr(), BL, p_big.data_ptr(), 0, a_big.data_ptr(), stream.cuda_stream)
                ev3 = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
                torch.cuda.synchronize()
                ev3[0].record(stream)
                for i in range(10):
                    engL.forward_device(x_big.data_ptr(), BL, p_big.data_ptr(), 0, a_big.data_ptr(), stream.cuda_stream)
                    ev3[i + 1].record(stream)
                torch.cuda.synchronize()
            big_ms = statistics.median(ev3[i].elapsed_time(ev3[i + 1]) for i in range(10))
            engL.close()
            assert abs(float(p_big.sum().item()) - BL) < 1e-2 * BL
            large = {"images_per_launch": BL, "kernel_ms": big_ms, "launches": 10}
            del x_big, p_big, a_big
        except Exception as ex:  # never lose the bench line to this leg
            large = {"error": str(ex)[:200]}

    # ---- T1: host buffers through the C ABI. Like a Storm worker with several executors feeding one GPU, a few
    # host threads submit chunks of fp32 images (sib_submit_f32 copies them into the pinned slot outside the engine
    # lock; ctypes releases the GIL for the call) while this thread polls results.
    e2e = None
    if not args.skip_e2e:
        import threading
        chunk = min(1024, B)
        per_step = B // chunk
        cores = os.cpu_count() or 8
        dflt_thr = 8 if args.e2e_copy else 2
        want_thr = args.submit_threads if args.submit_threads > 0 else min(dflt_thr, max(1, cores // max(1, world) - 1))
        n_thr = max(1, min(want_thr, per_step))

        def e2e_steps(n_steps, first):
            total = n_steps * per_step * chunk
            err = []

            def worker(t):
                try:
                    for s_ in range(n_steps):
                        h = host_batches[(first + s_) % R]
                        base = h.data_ptr()
                        hn = h.numpy()
                        for c in range(t, per_step, n_thr):
                            while True:
                                if args.e2e_copy:
                                    rc = eng.submit_f32(s_ * per_step + c, hn[c * chunk:(c + 1) * chunk])
                                else:
                                    rc = eng.submit_pinned(s_ * per_step + c, base + c * chunk * img * 4, chunk)
                                if rc == native.EAGAIN:
                                    time.sleep(0.00005)
                                    continue
                                if rc < 0:
                                    raise native.SibError(rc, "sib_submit")
                                break
                except Exception as ex:  # noqa: BLE001
                    err.append(ex)

            ths = [threading.Thread(target=worker, args=(t,)) for t in range(n_thr)]
            for t in ths:
                t.start()
            got = 0
            while got < total and not err:
                got += _poll_count(eng, 2000)
                if not any(t.is_alive() for t in ths):
                    eng.flush()
            for t in ths:
                t.join()
            if err:
                raise err[0]
            return got

        # the link's own ceiling, measured now: pinned cudaMemcpyAsync H2D of the same batches, every rank at once
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            d_in[0].copy_(host_batches[0], non_blocking=True)
            c0.record(stream)
            reps = max(2, int(2e9 // (B * img * 4)))
            for r_ in range(reps):
                d_in[r_ % R].copy_(host_batches[r_ % R], non_blocking=True)
            c1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ub, tb = sharding.aggregate(reps * B * img * 4, c0.elapsed_time(c1) * 1e-3, dist if world > 1 else None)
        h2d_roof_gbs = ub / tb * 1e-9

        e2e_steps(max(1, min(args.warmup, 3)), 0)
        st0 = eng.stats()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t_a = time.perf_counter()
        n_e2e = max(2, min(args.steps, 10))
        done = e2e_steps(n_e2e, 1)
        torch.cuda.synchronize()
        t_b = time.perf_counter()
        if world > 1:
            dist.barrier()
        st1 = eng.stats()
        u2, s2 = sharding.aggregate(done, t_b - t_a, dist if world > 1 else None)
        e2e = {"value": u2 / s2, "unit": "images/s",
               "level": (f"T1: sib_submit_f32 (staged copy)" if args.e2e_copy else "T1: sib_submit_pinned (zero-copy, page-locked host fp32)")
                        + f" + sib_poll, {chunk} images per tuple, {n_thr} submitting threads per GPU",
               "h2d_bytes_per_step": (st1.h2d_bytes - st0.h2d_bytes) // n_e2e,
               "d2h_bytes_per_step": (st1.d2h_bytes - st0.d2h_bytes) // n_e2e, "steps": n_e2e,
               "images_per_tuple": chunk,
               # T1's binding roof is the host->device link, not HBM: all ranks' pinned cudaMemcpyAsync at once, this run
               "roof_gbs": h2d_roof_gbs, "roof": "pinned H2D copy of the same batches, all ranks concurrently, measured in this run",
               "achieved_gbs": u2 / s2 * img * 4 * 1e-9, "frac": (u2 / s2 * img * 4 * 1e-9) / h2d_roof_gbs}

    guard_n = eng.stats().guard_recomputed
    eng.close()
    tuples = None
    if not args.skip_latency:
        tuples = tuple_rates(sib, model, blob, prec, not args.no_guard, dev, rank, world, dist if world > 1 else None)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    bpi = model.bytes_per_image()
    step_med = statistics.median(step_ms)   # one step = forward kernel launch (+ the guard's recompute launch)
    k_ms = tc_ms if tc_ms is not None else step_med
    achieved = B * bpi / (k_ms * 1e-3) * 1e-9
    traffic, traffic_src = None, None
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):  # NOT measured by this run: dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture
        try:
            tj = json.load(open(tr))
            traffic = tj.get(f"{model.name}:{prec_name}:{B}")
            traffic_src = tj.get("_source")
        except Exception:
            traffic = None
    # which roof binds this model: HBM (bytes per image) or the tensor pipe (flop per image), SURVEY.md 8d
    tf_peak = float(peaks.get("bf16_tflops", 1590.0))
    fpi = model.flops_per_image()
    tensor_bound = fpi / (tf_peak * 1e12) > bpi / (hbm_peak * 1e9)
    if tensor_bound:
        ach_tf = B * fpi / (k_ms * 1e-3) * 1e-12
        roof = {"bound": "tensor", "achieved": ach_tf, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach_tf / tf_peak,
                "peak_source": ("measured (MEASURED_PEAKS.json bf16_tflops, burst)" if "bf16_tflops" in peaks else "fallback 1590 TFLOP/s"),
                "hbm_frac": achieved / hbm_peak}
    else:
        roof = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "peak_source": peak_src}
    kernel_name = ("forward_fp32_kernel" if prec_name == "fp32" else
                   "forward_tc_kernel" if model.name in ("mnist-lenet5", "cifar-lenet5") else "forward_tcgen_kernel")
    line = {
        "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": {"fp32": "f32", "fp16": "f16 operands, f32 accumulate",
                                       "bf16": "bf16 operands, f32 accumulate"}[prec_name],
        "data": "synthetic", "config": workload_config(args, model),
        "e2e": e2e, "gpu_launches": int(launches),
        "roofline": {**roof, "traffic": traffic, "traffic_source": traffic_src,
                     "step_frac": roof["frac"] * k_ms / step_med,   # forward + guard kernels: what `value` is made of
                     "kernel": kernel_name,
                     "kernel_ms": k_ms, "step_ms": step_med,
                     "guard_kernel_ms": (step_med - tc_ms) if tc_ms is not None else 0.0,
                     "algorithmic_bytes_per_image": bpi, "images_per_launch": B,
                     "flops_per_image": model.flops_per_image(),
                     "large_launch": _large_launch(large, roof, bpi, model.flops_per_image())},
        "clocks": clocks, "guard_recomputed_images": int(guard_n),
        # (counted on the submit/poll path, i.e. over the e2e leg's images)
        "guard_recomputed_fraction": (float(guard_n) / float(B * (e2e["steps"] + max(1, min(args.warmup, 3)))) if e2e else None),
        "value_sustained": sustained, "tuples": tuples,
    }
    if not args.skip_cpu:
        v, cores, sample = cpu_per_tuple(model, args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample}
        # the other CPU shapes BASELINE.md asks for, a few seconds each
        v1, _, _ = cpu_per_tuple(model, 2.0, threads=1)
        v4, _, _ = cpu_per_tuple(model, 2.0, threads=4)  # INFERENCE_BOLT_PARAL = 4 (MainTopology.java:27)
        line["cpu_baseline"]["per_tuple_1_thread"] = v1
        line["cpu_baseline"]["per_tuple_4_threads"] = v4
        line["cpu_baseline"]["generous_torch_b256_all_cores"] = cpu_generous(model, 3.0)
        vt, ct, st_ = cpu_t2_json(model, 5.0)   # the bolt's real unit of work: JSON tuple in -> JSON message out
        line["cpu_baseline"]["t2_json_tuples_per_s_all_cores"] = vt
        line["cpu_baseline"]["t2_sample"] = st_
    if not args.skip_latency and world == 1:
        line["t2_and_latency"] = t2_and_latency(sib, model, blob, prec, not args.no_guard)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _large_launch(large, roof, bpi, fpi):
    """roofline of the dominant kernel at the large launch size, against the same peak as roofline.frac"""
    if not large or "kernel_ms" not in large:
        return large
    per = fpi * 1e-12 if roof["bound"] == "tensor" else bpi * 1e-9
    ach = large["images_per_launch"] * per / (large["kernel_ms"] * 1e-3)
    return {**large, "achieved": ach, "unit": roof["unit"], "frac": ach / roof["peak"]}


def _poll_count(eng, timeout_us):
    n, res = eng.poll_raw(4096, timeout_us)
    return sum(res[i].n for i in range(n))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
