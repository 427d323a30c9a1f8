"""Host-side sharding rules for the multi-GPU path (no collective on the data path).

The reference scales by `setBolt("inference-bolt", bolt, INFERENCE_BOLT_PARAL).shuffleGrouping("kafka-spout")`
(MainTopology.java:27,62): each executor owns a full model replica (InferenceBolt.java:57-58) and tuples are dealt
to executors with no key affinity.  Here one executor-engine is pinned per GPU (task index mod GPU count) and the
synthetic spout deals tuples round-robin, which is what Storm's shuffle grouping converges to.
torch.distributed is used by bench.py only for the start barrier and the max-over-ranks of the elapsed time.
"""
from __future__ import annotations


def device_for_task(task_index: int, n_gpus: int) -> int:
    """executor task index -> CUDA ordinal (sib_config.device < 0 applies the same rule natively)."""
    if n_gpus <= 0:
        raise ValueError("no GPUs")
    return task_index % n_gpus


def deal_round_robin(n_tuples: int, world: int, rank: int) -> range:
    """Indices of the tuples executor `rank` of `world` receives under shuffle grouping (round-robin deal)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    return range(rank, n_tuples, world)


def aggregate(local_units: int, local_seconds: float, dist=None):
    """Whole-job throughput = units all ranks processed / max-over-ranks elapsed. Returns (units, seconds)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_units, local_seconds
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([local_seconds], dtype=torch.float64, device=dev)
    u = torch.tensor([float(local_units)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return int(u.item()), float(t.item())
