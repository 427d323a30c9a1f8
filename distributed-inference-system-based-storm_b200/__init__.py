"""B200-native InferenceBolt hot path (drop-in for HyoJong-Moon/Distributed-Inference-System-based-Storm).

The product is libsib.so (csrc/, C ABI in include/sib.h): pinned-host micro-batcher + native JSON codec + sm_100a
CUDA kernels.  This Python layer only binds that ABI for tests and benchmarks; it contains no compute and no
fallback: importing `native` raises if the CUDA library has not been built.

The directory name carries hyphens (the task's naming), so import it through the repo-root shim `sib_b200`.
"""
from . import modelspec  # noqa: F401
from . import native  # noqa: F401
from .bolt import InferenceBolt, OutputCollector, Tuple  # noqa: F401


def smoke_precisions():
    """(precision, max-abs probability tolerance) pairs the smoke run covers."""
    return [(native.PREC_FP32, 1e-5), (native.PREC_FP16, 1e-2)]


def default_bench_precision() -> str:
    """Precision bench.py uses when none is given: the fastest path that is parity-green."""
    return "fp16"
