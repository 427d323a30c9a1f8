"""world_size-2 gloo test of the N>1 host logic: tuple dealing covers every tuple exactly once, executor->GPU
mapping, and the max-over-ranks / sum-of-units aggregation bench.py uses (MainTopology.java:27,62)."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_deal_and_device_mapping(sib):
    from importlib import import_module
    sh = import_module("sib_b200.sharding")
    for world in (1, 2, 4, 8):
        seen = sorted(i for r in range(world) for i in sh.deal_round_robin(1003, world, r))
        assert seen == list(range(1003))
        sizes = [len(sh.deal_round_robin(1003, world, r)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
    assert [sh.device_for_task(t, 8) for t in range(10)] == [0, 1, 2, 3, 4, 5, 6, 7, 0, 1]
    assert [sh.device_for_task(t, 4) for t in (0, 1, 2, 3)] == [0, 1, 2, 3]  # INFERENCE_BOLT_PARAL = 4


WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "oracle"))
    import numpy as np, torch.distributed as dist
    from importlib import import_module
    import sib_b200 as sib
    sh = import_module("sib_b200.sharding")
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    m = sib.modelspec.make_preset("mnist-lenet5")
    x = sib.modelspec.synthetic_images(m, 0, 37)
    mine = list(sh.deal_round_robin(len(x), world, rank))
    # each "executor" runs its share through the CPU oracle (there is no GPU here); results are gathered and
    # must equal the unsharded run: sharding changes nothing about any single tuple
    import oracle
    om = oracle.Model(m.to_blob())
    _, pr = om.forward(x[mine])
    out = [None] * world
    dist.all_gather_object(out, (mine, pr))
    units, secs = sh.aggregate(len(mine), 0.5 + rank, dist)
    if rank == 0:
        full = np.empty((len(x), 10), np.float32)
        for idx, p in out:
            full[idx] = p
        _, want = om.forward(x)
        assert full.tobytes() == want.tobytes()
        assert units == len(x) and secs == 0.5 + (world - 1), (units, secs)
        print("GLOO_OK")
    dist.destroy_process_group()
""")


def test_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % (ROOT, ROOT))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "GLOO_OK" in r.stdout
