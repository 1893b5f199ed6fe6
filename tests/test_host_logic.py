"""CPU tier: host-side logic -- the synthetic generator, the shard split used by bench.py / the multi-device driver, and a
world_size-2 gloo run of the per-rank bookkeeping (one process per GPU; no data-path collective exists to test)."""
import os
import subprocess
import sys

import numpy as np

from teqp_b200.workloads import WORKLOADS, shard_range, synth_points, workload_points

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_generator_is_deterministic_and_offsets_compose():
    T, rho, x = workload_points("C3", 5000)
    T2, rho2, x2 = workload_points("C3", 5000)
    assert np.array_equal(T, T2) and np.array_equal(x, x2)
    Ta, ra, xa = workload_points("C3", 2000, k0=0)
    Tb, rb, xb = workload_points("C3", 3000, k0=2000)
    assert np.array_equal(np.concatenate([Ta, Tb]), T) and np.array_equal(np.concatenate([xa, xb], axis=1), x)
    w = WORKLOADS["C3"]
    assert T.min() >= w["T"][0] and T.max() <= w["T"][1] and rho.min() >= w["rho"][0] and rho.max() <= w["rho"][1]
    np.testing.assert_allclose(x.sum(axis=0), 1.0, rtol=0, atol=2e-15)
    assert abs(x[0].mean() - w["z0"][0]) < 0.02


def test_shard_range_covers_batch():
    for npts in (0, 1, 7, 1000, 10**8 + 3):
        for world in (1, 2, 4, 8):
            spans = [shard_range(npts, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == npts
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [b - a for a, b in spans]
            assert max(sizes[:-1] or [0]) - min(sizes[:-1] or [0]) == 0      # even split, remainder to the last


def test_two_ranks_gloo():
    """torchrun-style world of 2 on CPU (gloo): each rank generates its shard of the stream, the shards tile the batch, and the
    max-over-ranks timing reduction bench.py uses works."""
    code = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %r)
from teqp_b200.workloads import shard_range, workload_points
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 10001
a, b = shard_range(n, rank, world)
T, rho, x = workload_points("C3", b - a, k0=a)
parts = [None] * world
dist.all_gather_object(parts, (a, b, T, x))
t = torch.tensor([1.0 + rank], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    Tfull, _, xfull = workload_points("C3", n)
    assert np.array_equal(np.concatenate([p[2] for p in parts]), Tfull)
    assert np.array_equal(np.concatenate([p[3] for p in parts], axis=1), xfull)
    assert t.item() == world
    print("GLOO_OK")
dist.destroy_process_group()
''' % ROOT
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    import tempfile
    with tempfile.NamedTemporaryFile("w", suffix=".py", delete=False) as f:
        f.write(code)
        path = f.name
    try:
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                            "--master-port", "29531", path], env=env, capture_output=True, text=True, timeout=300)
    finally:
        os.unlink(path)
    assert r.returncode == 0 and "GLOO_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
