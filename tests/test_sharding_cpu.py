"""N > 1 host logic on the CPU: query sharding and the max/sum reduction of timings, world size 2,
gloo backend (the GPU box runs the same code over NCCL)."""
import os
import subprocess
import sys
import textwrap

import numpy as np

from ga_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shards_partition_the_queries():
    for n, w in ((4096, 1), (4096, 2), (4096, 8), (4000, 3), (64, 4), (1, 2), (0, 2)):
        parts = [sharding.shard_queries(n, w, r) for r in range(w)]
        allq = np.concatenate(parts)
        assert np.array_equal(np.sort(allq), np.arange(n))
        # whole maps per rank: no map is split over two ranks
        maps = [set(p // sharding.QUERIES_PER_MAP) for p in parts]
        for a in range(w):
            for b in range(a + 1, w):
                assert not (maps[a] & maps[b])
        sizes = [len(set(p // sharding.QUERIES_PER_MAP)) for p in parts]
        assert max(sizes) - min(sizes) <= 1
    assert sharding.map_seed(0) == 50000 and sharding.map_seed(63) == 50000 and sharding.map_seed(64) == 50001


def test_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(textwrap.dedent("""
        import os, sys, json
        sys.path.insert(0, %r)
        import torch, torch.distributed as dist
        from ga_b200 import sharding
        dist.init_process_group("gloo")
        r, w = dist.get_rank(), dist.get_world_size()
        q = sharding.shard_queries(4096, w, r)
        ms, units = sharding.reduce_timing(dist, "cpu", 10.0 * (r + 1), len(q))
        dist.barrier()
        if r == 0:
            print(json.dumps({"ms": ms, "units": units, "n0": len(q)}))
        dist.destroy_process_group()
    """ % os.path.join(ROOT, "guided-extended-hybrid-astar_b200")))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29613", str(script)],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res == {"ms": 20.0, "units": 4096.0, "n0": 2048}
