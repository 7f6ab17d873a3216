"""CPU (gloo, world_size 2) tests of the N>1 host-side plumbing: read splitting, unique-id bootstrap, and the sharding
contract itself checked with the oracle (shard r holds exactly the survivors whose owner is r)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, os.environ["KC_ROOT"])
from mhm2_proxy_b200 import synth
from mhm2_proxy_b200.dist import split_range, share_unique_id
from oracle import harness
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
uid = share_unique_id(rank, lambda: bytes(range(128)))
assert uid == bytes(range(128))
k = 31
com = synth.make_community(40000, n_genomes=2, seed=3)
rb, ro = synth.make_reads(1500, com, error_rate=0.01, seed=4)
n = len(ro) - 1
r0, r1 = split_range(n, rank, world)
sizes = [None] * world
dist.all_gather_object(sizes, (r0, r1))
assert sizes[0][0] == 0 and sizes[-1][1] == n and all(sizes[i][1] == sizes[i + 1][0] for i in range(world - 1))
# the sharding contract: union over shards of (survivors with owner == shard) == the single-table result
full = harness.run("port", k, rb, ro, num_shards=world)
mine = full.shard == rank
counts = [None] * world
dist.all_gather_object(counts, int(mine.sum()))
assert sum(counts) == len(full.keys)
one = harness.run("port", k, rb, ro, num_shards=1)
assert np.array_equal(one.keys, full.keys) and np.array_equal(one.counts, full.counts)
# instances are conserved by the split
inst = [None] * world
dist.all_gather_object(inst, synth.count_instances(ro[r0:r1 + 1], k))
assert sum(inst) == synth.count_instances(ro, k)
dist.barrier()
if rank == 0:
    print("GLOO-OK")
dist.destroy_process_group()
'''


def test_two_rank_plumbing_gloo(built, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, KC_ROOT=ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29511", str(script)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-3000:]
    assert "GLOO-OK" in out.stdout


def test_split_range_partitions_everything():
    from mhm2_proxy_b200.dist import split_range
    for n in (0, 1, 7, 100, 13333334):
        for w in (1, 2, 3, 8):
            edges = [split_range(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            assert max(e[1] - e[0] for e in edges) - min(e[1] - e[0] for e in edges) <= 1
