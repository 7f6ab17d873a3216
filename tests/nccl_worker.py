"""torchrun worker: every rank parses its slice of the reads, exchanges supermers over NCCL inside the library, and checks
its shard of the table against the oracle. Prints NCCL-PARITY-OK on rank 0 when all ranks agree."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mhm2_proxy_b200 import kcount, synth  # noqa: E402
from mhm2_proxy_b200.dist import share_unique_id, split_range  # noqa: E402
from oracle import harness  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = share_unique_id(rank, kcount.KmerCounter.comm_unique_id)
    com = synth.make_community(120000, n_genomes=3, seed=5)
    ok = True
    for k in (21, 55):
        rb, ro = synth.make_reads(9000, com, error_rate=0.01, seed=40 + k, n_rate=1e-3)
        cs, co, cd = synth.make_contigs(com, k, 30, seed=k, depth_max=6)
        ref = harness.run("port", k, rb, ro, num_shards=world, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
        r0, r1 = split_range(len(ro) - 1, rank, world)
        c0, c1 = split_range(len(cd), rank, world)
        with kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=local, num_partitions=8, batch_reads=2000) as kc:
            kc.comm_init(uid)
            kc.add_reads(rb[int(ro[r0]):int(ro[r1])], ro[r0:r1 + 1] - ro[r0])
            if c1 > c0:
                kc.add_contigs(cs[int(co[c0]):int(co[c1])], co[c0:c1 + 1] - co[c0], cd[c0:c1])
            kc.finish()
            keys, counts, left, right = kc.survivors()
            sel = ref.shard == rank
            good = (np.array_equal(keys, ref.keys[sel]) and np.array_equal(counts, ref.counts[sel])
                    and np.array_equal(left, ref.left[sel]) and np.array_equal(right, ref.right[sel]))
            st = kc.stats()
            print(f"rank {rank} k={k}: {len(keys)} survivors (oracle {int(sel.sum())}), sent {st['supermer_bytes_sent']} B, "
                  f"exchange {st['ms_exchange']:.2f} ms, ok={good}", flush=True)
            ok = ok and good
    # ragged case: unequal numbers of batches per rank, the last rank has nothing to parse and learns the partition count
    # from its peers during the exchange rounds
    k = 33
    rb, ro = synth.make_reads(7000, com, error_rate=0.01, seed=77, n_rate=1e-3)
    ref = harness.run("port", k, rb, ro, num_shards=world)
    n = len(ro) - 1
    cuts = [0] + [int(n * (1 - 0.5 ** (r + 1))) for r in range(world - 1)] + [n]
    cuts[-2] = n  # last rank: empty
    r0, r1 = cuts[rank], cuts[rank + 1]
    with kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=local, expected_instances=4_000_000,
                            batch_reads=1000) as kc:
        kc.comm_init(uid)
        if r1 > r0:
            kc.add_reads(rb[int(ro[r0]):int(ro[r1])], ro[r0:r1 + 1] - ro[r0])
        kc.finish()
        keys, counts, left, right = kc.survivors()
        sel = ref.shard == rank
        good = (np.array_equal(keys, ref.keys[sel]) and np.array_equal(counts, ref.counts[sel])
                and np.array_equal(left, ref.left[sel]) and np.array_equal(right, ref.right[sel]))
        print(f"rank {rank} ragged k={k}: reads {r1 - r0}, {len(keys)} survivors (oracle {int(sel.sum())}), ok={good}", flush=True)
        ok = ok and good
    # the reference's flow (kcount.cpp:93-96,148; kmer_dht.cpp:227-231): ranks add DIFFERENT numbers of batches, then a flush,
    # then host-side collectives (a reduction and a barrier on the bootstrap group), then the contig pass on some ranks only,
    # flush again, finish. Without kc_flush the rank with more batches would wait in an exchange round for a peer that
    # already sits in the barrier.
    k = 31
    rb, ro = synth.make_reads(9000, com, error_rate=0.01, seed=91, n_rate=1e-3)
    cs, co, cd = synth.make_contigs(com, k, 30, seed=92, depth_max=6)
    ref = harness.run("port", k, rb, ro, num_shards=world, ctg_seqs=cs, ctg_offsets=co, ctg_depths=cd)
    n = len(ro) - 1
    cuts = [0] + [int(n * (1 - 0.25 ** (r + 1))) for r in range(world - 1)] + [n]  # rank 0 parses 3/4 of the reads
    r0, r1 = cuts[rank], cuts[rank + 1]
    with kcount.KmerCounter(k, num_shards=world, shard_id=rank, device=local, num_partitions=16, batch_reads=500) as kc:
        kc.comm_init(uid)
        kc.add_reads(rb[int(ro[r0]):int(ro[r1])], ro[r0:r1 + 1] - ro[r0])  # 14 batches on rank 0, 4 on rank 1 (world 2)
        kc.flush()
        tot = torch.tensor([r1 - r0], device="cuda")
        dist.all_reduce(tot)  # a host-side collective between the passes, like reduce_one + barrier in count_kmers
        dist.barrier()
        assert int(tot.item()) == n
        if rank == 0:  # contigs only on one rank: the others still flush
            kc.add_contigs(cs, co, cd)
        kc.flush()
        dist.barrier()
        kc.finish()
        keys, counts, left, right = kc.survivors()
        sel = ref.shard == rank
        good = (np.array_equal(keys, ref.keys[sel]) and np.array_equal(counts, ref.counts[sel])
                and np.array_equal(left, ref.left[sel]) and np.array_equal(right, ref.right[sel]))
        good = good and kc.digest() == kcount.table_digest(ref.keys[sel], ref.counts[sel], ref.left[sel], ref.right[sel])
        st = kc.stats()
        print(f"rank {rank} unequal batches + flush k={k}: reads {r1 - r0}, rounds {st['exchange_rounds']}, {len(keys)} survivors "
              f"(oracle {int(sel.sum())}), recv {st['supermer_bytes_recv']} B, comm {st['ms_comm']:.2f} ms, ok={good}", flush=True)
        ok = ok and good
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0 and int(t.item()) == 1:
        print("NCCL-PARITY-OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
