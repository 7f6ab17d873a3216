"""kc_table_digest_static — the allocation-free entry point (no counterpart in the reference):
its scratch is a file-scope __device__ array, so it can be called where cudaMalloc is forbidden. Checked through
the C ABI against the existing entry point kc_table_digest, bit-exact (the digest is integer arithmetic), on calls in a row
that share the scratch array. Collected last on purpose."""
import numpy as np
import pytest

from mhm2_proxy_b200 import synth

pytestmark = pytest.mark.gpu


def _static_digest(kcount, torch, kc, out):
    d_keys, d_counts, d_left, d_right = kc.survivors_device()
    n = kc.num_survivors()
    kcount.table_digest_static(0, d_keys or 0, d_counts or 0, d_left or 0, d_right or 0, n, kc.n_longs, out.data_ptr())
    torch.cuda.synchronize()
    return tuple(int(x) & 0xFFFFFFFFFFFFFFFF for x in out.cpu().tolist())


def test_static_digest_matches_table_digest_on_calls_in_a_row(built):
    import torch
    from mhm2_proxy_b200 import kcount
    com = synth.make_community(400000, n_genomes=3, seed=71)
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda")  # the caller's output buffer (allocated by the caller)
    seen = []
    # two different tables, then the first shape again: every call must clear what the previous one left in the scratch.
    # 60 k reads at k=21 give ~350 k survivors = several CTAs' partial results; k=55 has two key words per row.
    for k, n_reads, seed in ((21, 60000, 1), (55, 9000, 2), (21, 60000, 1)):
        rb, ro = synth.make_reads(n_reads, com, error_rate=0.01, seed=seed)
        with kcount.KmerCounter(k) as kc:
            kc.add_reads(rb, ro)
            kc.finish()
            want = kc.digest()  # the existing entry point (allocates its accumulator)
            got = _static_digest(kcount, torch, kc, out)
            again = _static_digest(kcount, torch, kc, out)  # twice in a row on the same table
            assert want[0] == kc.num_survivors() > 0
            assert got == want, f"k={k}: static digest {got} != kc_table_digest {want}"
            assert again == want
            seen.append(want)
    assert seen[0] == seen[2] and seen[0] != seen[1]
    assert seen[0][0] > 65536, "the first table must span more than one CTA's rows"


def test_static_digest_empty_table_and_limits(built):
    import torch
    from mhm2_proxy_b200 import kcount
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    kcount.table_digest_static(0, 0, 0, 0, 0, 0, 1, out.data_ptr())
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0, 0, 0, 0]
    dummy = torch.zeros(16, dtype=torch.int64, device="cuda")
    with pytest.raises(kcount.KcountError, match="exceed the static scratch's limit"):
        kcount.table_digest_static(0, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                   kcount.DIGEST_STATIC_MAX_ROWS + 1, 1, out.data_ptr())
    with pytest.raises(kcount.KcountError, match="n_longs"):
        kcount.table_digest_static(0, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 4, 5, out.data_ptr())
