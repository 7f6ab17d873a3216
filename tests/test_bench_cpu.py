"""bench.py without a GPU: the workload generators of BASELINE.json's configs and the contract of the reference arm."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.mark.parametrize("name", ["C2", "C4", "C5"])
def test_workload_generators_are_deterministic_and_shaped(name):
    import bench
    a = bench.Workload(name, 1, 0, reads=20000, pinned=False)
    b = bench.Workload(name, 1, 0, reads=20000, pinned=False)
    assert np.array_equal(a.tb.numpy(), b.tb.numpy()) and np.array_equal(a.to.numpy(), b.to.numpy())
    assert a.tb.numel() == 20000 * bench.READ_LEN and int(a.to[-1]) == a.tb.numel()
    codes = a.tb.numpy() & 7
    assert codes.max() <= 4  # PackedRead bytes: A0 C1 G2 T3 N4 in the low three bits (packed_reads.cpp:73-109)
    other = bench.Workload(name, 2, 1, reads=20000, pinned=False)  # another rank draws other reads from the same community
    assert not np.array_equal(a.tb.numpy(), other.tb.numpy())
    cfg = a.config(1)
    assert cfg["workload"].startswith(name + ":") and cfg["k"] == list(bench.WORKLOADS[name]["ks"])
    if name == "C4":
        cs, co, cd = a.ctg
        assert len(cd) == len(co) - 1 > 0 and int(co[-1]) == len(cs)
        assert cd.min() >= 2 and cd.max() <= 200  # depths 2..200
        assert a.ctg_instances(99) > 0
        assert set(np.unique(np.frombuffer(cs.tobytes(), dtype=np.uint8))) <= set(b"ACGTNacgtn")
    if name == "C5":
        reads = a.tb.numpy().reshape(20000, bench.READ_LEN) & 7
        assert ((reads == 0).sum(axis=1) >= bench.READ_LEN - 1).any()  # poly-A reads (one variable base): the saturating-counter case


def test_c3_community_grows_with_the_number_of_gpus():
    import bench
    g1, n1 = bench.community_size("C3", 1)
    g8, n8 = bench.community_size("C3", 8)
    assert g8 == 8 * g1 == 1_500_000_000 and n8 == 300  # 1.5 Gbp / 300 genomes over 8 GPUs (BASELINE.json configs[2])
    assert bench.community_size("C2", 8) == bench.community_size("C2", 1)
    assert bench.default_config(8) == "C3" and bench.default_config(1) == "C2"


def test_b_alg_follows_the_survey_formula():
    import bench
    assert abs(bench.b_alg(21) - (150 / 128 + 8 + 12)) < 1e-12
    assert abs(bench.b_alg(33) - (150 / 116 + 16 + 12)) < 1e-12
    assert abs(bench.b_alg(99) - (150 / 50 + 32 + 12)) < 1e-12


def test_reference_arm_prints_the_contract_line(built):
    """`bench.py --impl reference`: the reference's own CPU kcount on a bounded sample, one JSON line with the keys the
    driver reads (impl, metric, unit, value, cpu_baseline, e2e with zero copy bytes)."""
    from oracle import harness
    if not harness.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "C2", "--ref-reads", "4000",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "kcount k-mers inserted/sec" and d["unit"] == "kmers/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and "4000 reads" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "kmers/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["k"] == [21, 33] and d["config"]["workload"].startswith("C2:")


def test_traffic_json_is_what_the_committed_ncu_pages_say(tmp_path):
    """profiles/traffic.json (read back into roofline.traffic) must be regenerable from the committed ncu raw pages: no
    hand-edited constants (evidence hygiene)."""
    import shutil
    prof = os.path.join(ROOT, "profiles")
    cur = json.load(open(os.path.join(prof, "traffic.json")))
    tag = cur["_comment"].split("capture ")[1].split(" ")[0]
    raws = [os.path.join(prof, f"{tag}_count_raw.csv"), os.path.join(prof, f"{tag}_parse_raw.csv")]
    assert all(os.path.exists(r) for r in raws), raws
    # run the generator on a scratch copy of the tree layout (it writes <root>/profiles/traffic.json next to tools/)
    scratch = tmp_path / "repo"
    (scratch / "tools").mkdir(parents=True)
    (scratch / "profiles").mkdir()
    shutil.copy(os.path.join(ROOT, "tools", "make_traffic.py"), scratch / "tools" / "make_traffic.py")
    out = subprocess.run([sys.executable, str(scratch / "tools" / "make_traffic.py"), tag] + raws, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    new = json.load(open(scratch / "profiles" / "traffic.json"))
    for k in ("count_kernel", "parse_kernel", "fill_kernel"):
        assert new[k] == cur[k], (k, new[k], cur[k])
