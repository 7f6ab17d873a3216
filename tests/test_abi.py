"""CPU tests of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/kcount_b200.h declares. No compute calls here (no GPU in this container)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(built):
    from mhm2_proxy_b200 import kcount
    lib = kcount.load_library()
    hdr = open(os.path.join(ROOT, "include", "kcount_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(kc_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 18
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in kcount_b200.h but not exported"
    assert declared == set(kcount.EXPORTS)


def test_struct_layouts_match_header(built):
    from mhm2_proxy_b200 import kcount
    lib = kcount.load_library()
    cfg = kcount.KcConfig()
    lib.kc_config_default(ctypes.byref(cfg), 31)
    # defaults are the reference's constants (CMakeDefinitions.txt:46,60; options.hpp:96)
    assert (cfg.kmer_len, cfg.qual_cutoff, cfg.dmin_thres, cfg.num_shards) == (31, 20, 2, 1)
    assert abs(cfg.dyn_min_depth - 0.9) < 1e-15
    a, b = ctypes.c_uint32(), ctypes.c_uint32()
    lib.kc_struct_sizes(ctypes.byref(a), ctypes.byref(b))
    assert ctypes.sizeof(kcount.KcConfig) == a.value == 80
    assert ctypes.sizeof(kcount.KcStats) == b.value == (17 + 9 + 4 + 4) * 8


def test_no_gpu_means_loud_failure(built):
    """Without a CUDA device the product path must fail with an error, never fall back to the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mhm2_proxy_b200 import kcount
    with pytest.raises(kcount.KcountError):
        kcount.KmerCounter(31)


def test_dropin_binding_builds_against_reference_headers(built):
    """integration/kcount_b200.cpp (the reference-side binding) compiles together with the reference's unmodified kcount.cpp,
    kmer_dht.cpp and kmer.cpp, and the result links against libkcount_b200.so (built only where /root/reference exists)."""
    from oracle import harness
    if not harness.have_dropin():
        pytest.skip("drop-in build not present (no /root/reference)")
    lib = ctypes.CDLL(harness.DROPIN_SO)
    for name in ("odrop_run", "odrop_num_survivors", "odrop_copy_survivors", "odrop_free"):
        assert hasattr(lib, name)


def test_product_never_imports_oracle():
    for pkg in (os.path.join(ROOT, "mhm2_proxy_b200"), os.path.join(ROOT, "integration")):
      for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".inl")):
                src = open(os.path.join(dirpath, f)).read()
                for pat in ("import oracle", "from oracle", "oracle/", "oracle.harness", "libkcount_oracle", "libkcount_ref", "ko_run", "oref_"):
                    assert pat not in src, f"{f} references the oracle ({pat})"
