"""mhm2_proxy_b200 — B200-native kcount (k-mer analysis) hot path of MetaHipMer2 behind a C ABI."""
