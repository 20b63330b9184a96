"""gappa_b200 — B200-native pairwise KR-distance path of `gappa analyze krd`.

Only what the hot path needs lives here: csrc/ (CUDA kernels + C ABI -> libkrd_b200.so),
krd.py (ctypes binding + host mirror of genesis' earth_movers_distance operator),
synth.py (synthetic inputs of BASELINE.json's shapes), shard.py (multi-GPU tile sharding).
"""
from .krd import KrdPlan, KrdError, earth_movers_distance, lib, LIB_PATH, EXPORTS, measure_fp64_peak  # noqa: F401
