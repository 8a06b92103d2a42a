"""viso_b200 — B200-native bundle-adjustment backend behind viso's `bundle_adjuster.h` surface.

Only what the hot path needs lives here: csrc/ (CUDA kernels + the C ABI of
include/viso_ba.h) and api.py (ctypes binding + the Python mirror of the reference class).
"""
from .api import (  # noqa: F401
    BundleAdjuster, Config, IterStats, KernelTime, Summary, SynthSpec, VisoBaError, LIB_PATH, SYNTH_LIB_PATH, SYMBOLS, SYNTH_SYMBOLS,
    batch_problems, default_config, load_library, load_synth_library, nccl_unique_id, shard_point_ranges, synth_config, synth_preset, synth_problem,
)
from .stream import optimize_stream  # noqa: F401,E402
