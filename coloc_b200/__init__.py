"""coloc_b200 -- B200-native (sm_100a) front end of CoLoC: pyramid resize, FAST+NMS+orientation,
512-bit LATCH, brute-force Hamming 2-NN, behind the C ABI in include/coloc_b200.h.

This package is a thin ctypes binding over libcoloc_b200.so for tests and bench.py; the drop-in
host side for the reference is the C++ shim in coloc_b200/shim/coloc/ (see INTEGRATION.md).
There is no CPU fallback: using `coloc_b200.api` without the built library raises.
"""
from .api import (Detector, Matcher, ShardedMatcher, FrameStream, FeatureMap, is_lab, ClbError, KP_DTYPE, PARTIAL_DTYPE, PAIR_DTYPE, OBS_DTYPE, UNDEFINED_FEAT, lib, lib_path, build,  # noqa: F401
                  launch_count, set_timing, timeline, last_kernel_ms, set_param, get_param, microbench_pipes, microbench_mma, match_2nn_noalloc, PinnedArray, pinned_like)
