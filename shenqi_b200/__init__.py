"""shenqi_b200: B200-native (sm_100a) short-range TreePM gravity for MP-Gadget/shenqi.

Python is only the test / benchmark harness around libshenqi_b200.so (hand-written CUDA behind the C ABI
of include/shenqi_b200.h).  There is no CPU fallback: importing works without a GPU (so the build can be
checked), creating a Context does not.
"""
from .api import (Context, Params, WalkStats, TreeInfo, Timings, ShenqiError, load_library, library_path,
                  NODE_DTYPE, PARTICLE_DTYPE, TimebinParams, pushdown_bin, Domain, LetInfo, domain_from_histogram, domain_assign_balanced,
                  comm_unique_id, comm_loopback_id, bigfile_block_info, bigfile_read_block, bigfile_read_attr, TOPNODE_DTYPE, TOPLEAF_DTYPE, TASK_DTYPE)

__all__ = ["Context", "Params", "WalkStats", "TreeInfo", "Timings", "ShenqiError", "load_library", "library_path",
           "NODE_DTYPE", "PARTICLE_DTYPE", "TimebinParams", "pushdown_bin", "Domain", "LetInfo", "domain_from_histogram", "domain_assign_balanced",
           "comm_unique_id", "comm_loopback_id", "bigfile_block_info", "bigfile_read_block", "bigfile_read_attr", "TOPNODE_DTYPE", "TOPLEAF_DTYPE", "TASK_DTYPE"]
