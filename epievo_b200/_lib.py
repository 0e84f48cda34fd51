"""ctypes binding of libepievo_b200.so (the C-ABI in include/epievo_b200.h).

There is no CPU fallback: if the library has not been built, or no CUDA device is
usable, loading / context creation raises."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# EPIEVO_B200_LIB: A/B-test another build of the same library (tools/build_variants.sh)
LIB_PATH = os.environ.get("EPIEVO_B200_LIB") or os.path.join(HERE, "libepievo_b200.so")

dp = C.POINTER(C.c_double)
u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
ctx_p = C.c_void_p
u64 = C.c_uint64

# name -> (restype, argtypes); mirrors include/epievo_b200.h one to one
SIGNATURES = {
    "epv_create": (C.c_int, [C.POINTER(ctx_p), C.c_int]),
    "epv_destroy": (None, [ctx_p]),
    "epv_last_error": (C.c_char_p, [ctx_p]),
    "epv_set_stream": (C.c_int, [ctx_p, C.c_void_p]),
    "epv_synchronize": (C.c_int, [ctx_p]),
    "epv_set_tree": (C.c_int, [ctx_p, u64, u64p, u64p, dp]),
    "epv_set_model": (C.c_int, [ctx_p, dp, dp]),
    "epv_set_shard": (C.c_int, [ctx_p, u64, u64, u64, u64, u64]),
    "epv_upload_paths": (C.c_int, [ctx_p, u64, u64, u8p, u64p, dp, dp]),
    "epv_paths_size": (C.c_int, [ctx_p, u64p, u64p, u64p]),
    "epv_download_paths": (C.c_int, [ctx_p, u8p, u64p, dp, dp]),
    "epv_upload_paths_compact": (C.c_int, [ctx_p, u64, u64, u16p, dp, u64, dp]),
    "epv_download_paths_compact": (C.c_int, [ctx_p, u16p, dp, u64, u64p, dp]),
    "epv_reset": (C.c_int, [ctx_p]),
    "epv_sweep": (C.c_int, [ctx_p, u64, u64, u64p]),
    "epv_run_mcmc": (C.c_int, [ctx_p, u64, u64, u64, u64, dp, dp, dp]),
    "epv_run_mcmc_raw": (C.c_int, [ctx_p, u64, u64, u64, u64, dp, dp, u64p]),
    "epv_suffstats": (C.c_int, [ctx_p, dp, dp]),
    "epv_llh_grad": (C.c_int, [ctx_p, dp, dp, dp]),
    "epv_scale_jump_times": (C.c_int, [ctx_p, dp]),
    "epv_mcem_iteration": (C.c_int, [ctx_p, u64, u64, u64, u64, C.c_int, C.c_double, dp, dp, dp, dp,
                                     dp, dp]),
    "epv_simulate": (C.c_int, [ctx_p, u64, u8p, u64, u64p]),
    "epv_simulate_fetch": (C.c_int, [ctx_p, u8p, u64p, dp]),
    "epv_simulate_adopt": (C.c_int, [ctx_p]),
    "epv_pass": (C.c_int, [ctx_p, C.c_int, u64, u64]),
    "epv_accept_count": (C.c_int, [ctx_p, C.c_int, u64p]),
    "epv_halo_record_bytes": (C.c_int, [ctx_p, u64p]),
    "epv_halo_pack": (C.c_int, [ctx_p, C.c_int, C.c_void_p, C.c_void_p]),
    "epv_halo_unpack": (C.c_int, [ctx_p, C.c_int, C.c_void_p, C.c_void_p]),
    "epv_halo_capacity": (C.c_int, [ctx_p, u64p, u64p]),
    "epv_halo_set_capacity": (C.c_int, [ctx_p, u64]),
    "epv_halo_need_async": (C.c_int, [ctx_p, C.c_void_p, C.c_void_p]),
    "epv_set_side_stream": (C.c_int, [ctx_p, C.c_void_p]),
    "epv_pass_begin": (C.c_int, [ctx_p, C.c_int, u64, u64, C.c_void_p, C.c_void_p]),
    "epv_pass_end": (C.c_int, [ctx_p, C.c_int, C.c_void_p, C.c_void_p]),
    "epv_stats_words": (C.c_int, [ctx_p, u64p]),
    "epv_stats_begin": (C.c_int, [ctx_p]),
    "epv_stats_accumulate": (C.c_int, [ctx_p]),
    "epv_stats_export": (C.c_int, [ctx_p, C.c_void_p]),
    "epv_stats_add": (C.c_int, [ctx_p, C.c_void_p]),
    "epv_stats_fold": (C.c_int, [ctx_p, C.c_void_p, dp, dp]),
    "epv_suffstats_static": (C.c_int, [ctx_p, dp, dp]),
    "epv_static_limits": (C.c_int, [u64p, u64p]),
    "epv_suffstats_async": (C.c_int, [ctx_p]),
    "epv_suffstats_fetch": (C.c_int, [ctx_p, dp, dp]),
    "epv_multi_create": (C.c_int, [C.POINTER(ctx_p), C.c_int, C.POINTER(C.c_int)]),
    "epv_multi_destroy": (None, [ctx_p]),
    "epv_multi_last_error": (C.c_char_p, [ctx_p]),
    "epv_multi_shards": (C.c_int, [ctx_p]),
    "epv_multi_context": (ctx_p, [ctx_p, C.c_int]),
    "epv_multi_partition": (C.c_int, [u64, C.c_int, u64p, u64p]),
    "epv_multi_set_tree": (C.c_int, [ctx_p, u64, u64p, u64p, dp]),
    "epv_multi_set_model": (C.c_int, [ctx_p, dp, dp]),
    "epv_multi_upload_paths": (C.c_int, [ctx_p, u64, u64, u8p, u64p, dp, dp]),
    "epv_multi_paths_size": (C.c_int, [ctx_p, u64p, u64p, u64p]),
    "epv_multi_download_paths": (C.c_int, [ctx_p, u8p, u64p, dp, dp]),
    "epv_multi_reset": (C.c_int, [ctx_p]),
    "epv_multi_sweep": (C.c_int, [ctx_p, u64, u64, u64p]),
    "epv_multi_run_mcmc": (C.c_int, [ctx_p, u64, u64, u64, u64, dp, dp, dp]),
    "epv_multi_suffstats": (C.c_int, [ctx_p, dp, dp]),
    "epv_multi_scale_jump_times": (C.c_int, [ctx_p, dp]),
    "epv_multi_mcem_iteration": (C.c_int, [ctx_p, u64, u64, u64, u64, C.c_int, C.c_double, dp, dp, dp, dp,
                                           dp, dp]),
    "epv_multi_launch_count": (u64, [ctx_p]),
    "epv_site_debug": (C.c_int, [ctx_p, u64, u64, u64p, u64p, dp, dp, u64p, u64p, dp, dp, dp,
                                 dp, dp]),
    "epv_get_tri_llh": (C.c_int, [ctx_p, dp]),
    "epv_end_cond_sample": (C.c_int, [ctx_p, C.c_double, C.c_double, C.c_int, C.c_int,
                                      C.c_double, u64, u64, u64, u32p, dp]),
    "epv_model_from_params": (None, [C.c_double] * 4 + [C.c_int] + [dp] * 4),
    "epv_model_rebuild": (None, [dp] * 4),
    "epv_rate_scaling_factor": (C.c_double, [dp]),
    "epv_estimate_rates": (C.c_double, [u64, dp, dp, C.c_double, dp, dp, dp]),
    "epv_estimate_rates_and_branches": (C.c_double, [u64, dp, dp, C.c_double, dp, dp, dp, dp]),
    "epv_set_one_change": (None, [dp, u64, dp]),
    "epv_indep_suffstats": (C.c_int, [ctx_p, dp, dp]),
    "epv_indep_expectation": (C.c_int, [ctx_p, dp, dp, dp]),
    "epv_indep_update_paths": (C.c_int, [ctx_p, dp, u64, u64]),
    "epv_indep_tables": (None, [C.c_double, C.c_double, C.c_double, dp, dp, dp, dp, dp]),
    "epv_indep_estimate_rates": (None, [u64, dp, dp, dp]),
    "epv_indep_estimate_rates_and_branches": (None, [u64, dp, dp, dp, dp]),
    "epv_pathfile_read": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_void_p), u64p, u64p, u64p]),
    "epv_pathfile_get": (C.c_int, [C.c_void_p, u8p, u64p, dp, dp]),
    "epv_pathfile_node_name": (C.c_char_p, [C.c_void_p, u64]),
    "epv_pathfile_free": (None, [C.c_void_p]),
    "epv_pathfile_write": (C.c_int, [C.c_char_p, u64, u64, C.POINTER(C.c_char_p), u8p, u64p, dp, dp, C.c_int]),
    "epv_pathfile_write_binary": (C.c_int, [C.c_char_p, u64, u64, C.POINTER(C.c_char_p), u8p, u64p, dp, dp, C.c_int]),
    "epv_io_last_error": (C.c_char_p, []),
    "epv_launch_count": (u64, [ctx_p]),
    "epv_recovery_count": (u64, [ctx_p]),
    "epv_mh_kernel_time": (C.c_int, [ctx_p, C.c_int, dp, u64p]),
    "epv_rebuild_kernel_time": (C.c_int, [ctx_p, dp]),
    "epv_enable_kernel_timing": (C.c_int, [ctx_p, C.c_int]),
    "epv_jump_counts": (C.c_int, [ctx_p, u64p]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "epievo_b200: %s is missing -- build it with `python -m epievo_b200.build` "
                "(or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
