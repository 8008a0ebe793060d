"""ctypes declarations of the libfgs C ABI (include/fgs.h, include/fgs_types.h).

This module only loads the shared library and mirrors the C structs; it contains no arithmetic.
Loading fails loudly when libfgs.so has not been built (run ``python -c 'import __graft_entry__ as g; g.build()'``):
there is no Python or CPU substitute for the CUDA path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfgs.so")

FGS_MAX_SCALARS = 8
FGS_MAX_DEVICES = 16
FGS_MAX_SCORE_RES = 8
FGS_RES_CPU, FGS_RES_MEM, FGS_RES_EPH, FGS_RES_SCALAR0, FGS_RES_UNKNOWN = 0, 1, 2, 3, -1
FGS_ANNOT_NONE = -(2 ** 31)
FGS_PL_FLEXGPU, FGS_PL_FIT, FGS_PL_ALLOCATABLE = 1, 2, 4
FGS_DIAG_FLEX0, FGS_DIAG_FLEX_PANIC, FGS_DIAG_FIT0, FGS_DIAG_NOT_A_NODE, FGS_DIAG_SLOTS = 0, 5, 6, 30, 32
(FGS_PREEMPT_CANDIDATE, FGS_PREEMPT_NO_VICTIMS, FGS_PREEMPT_STILL_UNFIT, FGS_PREEMPT_QUOTA, FGS_PREEMPT_ERROR, FGS_PREEMPT_NOT_A_NODE,
 FGS_PREEMPT_PANIC) = range(7)
FGS_CODE_SUCCESS, FGS_CODE_ERROR, FGS_CODE_UNSCHEDULABLE, FGS_CODE_UNRESOLVABLE, FGS_CODE_WAIT = 0, 1, 2, 3, 4
FGS_PODERR_SCORE_RANGE, FGS_PODERR_REF_PANIC = 1, 2
FGS_REDUCE_ROWS = 6
FGS_OK, FGS_E_INVALID, FGS_E_CUDA, FGS_E_RANGE, FGS_E_STATE, FGS_E_TOO_LARGE = 0, -1, -2, -3, -4, -5

p_i64 = C.POINTER(C.c_int64)
p_i32 = C.POINTER(C.c_int32)
p_u32 = C.POINTER(C.c_uint32)
p_u8 = C.POINTER(C.c_uint8)


class Podset(C.Structure):
    _fields_ = [
        ("n_pods", C.c_int64), ("n_scalars", C.c_int32), ("_pad0", C.c_int32),
        ("ns", p_i32), ("pg", p_i32), ("priority", p_i32), ("has_node_name", p_u8), ("index_annot", p_i32),
        ("ctr_off", p_i64), ("n_init", p_i32), ("overhead_present", p_u32), ("overhead", p_i64),
        ("n_ctrs", C.c_int64), ("c_req", p_i64), ("c_req_present", p_u32), ("c_lim_flex", p_i64), ("c_lim_present", p_u32),
        ("start_time", p_i64),
    ]


class Snapshot(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int64), ("n_scalars", C.c_int32), ("_pad0", C.c_int32),
        ("node_valid", p_u8), ("alloc", p_i64), ("alloc_scalar_present", p_u32), ("requested", p_i64), ("nonzero", p_i64),
        ("pod_off", p_i64), ("pods", Podset), ("generation", p_i64),
    ]


class Podgroups(C.Structure):
    _fields_ = [
        ("n_groups", C.c_int64), ("n_scalars", C.c_int32), ("_pad0", C.c_int32),
        ("exists", p_u8), ("ns", p_i32), ("min_member", p_i32), ("has_min_resources", p_u8), ("min_res", p_i64),
        ("min_res_present", p_u32), ("denied", p_u8), ("permitted", p_u8),
    ]


class Quotas(C.Structure):
    _fields_ = [
        ("n_ns", C.c_int64), ("n_scalars", C.c_int32), ("_pad0", C.c_int32),
        ("has_eq", p_u8), ("min", p_i64), ("max", p_i64), ("used", p_i64), ("used_scalar_present", p_u32),
    ]


class PluginCfg(C.Structure):
    _fields_ = [
        ("filter_mask", C.c_uint32), ("score_mask", C.c_uint32), ("filter_order", C.c_int32 * 2),
        ("weight_flex", C.c_int64), ("weight_fit", C.c_int64), ("weight_alloc", C.c_int64),
        ("fit_strategy", C.c_int32), ("fit_n_res", C.c_int32), ("fit_res", C.c_int32 * FGS_MAX_SCORE_RES),
        ("fit_res_w", C.c_int64 * FGS_MAX_SCORE_RES), ("fit_ignored_scalar_mask", C.c_uint32),
        ("alloc_mode", C.c_int32), ("alloc_n_res", C.c_int32), ("_pad0", C.c_int32),
        ("alloc_res", C.c_int32 * FGS_MAX_SCORE_RES), ("alloc_res_w", C.c_int64 * FGS_MAX_SCORE_RES),
    ]


class NodeTable(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int64), ("n_scalars", C.c_int32), ("n_devices", C.c_int32),
        ("alloc_mcpu", p_i64), ("alloc_mem", p_i64), ("alloc_eph", p_i64), ("alloc_pods", p_i64),
        ("req_mcpu", p_i64), ("req_mem", p_i64), ("req_eph", p_i64), ("num_pods", p_i64), ("nz_mcpu", p_i64), ("nz_mem", p_i64),
        ("alloc_scalar", p_i64), ("req_scalar", p_i64), ("alloc_scalar_present", p_u32),
        ("flex_gpu_sum", p_i64), ("flex_mem_sum", p_i64), ("used_mem", p_i64), ("monopoly_mask", p_u32), ("flags", p_u32),
    ]


class PodStats(C.Structure):
    _fields_ = [("or_bits", C.c_uint64 * (8 + FGS_MAX_SCALARS)), ("max_val", C.c_int64 * (8 + FGS_MAX_SCALARS)), ("present_or", C.c_uint32),
                ("flags", C.c_uint32)]


class PodTable(C.Structure):
    _fields_ = [
        ("n_pods", C.c_int64), ("n_scalars", C.c_int32), ("_pad0", C.c_int32),
        ("flex_gpu_limit", p_i64), ("flex_mem_limit", p_i64), ("flags", p_u32),
        ("req_mcpu", p_i64), ("req_mem", p_i64), ("req_eph", p_i64), ("req_scalar", p_i64), ("req_scalar_present", p_u32),
        ("nz_mcpu", p_i64), ("nz_mem", p_i64), ("ns", p_i32), ("pg", p_i32), ("priority", p_i32),
        ("acct_nz_mcpu", p_i64), ("acct_nz_mem", p_i64), ("index_annot", p_i32), ("uid", p_i64), ("stats", C.POINTER(PodStats)),
    ]


class Config(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("n_scalars", C.c_int32), ("node_index_base", C.c_int64), ("total_nodes", C.c_int64),
        ("select_seed", C.c_uint64), ("stream", C.c_void_p),
    ]


class Results(C.Structure):
    _fields_ = [
        ("feasible_count", p_i32), ("selected_node", p_i32), ("selected_score", p_i64), ("pod_err", p_u32),
        ("feasible_bitmap", p_u32), ("scores", p_i64), ("words_per_row", C.c_int64), ("score_row_stride", C.c_int64),
        ("next_start_node_index", C.c_int64), ("processed_nodes", p_i32),
    ]


class EvalOpts(C.Structure):
    _fields_ = [("emit_mask", C.c_uint32), ("emit_raw", C.c_uint32), ("skip_pass_b", C.c_uint32), ("sample", C.c_uint32),
                ("percentage_of_nodes_to_score", C.c_int32), ("_pad", C.c_int32), ("start_node_index", C.c_int64)]


class DeviceView(C.Structure):
    _fields_ = [
        ("reduce_max", C.c_void_p), ("reduce_sum", C.c_void_p), ("select_key", C.c_void_p), ("feasible_bitmap", C.c_void_p),
        ("scores", C.c_void_p), ("p_pad", C.c_int64), ("words_per_row", C.c_int64), ("score_row_stride", C.c_int64),
        ("stream", C.c_void_p),
    ]


FGS_PEER_MAX_RANKS = 8


class PeerGroup(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("buffers", C.c_void_p * FGS_PEER_MAX_RANKS), ("multicast", C.c_void_p),
                ("bytes", C.c_uint64)]


class Timings(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("reduce_ms", C.c_float), ("emit_ms", C.c_float), ("d2h_ms", C.c_float),
                ("kernel_launches", C.c_int64), ("reduce_int32", C.c_int32), ("emit_int32", C.c_int32),
                ("seq_ms", C.c_float), ("seq_ctas", C.c_int32), ("fused", C.c_int32), ("reduce_lookup", C.c_int32), ("small", C.c_int32), ("_pad", C.c_int32),
                ("node_patches", C.c_int64), ("node_patch_redos", C.c_int64)]


class SeqOpts(C.Structure):
    _fields_ = [("reserve_flex", C.c_int32), ("_pad", C.c_int32), ("quotas", C.c_void_p)]


class SeqResults(C.Structure):
    _fields_ = [("feasible_count", p_i32), ("selected_node", p_i32), ("selected_score", p_i64), ("pod_err", p_u32),
                ("device_index", p_i32), ("committed", p_u8), ("quota_code", p_u8), ("quota_used", p_i64),
                ("quota_used_present", p_u32)]


class SynthParams(C.Structure):
    _fields_ = [
        ("seed_nodes", C.c_uint64), ("seed_pods", C.c_uint64), ("n_nodes", C.c_int64), ("n_pending", C.c_int64),
        ("n_scalars", C.c_int32), ("n_devices", C.c_int32), ("max_existing_pods", C.c_int32),
        ("pm_fractional", C.c_int32), ("pm_whole", C.c_int32), ("pm_nogpu", C.c_int32), ("pm_degenerate", C.c_int32),
        ("pm_missing_req", C.c_int32), ("pm_init_ctr", C.c_int32), ("pm_overhead", C.c_int32), ("pm_two_ctrs", C.c_int32),
        ("pm_extra_scalar", C.c_int32),
        ("pm_quirk_nodes", C.c_int32), ("pm_nongpu_nodes", C.c_int32), ("pm_panic_nodes", C.c_int32), ("pm_invalid_nodes", C.c_int32),
        ("n_groups", C.c_int64),
        ("min_member_lo", C.c_int32), ("min_member_hi", C.c_int32), ("pm_groups_placed", C.c_int32), ("pm_groups_short", C.c_int32),
        ("pm_neg_left_nodes", C.c_int32),
        ("n_namespaces", C.c_int64), ("pm_pods_without_eq", C.c_int32), ("n_nominated", C.c_int32),
    ]


# every entry point include/fgs.h declares: name -> (restype, argtypes)
ABI = {
    "fgs_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "fgs_destroy": (None, [C.c_void_p]),
    "fgs_last_error": (C.c_char_p, [C.c_void_p]),
    "fgs_set_plugins": (C.c_int, [C.c_void_p, C.POINTER(PluginCfg)]),
    "fgs_packer_create": (C.c_int, [C.c_int32, C.POINTER(C.c_void_p)]),
    "fgs_packer_destroy": (None, [C.c_void_p]),
    "fgs_packer_error": (C.c_char_p, [C.c_void_p]),
    "fgs_packer_nodes": (C.c_int, [C.c_void_p, C.POINTER(Snapshot), C.c_int64, C.c_int64, C.POINTER(C.POINTER(NodeTable))]),
    "fgs_packer_pods": (C.c_int, [C.c_void_p, C.POINTER(Podset), C.c_int64, C.c_int64, C.POINTER(C.POINTER(PodTable))]),
    "fgs_upload_nodes": (C.c_int, [C.c_void_p, C.POINTER(NodeTable)]),
    "fgs_update_nodes": (C.c_int, [C.c_void_p, C.POINTER(NodeTable), p_i64]),
    "fgs_evaluate": (C.c_int, [C.c_void_p, C.POINTER(PodTable), C.POINTER(EvalOpts), C.POINTER(Results)]),
    "fgs_select_victims": (C.c_int, [C.c_void_p, C.POINTER(Snapshot), C.POINTER(Podset), C.c_int64, C.POINTER(Quotas), C.c_int32, p_u8, p_u8,
                                     p_i32, p_i32]),
    "fgs_set_nominated": (C.c_int, [C.c_void_p, C.POINTER(PodTable), p_i32, p_i64]),
    "fgs_submit": (C.c_int, [C.c_void_p, C.POINTER(PodTable), C.POINTER(EvalOpts)]),
    "fgs_collect": (C.c_int, [C.c_void_p, C.POINTER(Results)]),
    "fgs_packer_pin": (C.c_int, [C.c_void_p]),
    "fgs_packer_node_rows": (C.c_int, [C.c_void_p, C.POINTER(Snapshot), p_i64, C.c_int64, C.POINTER(C.POINTER(NodeTable))]),
    "fgs_sync_nodes": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Snapshot), C.c_int64, C.c_int64, p_i64]),
    "fgs_gang_postfilter": (C.c_int, [C.c_void_p, C.POINTER(Snapshot), C.POINTER(Podset), C.POINTER(Podgroups), p_u8, p_i32]),
    "fgs_score_masked": (C.c_int, [C.c_void_p, p_u32, C.c_int32, C.POINTER(EvalOpts), C.POINTER(Results)]),
    "fgs_resubmit": (C.c_int, [C.c_void_p, C.POINTER(EvalOpts)]),
    "fgs_set_resident_batches": (C.c_int, [C.c_void_p, C.c_int32]),
    "fgs_max_in_flight": (C.c_int, [C.c_void_p]),
    "fgs_stage_pods": (C.c_int, [C.c_void_p, C.POINTER(PodTable)]),
    "fgs_pass_reduce": (C.c_int, [C.c_void_p]),
    "fgs_fused_eligible": (C.c_int, [C.c_void_p, C.POINTER(EvalOpts)]),
    "fgs_pass_fused": (C.c_int, [C.c_void_p, C.POINTER(EvalOpts)]),
    "fgs_pass_fused_prepass": (C.c_int, [C.c_void_p, C.POINTER(EvalOpts)]),
    "fgs_pass_fused_main": (C.c_int, [C.c_void_p, C.POINTER(EvalOpts)]),
    "fgs_pass_emit": (C.c_int, [C.c_void_p, C.POINTER(EvalOpts)]),
    "fgs_fetch_results": (C.c_int, [C.c_void_p, C.POINTER(Results)]),
    "fgs_get_device_view": (C.c_int, [C.c_void_p, C.POINTER(DeviceView)]),
    "fgs_peer_bytes": (C.c_uint64, [C.c_int64]),
    "fgs_peer_attach": (C.c_int, [C.c_void_p, C.POINTER(PeerGroup)]),
    "fgs_peer_merge_reduce": (C.c_int, [C.c_void_p]),
    "fgs_peer_merge_select": (C.c_int, [C.c_void_p]),
    "fgs_peer_status": (C.c_int, [C.c_void_p, p_i32]),
    "fgs_synchronize": (C.c_int, [C.c_void_p]),
    "fgs_status_plane": (C.c_int, [C.c_void_p, C.c_uint32, p_u8]),
    "fgs_reason_plane": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint16)]),
    "fgs_fit_error_histogram": (C.c_int, [C.c_void_p, p_i32]),
    "fgs_schedule_sequential": (C.c_int, [C.c_void_p, C.POINTER(PodTable), C.POINTER(SeqOpts), C.POINTER(SeqResults)]),
    "fgs_reserve": (C.c_int, [C.c_void_p, C.POINTER(PodTable), p_i32, p_i32]),
    "fgs_gang_check": (C.c_int, [C.c_void_p, C.POINTER(Snapshot), C.POINTER(Podset), C.POINTER(Podgroups), p_u8, p_u8, p_i32]),
    "fgs_quota_check": (C.c_int, [C.c_void_p, C.POINTER(PodTable), C.POINTER(Quotas), C.POINTER(PodTable), p_i64, p_i64, p_u8]),
    # device pointers are passed as plain integers (c_void_p): the caller owns that memory
    "fgs_quota_check_static": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int64]),
    "fgs_get_timings": (C.c_int, [C.c_void_p, C.POINTER(Timings)]),
    "fgs_synth_create": (C.c_int, [C.POINTER(SynthParams), C.POINTER(C.c_void_p)]),
    "fgs_synth_destroy": (None, [C.c_void_p]),
    "fgs_synth_snapshot": (C.POINTER(Snapshot), [C.c_void_p]),
    "fgs_synth_pending": (C.POINTER(Podset), [C.c_void_p]),
    "fgs_synth_nominated": (C.POINTER(Podset), [C.c_void_p]),
    "fgs_synth_podgroups": (C.POINTER(Podgroups), [C.c_void_p]),
    "fgs_synth_quotas": (C.POINTER(Quotas), [C.c_void_p]),
}

_lib = None


def load():
    """Load libfgs.so and bind the ABI. Raises when the library is missing: no fallback exists."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA extension first (__graft_entry__.build()). "
            "flex-gpu-scheduler_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in ABI.items():
        fn = getattr(lib, name)  # AttributeError here = the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
