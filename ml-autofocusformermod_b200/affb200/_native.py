"""ctypes binding of libaffb200.so (C ABI in include/affb200.h).

There is deliberately no CPU or pure-PyTorch fallback here: if the shared library is missing the import raises,
mirroring the reference's behaviour when its extension cannot be imported (reference clusten/clusten.py:8-14).
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libaffb200.so")

_i64 = ctypes.c_int64
_ptr = ctypes.c_void_p
_int = ctypes.c_int



class NbhdPlanStruct(ctypes.Structure):
    """affb200_nbhd_plan (include/affb200.h): device pointers of the structures derived from one nbhd_idx tensor."""
    _fields_ = [("csr_off", _ptr), ("csr_ent", _ptr),
                ("tq", ctypes.c_int32), ("q_hdr", _ptr), ("q_rows", _ptr), ("q_loc", _ptr),
                ("tk", ctypes.c_int32), ("k_hdr", _ptr), ("k_utok", _ptr), ("k_rloc", _ptr)]


_plan_p = ctypes.POINTER(NbhdPlanStruct)

PLAN_UMAX = 256
PLAN_MAX_NK = 131072
PLAN_MAX_M = 192
FUSED_IMG_MAX_N = 255

# name -> (restype, argtypes); kept in one table so tests can check it against include/affb200.h
SIGNATURES = {
    "affb200_abi_version": (_int, []),
    "affb200_last_error": (ctypes.c_char_p, []),
    "affb200_target_arch": (ctypes.c_char_p, []),
    "affb200_launch_count": (_i64, []),
    "affb200_csr_off_bytes": (ctypes.c_size_t, [_i64, _i64]),
    "affb200_csr_ent_bytes": (ctypes.c_size_t, [_i64, _i64, _i64]),
    "affb200_csr_work_bytes": (ctypes.c_size_t, [_i64, _i64]),
    "affb200_csr_build": (_int, [_ptr, _i64, _i64, _i64, _i64, _ptr, _ptr, _ptr, _int, _ptr]),
    "affb200_plan_tiles": (_i64, [_i64, _i64]),
    "affb200_qplan_build": (_int, [_ptr] + [_i64] * 5 + [_ptr] * 5),
    "affb200_kplan_build": (_int, [_ptr] + [_i64] * 5 + [_ptr] * 6),
    "affb200_sfc_work_bytes": (ctypes.c_size_t, [_i64, _i64]),
    "affb200_sfc_cluster": (_int, [_ptr] + [_i64] * 5 + [_ptr] * 5),
    "affb200_curve_key": (_int, [_ptr, _i64, _i64, _i64, _i64, _int, _ptr, _ptr]),
    "affb200_sfc_cluster_tables": (_int, [_ptr] + [_i64] * 5 + [ctypes.c_float, ctypes.c_float] + [_ptr] * 8),
    "affb200_permute_rows": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _int, _ptr]),
    "affb200_argsort_work_bytes": (ctypes.c_size_t, [_i64, _i64]),
    "affb200_argsort_f32": (_int, [_ptr, _i64, _i64, _int, _ptr, _ptr, _ptr]),
    "affb200_nearest_l1": (_int, [_ptr, _ptr] + [_i64] * 3 + [_ptr, _ptr]),
    "affb200_knn2d": (_int, [_ptr, _ptr] + [_i64] * 4 + [_ptr] * 3),
    "affb200_nbhd_build": (_int, [_ptr, _ptr] + [_i64] * 6 + [_int] + [_ptr] * 4),
    "affb200_merge_select_work_bytes": (ctypes.c_size_t, [_i64, _i64]),
    "affb200_merge_select": (_int, [_ptr, _ptr, _int, _i64, _i64, _i64, ctypes.c_double, _i64, _i64, _int] + [_ptr] * 5),
    "affb200_gather_rows_i64": (_int, [_ptr, _ptr] + [_i64] * 4 + [_ptr, _ptr]),
    "affb200_attn_bias_add": (_int, [_ptr] * 4 + [_i64] * 4 + [_int, _ptr]),
    "affb200_attn_bias_bwd": (_int, [_ptr, _ptr] + [_i64] * 5 + [_int, _ptr, _int, _ptr]),
    "affb200_fused_attn_pack_pe": (_int, [_ptr, _ptr, _i64, _ptr, _ptr, _ptr]),
    "affb200_fused_attn_fwd": (_int, [_ptr] * 4 + [_ptr] * 3 + [_plan_p, _ptr, ctypes.c_int32, _ptr] + [_i64] * 5
                               + [ctypes.c_float, _int, _ptr]),
    "affb200_fused_attn_bwd": (_int, [_ptr] * 8 + [_ptr] * 3 + [_plan_p, _ptr, ctypes.c_int32, _ptr] + [_ptr] * 4
                               + [_i64] * 6 + [ctypes.c_float, _int, _ptr]),
    "affb200_fused_attn_img_prepare": (_int, [_ptr, _ptr] + [_i64] * 3 + [_ptr, _ptr, _ptr]),
    "affb200_fused_attn_img_fwd": (_int, [_ptr] * 4 + [_ptr] * 3 + [_ptr, _ptr, _ptr, _ptr] + [_i64] * 5
                                   + [ctypes.c_float, _int, _ptr]),
    "affb200_fused_attn_img_bwd": (_int, [_ptr] * 8 + [_ptr] * 3 + [_ptr, _ptr, _ptr, _ptr] + [_ptr] * 3
                                   + [_i64] * 6 + [ctypes.c_float, _int, _ptr]),
    "affb200_cl_tiles": (_i64, [_i64]),
    "affb200_cl_qplan_build": (_int, [_ptr] * 4 + [_i64] * 5 + [_ptr] * 6 + [_ptr]),
    "affb200_cl_qplan_build_split": (_int, [_ptr] * 4 + [_i64] * 5 + [_ptr] * 8 + [_ptr]),
    "affb200_cl_kplan_build": (_int, [_ptr] + [_i64] * 4 + [_ptr] * 3 + [_ptr]),
    "affb200_cl_pack_pe_sorted": (_int, [_ptr] * 3 + [_i64] * 3 + [_ptr] * 3),
    "affb200_fused_cld_fwd": (_int, [_ptr] * 4 + [_ptr, _i64] + [_ptr] * 4 + [_i64, _ptr] + [_i64] * 7
                              + [ctypes.c_float, _int, _ptr]),
    "affb200_fused_cld_bwd": (_int, [_ptr] * 8 + [_ptr, _i64] + [_ptr] * 4 + [_i64] + [_ptr] * 3 + [_i64] * 7
                              + [ctypes.c_float, _int, _ptr]),
    "affb200_fused_cl_bwd_grid": (_i64, [_i64] * 3),
    "affb200_fused_cl_bwd": (_int, [_ptr] * 8 + [_ptr] * 3 + [_ptr] + [_ptr] * 5 + [_i64] * 7
                             + [ctypes.c_float, _int, _ptr]),
    "affb200_fused_cl_fwd": (_int, [_ptr] * 4 + [_ptr] * 3 + [_ptr, _ptr, _ptr] + [_i64] * 6
                             + [ctypes.c_float, _int, _ptr]),
    "affb200_table_grad_u16": (_int, [_ptr, _ptr] + [_i64] * 4 + [_ptr, _int, _ptr]),
    "affb200_pe_feat_grad_work_floats": (_i64, [_i64, _i64, _i64]),
    "affb200_pe_feat_grad": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _ptr, _ptr, _int, _ptr]),
    "affb200_merge_fwd": (_int, [_ptr] * 7 + [_i64] * 5 + [_int, _int, _ptr]),
    "affb200_merge_bwd": (_int, [_ptr] * 12 + [_i64] * 5 + [_int, _int, _ptr]),
    "affb200_table_gather": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _int, _ptr]),
    "affb200_conv_stem_fwd": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _i64, _int, _ptr]),
    "affb200_bias_rows": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _int, _ptr]),
    "affb200_scale_add": (_int, [_ptr, _ptr, _ptr, ctypes.c_float, _ptr, _i64, _i64, _int, _ptr]),
    "affb200_batch_norm_work_floats": (_i64, [_i64, _i64]),
    "affb200_batch_norm_fwd": (_int, [_ptr] * 5 + [ctypes.c_float, ctypes.c_float, _int] + [_ptr] * 4
                               + [_i64, _i64, _int, _ptr]),
    "affb200_batch_norm_apply": (_int, [_ptr] * 5 + [_int, _ptr, _i64, _i64, _int, _ptr]),
    "affb200_batch_norm_bwd": (_int, [_ptr] * 6 + [_int] + [_ptr] * 4 + [_i64, _i64, _int, _ptr, _ptr]),
    "affb200_gelu_tables": (_int, [_int, _ptr, _ptr, _ptr]),
    "affb200_gelu_fwd": (_int, [_ptr, _ptr, _i64, _int, _ptr, _ptr]),
    "affb200_gelu_bwd_work_floats": (_i64, [_i64]),
    "affb200_gelu_bwd": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _int, _ptr, _ptr, _int, _ptr, _ptr]),
    "affb200_weight_net_work_floats": (_i64, [_i64]),
    "affb200_weight_net_fwd": (_int, [_ptr] * 5 + [ctypes.c_float, _i64, _int, _ptr, _ptr]),
    "affb200_weight_net_bwd": (_int, [_ptr] * 5 + [ctypes.c_float, _i64, _int, _ptr, _ptr, _ptr, _ptr]),
    "affb200_col_sum_work_floats": (_i64, [_i64, _i64]),
    "affb200_col_sum": (_int, [_ptr, _ptr, _ptr, _i64, _i64, _int, _int, _ptr]),
    "affb200_col_sum_static_max_floats": (_i64, []),
    "affb200_col_sum_static": (_int, [_ptr, _ptr, _i64, _i64, _int, _int, _ptr]),
    "affb200_layer_norm_fwd": (_int, [_ptr] * 6 + [_i64, _i64, ctypes.c_float, _int, _int, _ptr]),
    "affb200_layer_norm_bwd_work_floats": (_i64, [_i64, _i64]),
    "affb200_layer_norm_bwd": (_int, [_ptr] * 10 + [_i64, _i64, _int, _int, _int, _ptr]),
    "affb200_qk_fwd": (_int, [_ptr, _ptr, _ptr, _plan_p, _ptr] + [_i64] * 6 + [_int, _ptr]),
    "affb200_qk_bwd": (_int, [_ptr] * 4 + [_plan_p] + [_ptr] * 2 + [_i64] * 6 + [_int, _ptr]),
    "affb200_av_fwd": (_int, [_ptr, _ptr, _ptr, _plan_p, _ptr] + [_i64] * 6 + [_int, _ptr]),
    "affb200_av_bwd": (_int, [_ptr] * 4 + [_plan_p] + [_ptr] * 2 + [_i64] * 6 + [_int, _ptr]),
    "affb200_wf_fwd": (_int, [_ptr, _ptr, _ptr, _plan_p, _ptr] + [_i64] * 6 + [_int, _ptr]),
    "affb200_wf_bwd": (_int, [_ptr] * 4 + [_plan_p] + [_ptr] * 2 + [_i64] * 6 + [_int, _ptr]),
}

DTYPE_ENUM = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2, torch.float64: 3}


def _load():
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "Could not load CLUSTEN CUDA extension. libaffb200.so is missing at %s; build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` or `make -C ml-autofocusformermod_b200/csrc`."
            % LIB_PATH)
    try:
        lib = ctypes.CDLL(LIB_PATH)
    except OSError as e:  # pragma: no cover - depends on the machine
        raise RuntimeError("Could not load CLUSTEN CUDA extension. " + str(e))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.affb200_abi_version() != 5:
        raise RuntimeError("libaffb200.so ABI version mismatch")
    return lib


LIB = _load()


def last_error():
    return LIB.affb200_last_error().decode()


def check(code, what):
    if code != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, code, last_error()))


def dtype_enum(t):
    try:
        return DTYPE_ENUM[t.dtype]
    except KeyError:
        raise RuntimeError('"%s" not implemented for \'%s\'' % ("clusten", str(t.dtype).replace("torch.", "")))


def stream_ptr(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def p(t):
    """device pointer of a tensor (None -> NULL)"""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())
