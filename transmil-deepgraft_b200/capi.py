"""ctypes binding of libtransmil_b200.so (the C ABI in include/transmil_b200.h).

The library is the product: there is no CPU path and no PyTorch fallback.  If the
shared object is missing this module raises at import of the first symbol.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtransmil_b200.so")

TM_FLAG_SIMT_ONLY = 1
TM_FLAG_KEEP_STAGES = 2
TM_FLAG_TF32_OPERANDS = 4
TM_FLAG_PINV_GEMM_CHAIN = 8
TM_FLAG_F16_ACT = 16
TM_FLAG_NO_OVERLAP = 32
TM_FLAG_X_F16 = 64
TM_FLAG_CLS_ONLY_L2 = 128
TM_FC1_SINGLE, TM_FC1_2048, TM_FC1_768, TM_FC1_TOKENS = 0, 1, 2, 3

_fp = C.c_void_p  # device/host float pointers travel as raw addresses


class Plan(C.Structure):
    _fields_ = [(k, C.c_int64) for k in ("n", "G", "N", "dup", "S", "l", "Sp", "P", "padding")]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


class LayerWeights(C.Structure):
    _fields_ = [(k, _fp) for k in ("norm_w", "norm_b", "qkv_w", "out_w", "out_b", "conv_w")]


class Weights(C.Structure):
    _fields_ = ([("fc1_kind", C.c_int32), ("in_features", C.c_int32), ("n_classes", C.c_int32), ("reserved", C.c_int32)]
                + [(k, _fp) for k in ("fc1_w0", "fc1_b0", "fc1_ln_w", "fc1_ln_b", "fc1_w1", "fc1_b1",
                                      "fc1_ln2_w", "fc1_ln2_b", "cls_token")]
                + [("layer1", LayerWeights), ("layer2", LayerWeights)]
                + [(k, _fp) for k in ("ppeg_w7", "ppeg_b7", "ppeg_w5", "ppeg_b5", "ppeg_w3", "ppeg_b3",
                                      "norm_w", "norm_b", "fc_w", "fc_b")])


class ShardPlanC(C.Structure):
    """tm_shard_plan: what one rank of a token-sharded bag owns (same integers as sharded.RankShard)."""
    _fields_ = ([("bag", Plan)] + [(k, C.c_int64) for k in ("world", "rank", "row0", "rows", "P_local", "S_local", "has_cls", "tok0",
                                                           "ntok", "n_x_ranges")]
                + [("x_range", (C.c_int64 * 2) * 2)]
                + [(k, C.c_int64) for k in ("y_in0", "y_in1", "halo_lo", "halo_hi", "n_up", "n_dn")])

    def x_ranges(self):
        return [(int(self.x_range[i][0]), int(self.x_range[i][1])) for i in range(int(self.n_x_ranges))]


TM_SHARD_HANDLE_BYTES = 64
TM_SHARD_MAX_WORLD = 16

# symbol -> (restype, argtypes); every symbol declared in include/transmil_b200.h
_i64, _u32, _sz = C.c_int64, C.c_uint32, C.c_size_t
_PW = C.POINTER(Weights)
SIGNATURES = {
    "tm_version": (C.c_char_p, []),
    "tm_last_error": (C.c_char_p, []),
    "tm_device_supported": (C.c_int, []),
    "tm_launch_count": (_i64, []),
    "tm_reset_launch_count": (None, []),
    "tm_profile_enable": (C.c_int, [C.c_int]),
    "tm_profile_stage_count": (C.c_int, []),
    "tm_profile_stage_name": (C.c_char_p, [C.c_int]),
    "tm_profile_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(_i64), C.c_int]),
    "tm_make_plan": (C.c_int, [_i64, C.POINTER(Plan)]),
    "tm_dup_source": (_i64, [C.POINTER(Plan), _i64]),
    "tm_landmark_group": (_i64, [C.POINTER(Plan), _i64]),
    "tm_padded_row": (_i64, [C.POINTER(Plan), _i64]),
    "tm_ppeg_pixel": (None, [C.POINTER(Plan), _i64, C.POINTER(_i64), C.POINTER(_i64)]),
    "tm_workspace_bytes": (C.c_int, [_i64, _i64, _PW, _u32, C.POINTER(_sz)]),
    "tm_transmil_fwd": (C.c_int, [_PW, _fp, _i64, _i64, _fp, _fp, _sz, _u32, _fp]),
    "tm_transmil_fwd_host": (C.c_int, [_PW, _fp, _i64, _i64, _fp, _fp, _fp, _fp, _sz, _u32, _fp]),
    "tm_stage_ptr": (C.c_int, [C.c_char_p, _i64, _i64, _PW, _u32, _fp, C.POINTER(C.c_void_p), C.POINTER(_i64)]),
    "tm_fc1_fwd": (C.c_int, [_PW, _fp, _i64, _i64, _fp, _fp, _u32, _fp]),
    "tm_ln_pad_fwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _i64, _fp, _u32, _fp]),
    "tm_cvt_f16": (C.c_int, [_fp, _fp, _i64, _fp]),
    "tm_cvt_f32": (C.c_int, [_fp, _fp, _i64, _fp]),
    "tm_round_tf32": (C.c_int, [_fp, _fp, _i64, _fp]),
    "tm_qkv_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _fp, _fp, _fp, _u32, _fp]),
    "tm_landmarks_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _i64, _fp, _fp, _fp]),
    "tm_pinv_scratch_floats": (_i64, [_i64]),
    "tm_pinv_fwd": (C.c_int, [_fp, _fp, _i64, _fp, _fp, _fp, _u32, _fp]),
    "tm_attn3v_scratch_floats": (_i64, [_i64, _i64]),
    "tm_attn3v_fwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _fp, _u32, _fp]),
    "tm_zw_fwd": (C.c_int, [_fp, _fp, _i64, _fp, _fp]),
    "tm_nystrom_out_fwd": (C.c_int, [_fp, _fp, _fp, _fp, _fp, _i64, _i64, _fp, _u32, _fp]),
    "tm_qkv16_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _fp, _fp, _fp, _u32, _fp]),
    "tm_qkv16_cols_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _fp, _fp, _fp, _i64, _i64, _i64, _fp, _u32, _fp]),
    "tm_landmarks_parts_fwd": (C.c_int, [_fp, _i64, _i64, _i64, _fp, _fp, _fp, _fp, _fp]),
    "tm_landmarks16_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _i64, _fp, _fp, _fp, _fp, _fp]),
    "tm_stage_is_f16": (C.c_int, [C.c_char_p, _u32]),
    "tm_zw16_fwd": (C.c_int, [_fp, _fp, _i64, _fp, _fp, _fp]),
    "tm_mt16_fwd": (C.c_int, [_fp, _fp, _i64, _fp]),
    "tm_nystrom_out_tc_fwd": (C.c_int, [_fp, _fp, _fp, _fp, _fp, _i64, _i64, _fp, _u32, _fp]),
    "tm_attn3v_tc_scratch_floats": (_i64, [_i64, _i64]),
    "tm_attn3v_tc_fwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _fp, _u32, _fp]),
    "tm_attn3v_tc_fwd_static": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _u32, _fp]),
    "tm_to_out_residual_fwd": (C.c_int, [_fp, _fp, _fp, _fp, _i64, _i64, _i64, _fp, _u32, _fp]),
    "tm_ppeg_fwd": (C.c_int, [_PW, _fp, _i64, _i64, _fp, _fp, _fp]),
    "tm_head_fwd": (C.c_int, [_PW, _fp, _i64, _i64, _fp, _fp]),
    "tm_attn_rows_scratch_floats": (_i64, [_i64, _i64, _i64]),
    "tm_attn_rows_fwd": (C.c_int, [_fp, _fp, _fp, _fp, _fp, _i64, _i64, C.POINTER(_i64), _i64, _fp, _fp, _fp]),
    "tm_fc1_rows_fwd": (C.c_int, [_PW, _fp, _i64, _fp, _fp, _u32, _fp]),
    "tm_landmark_sums16_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _i64, _i64, _fp, _fp]),
    "tm_landmark_finish_fwd": (C.c_int, [_fp, _i64, _i64, _fp, _fp, _fp, _fp, _fp]),
    "tm_attn3v_partial_floats": (_i64, [_i64]),
    "tm_attn3v_tc_partial_fwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _fp, _fp]),
    "tm_attn3v_merge_fwd": (C.c_int, [_fp, _i64, _i64, _fp, _fp]),
    "tm_conv_halo_fix_fwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _fp]),
    "tm_ppeg_rows_fwd": (C.c_int, [_PW, _fp, _i64, _i64, _i64, _i64, _i64, _fp, _fp, _fp]),
    "tm_gemm_nt": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _i64, _u32, _fp]),
    "tm_make_shard_plan": (C.c_int, [_i64, _i64, _i64, C.POINTER(ShardPlanC)]),
    "tm_shard_comm_create": (C.c_int, [C.c_int, C.c_int, _i64, C.POINTER(C.c_void_p)]),
    "tm_shard_comm_bytes": (_i64, [C.c_void_p]),
    "tm_shard_comm_handle": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tm_shard_comm_attach": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tm_shard_comm_base": (C.c_void_p, [C.c_void_p]),
    "tm_shard_comm_attach_ptrs": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int)]),
    "tm_shard_comm_status": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32)]),
    "tm_shard_comm_destroy": (C.c_int, [C.c_void_p]),
    "tm_sharded_workspace_bytes": (C.c_int, [_PW, C.POINTER(ShardPlanC), C.POINTER(_sz)]),
    "tm_transmil_fwd_sharded": (C.c_int, [_PW, _fp, C.POINTER(ShardPlanC), C.c_void_p, _fp, _fp, _sz, _u32, _fp]),
}

# backward entry points (csrc/backward.cu); transmil-deepgraft_b200/training.py is their only caller
_f32, _u64 = C.c_float, C.c_uint64
BWD_SIGNATURES = {
    "tm_bgemm": (C.c_int, [_fp, _fp, _fp, _fp] + [_i64] * 14 + [_f32, _f32, _fp]),
    "tm_bgemm_ex": (C.c_int, [_fp, _fp, _fp, _fp] + [_i64] * 14 + [_f32, _f32, _fp, _i64, _fp]),
    "tm_bgemm_scratch_floats": (_i64, [_i64, _i64, _i64, _i64]),
    "tm_bgemm_engine": (C.c_int, [C.c_int]),
    "tm_softmax_rows": (C.c_int, [_fp, _i64, _i64, _fp]),
    "tm_softmax_rows_bwd": (C.c_int, [_fp, _fp, _i64, _i64, _fp]),
    "tm_ln_rows_fwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _fp]),
    "tm_ln_bwd_scratch_floats": (_i64, [_i64, _i64]),
    "tm_ln_rows_bwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _i64, _i64, _i64, _fp, _fp, _fp, _fp, _fp, _fp]),
    "tm_colsum_scratch_floats": (_i64, [_i64, _i64]),
    "tm_colsum": (C.c_int, [_fp, _i64, _i64, _i64, _fp, _fp, _fp]),
    "tm_gelu_fwd": (C.c_int, [_fp, _fp, _i64, _fp]),
    "tm_gelu_bwd": (C.c_int, [_fp, _fp, _fp, _i64, _fp]),
    "tm_dropout_resid_fwd": (C.c_int, [_fp, _fp, _i64, _u64, _f32, _fp]),
    "tm_dropout_fwd": (C.c_int, [_fp, _fp, _i64, _u64, _f32, _fp]),
    "tm_pad_front_bwd": (C.c_int, [_fp, _i64, _i64, _i64, _i64, _u64, _f32, _fp, _fp]),
    "tm_split_heads": (C.c_int, [_fp, _i64, _i64, _fp, _fp]),
    "tm_qkv_grad_merge": (C.c_int, [_fp, _fp, _fp, _fp, _fp, _i64, _i64, _i64, _fp, _fp]),
    "tm_conv33_bwd_scratch_floats": (_i64, [_i64, _i64]),
    "tm_conv33_bwd": (C.c_int, [_fp, _fp, _fp, _i64, _i64, _fp, _fp, _fp, _fp]),
    "tm_pinv_bwd_scratch_floats": (_i64, [_i64]),
    "tm_pinv_bwd": (C.c_int, [_fp, _fp, _i64, _fp, _fp, _fp]),
    "tm_ppeg_bwd_scratch_floats": (_i64, [_i64, _i64]),
    "tm_ppeg_bwd": (C.c_int, [_PW, _fp, _fp, _i64, _i64] + [_fp] * 9),
    "tm_head_bwd": (C.c_int, [_PW, _fp, _fp, _i64, _i64, _fp, _fp, _fp, _fp, _fp, _fp]),
    "tm_dup_cls_fwd": (C.c_int, [_fp, _fp, _i64, _i64, _i64, _fp, _fp]),
    "tm_dup_cls_bwd": (C.c_int, [_fp, _i64, _i64, _i64, _fp, _fp]),
}


class OptTensor(C.Structure):
    _fields_ = [("p", _fp), ("g", _fp), ("m", _fp), ("v", _fp), ("n", _i64), ("weight_decay", C.c_float), ("reserved", C.c_float)]


TM_OPT_MAX_TENSORS = 48
BWD_SIGNATURES.update({
    "tm_radam_step": (C.c_int, [C.POINTER(OptTensor), _i64, _i64, _f32, _f32, _f32, _f32, _fp]),
    "tm_lookahead_sync": (C.c_int, [C.POINTER(OptTensor), _i64, _f32, _fp]),
})
SIGNATURES.update(BWD_SIGNATURES)

_lib = None


class TransmilError(RuntimeError):
    pass


def lib():
    """The loaded shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise TransmilError(
                f"{LIB_PATH} is missing: build it with `python transmil-deepgraft_b200/build.py` "
                "(there is no CPU or PyTorch fallback for this path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


_STATUS = {-1: "bad shape", -2: "bad alignment", -3: "unsupported arch", -4: "workspace too small",
           -5: "CUDA error", -6: "bad argument"}


def check(status):
    if status != 0:
        msg = lib().tm_last_error().decode("utf-8", "replace")
        exc = ValueError if status in (-1, -2, -6) else TransmilError
        raise exc(f"libtransmil_b200: {_STATUS.get(status, status)}: {msg}")


def make_plan(n):
    p = Plan()
    check(lib().tm_make_plan(int(n), C.byref(p)))
    return p


def profile_read():
    """{stage: (total_ms, intervals)} since the last read (see tm_profile_enable)."""
    L = lib()
    n = L.tm_profile_stage_count()
    ms = (C.c_double * n)()
    cnt = (C.c_int64 * n)()
    check(L.tm_profile_read(ms, cnt, n))
    return {L.tm_profile_stage_name(i).decode(): (ms[i], int(cnt[i])) for i in range(n) if cnt[i] > 0}
