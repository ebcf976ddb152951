"""ctypes binding of libpcp_b200.so (include/pcp.h).

There is no CPU fallback: if the library is missing `load()` raises, and every compute entry point
returns PCP_ERR_CUDA when no CUDA device is present.  The oracle/ directory is never imported here.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libpcp_b200.so"
_lib = None

PRECISION_FP32 = 0
PRECISION_BF16 = 1
PRECISION_FP32_TC = 2
HEAD_CLS = 1
HEAD_SEG = 2
GROUP_SHARED, GROUP_INPUT_TRANSFORM, GROUP_CLS_HEAD, GROUP_SEG_HEAD = 0, 1, 2, 3


class PcpError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"libpcp_b200 error {status}: {msg}")
        self.status = status


class ModelDesc(C.Structure):
    _fields_ = [
        ("num_classes", C.c_int32), ("num_parts", C.c_int32), ("vanilla", C.c_int32),
        ("regularize_input_transform", C.c_int32), ("regularize_feature_transform", C.c_int32),
        ("dropout_rate", C.c_float), ("precision", C.c_int32),
    ]


class TrainConfig(C.Structure):
    _fields_ = [
        ("w_cls", C.c_float), ("w_seg", C.c_float), ("w_rot", C.c_float),
        ("lr0", C.c_float), ("decay_rate", C.c_float), ("decay_steps", C.c_int32),
        ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("dropout_seed", C.c_uint64), ("sample_offset", C.c_int64), ("use_dropout_masks", C.c_int32),
        ("skip_unweighted_heads", C.c_int32),
    ]


_vp, _i32, _i64, _u32, _u64, _sz, _f = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64, C.c_size_t, C.c_float
_fp = C.POINTER(C.c_float)

# name -> (restype, argtypes); every name here must be declared in include/pcp.h (tests check both ways)
# pcp_allreduce_fn (include/pcp.h): int fn(void *user, void *d_buf, size_t count, int dtype, pcp_stream stream)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p)

SIGNATURES = {
    "pcp_last_error": (C.c_char_p, []),
    "pcp_version": (_i32, []),
    "pcp_launch_count": (_u64, []),
    "pcp_device_info": (_i32, [C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_sz)]),
    "pcp_malloc": (_i32, [C.POINTER(_vp), _sz]),
    "pcp_free": (_i32, [_vp]),
    "pcp_malloc_host": (_i32, [C.POINTER(_vp), _sz]),
    "pcp_free_host": (_i32, [_vp]),
    "pcp_memcpy_h2d": (_i32, [_vp, _vp, _sz, _vp]),
    "pcp_memcpy_d2h": (_i32, [_vp, _vp, _sz, _vp]),
    "pcp_memset": (_i32, [_vp, _i32, _sz, _vp]),
    "pcp_stream_sync": (_i32, [_vp]),
    "pcp_normalize_f32": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp, _vp]),
    "pcp_normalize_workspace_bytes": (_i32, [_i32, _i64, C.POINTER(_sz)]),
    "pcp_normalize_ws_f32": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "pcp_normalize_noalloc_max_clouds": (_i32, []),
    "pcp_normalize_noalloc_f32": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp, _vp]),
    "pcp_jitter_f32": (_i32, [_vp, _i64, _fp, _u64, _u64, _vp, _vp, _vp]),
    "pcp_philox_normal_f32": (_i32, [_i64, _u64, _u64, _vp, _vp]),
    "pcp_rotate_f32": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp]),
    "pcp_resize_to_width_f32": (_i32, [_vp, _vp, _i32, _i64, _vp, _i64, _u64, _vp, _vp, _vp]),
    "pcp_fps_f32": (_i32, [_vp, _i32, _i64, _i32, _i32, _vp, _vp]),
    "pcp_fps_workspace_bytes": (_i32, [_i32, _i64, _i32, C.POINTER(_sz)]),
    "pcp_fps_ws_f32": (_i32, [_vp, _i32, _i64, _i32, _i32, _vp, _vp, _sz, _vp]),
    "pcp_fps_phase_cycles": (_i32, [C.POINTER(C.c_ulonglong)]),
    "pcp_gather_points_f32": (_i32, [_vp, _vp, _i32, _i64, _vp, _i32, _vp, _vp, _vp]),
    "pcp_voxel_workspace_bytes": (_i32, [_i32, _i64, C.POINTER(_sz)]),
    "pcp_voxel_f32": (_i32, [_vp, _i32, _i64, _fp, _f, _vp, _vp, _vp, _vp, _sz, _vp]),
    "pcp_voxel_compact_f32": (_i32, [_vp, _i32, _i64, _fp, _f, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "pcp_model_create": (_i32, [C.POINTER(ModelDesc), C.POINTER(_vp)]),
    "pcp_model_destroy": (_i32, [_vp]),
    "pcp_model_num_tensors": (_i32, [_vp]),
    "pcp_model_tensor_info": (_i32, [_vp, _i32, C.c_char_p, _sz, C.POINTER(_sz), C.POINTER(_i32)]),
    "pcp_model_set_tensor": (_i32, [_vp, C.c_char_p, _fp, _sz]),
    "pcp_model_get_tensor": (_i32, [_vp, C.c_char_p, _fp, _sz]),
    "pcp_model_prepare": (_i32, [_vp]),
    "pcp_model_reserve": (_i32, [_vp, _i32, _i64]),
    "pcp_model_forward": (_i32, [_vp, _vp, _i32, _i64, _u32, _vp, _vp, _vp, _vp, _vp, _vp]),
    "pcp_model_predict_host": (_i32, [_vp, _fp, _i32, _i64, _u32, _fp, _fp, _fp, _vp]),
    "pcp_model_get_pooled": (_i32, [_vp, _i32, _i32, _fp]),
    "pcp_model_enable_timing": (_i32, [_vp, _i32]),
    "pcp_model_chain_time": (_i32, [_vp, _fp, C.POINTER(_i32)]),
    "pcp_selftest_umma": (_i32, [_fp, _fp, _i32, _i32, _fp]),
    "pcp_selftest_umma2": (_i32, [_fp, _fp, _i32, _i32, _fp]),
    "pcp_selftest_wgrad": (_i32, [_fp, _fp, _i64, _i32, _i32, _fp]),
    "pcp_selftest_tc_mm": (_i32, [_fp, _fp, _i64, _i32, _i32, _i32, _i32, _fp]),
    "pcp_selftest_tc_mm_splitk": (_i32, [_fp, _fp, _i64, _i32, _i32, _i32, _i32, _fp]),
    "pcp_selftest_tc_gemm_bf16": (_i32, [_fp, _fp, _fp, _i64, _i32, _i32, _i32, _i32, _i32, _fp]),
    "pcp_bench_tc_gemm": (_i32, [_i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _fp]),
    "pcp_selftest_pooled_layer": (_i32, [_fp, _fp, _fp, _fp, _fp, _i32, _i64, _i32, _i32, _fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "pcp_dcm_from_rpy_f32": (_i32, [_vp, _i32, _vp, _vp]),
    "pcp_rpy_from_dcm_f32": (_i32, [_vp, _i32, _vp, _vp]),
    "pcp_reorthogonalize_f32": (_i32, [_vp, _i32, _vp, _vp]),
    "pcp_kabsch_f32": (_i32, [_vp, _vp, _i32, _i64, _vp, _vp, _vp, _vp]),
    "pcp_model_set_trainable": (_i32, [_vp, _i32, _i32]),
    "pcp_model_set_bn_allreduce": (_i32, [_vp, _vp, _vp]),
    "pcp_selftest_set_bn_allreduce": (_i32, [_vp, _vp, C.c_double]),
    "pcp_model_loss_backward": (_i32, [_vp, _vp, _i32, _i64, _vp, _vp, _vp, C.POINTER(TrainConfig), _i32, _vp, _vp, _fp, _vp]),
    "pcp_model_evaluate": (_i32, [_vp, _vp, _i32, _i64, _vp, _vp, _vp, C.POINTER(TrainConfig), _fp, _vp]),
    "pcp_model_grad_buffer": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "pcp_model_adam_step": (_i32, [_vp, C.POINTER(TrainConfig), _i64, _vp]),
    "pcp_model_reset_optimizer": (_i32, [_vp]),
    "pcp_crc32c": (_u32, [_vp, _sz]),
    "pcp_crc32c_masked": (_u32, [_vp, _sz]),
    "pcp_tfrecord_scan": (_i32, [_vp, _sz, _i32, _vp, _vp, _sz, C.POINTER(_sz)]),
    "pcp_example_decode": (_i32, [_vp, _sz, _i64, _vp, C.POINTER(_i64), _vp, _vp]),
    "pcp_example_decode_batch": (_i32, [_vp, _vp, _i32, _i64, _vp, _vp, _vp, _vp, _i32]),
    "pcp_parse_aftr_frame": (_i32, [C.c_char_p, _sz, C.c_char_p, C.c_char_p, _vp, _vp, _sz, C.POINTER(_sz), C.POINTER(_i32), C.POINTER(_sz)]),
    "pcp_format_aftr_frame": (_i32, [_vp, _i32, _sz, _vp, _i32, C.c_char_p, _vp, _sz, C.POINTER(_sz)]),
    "pcp_model_get_grad": (_i32, [_vp, C.c_char_p, _fp, _sz]),
    "pcp_model_get_train_output": (_i32, [_vp, _i32, _fp, _sz]),
}


def lib_path() -> Path:
    return Path(os.environ.get("PCP_B200_LIB", str(_LIB_PATH)))


def load():
    """Load the shared library (once) and bind every symbol of include/pcp.h."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not p.exists():
        raise ImportError(
            f"{p} is missing: build it with `python -m pointcloudprocessing_b200.build` "
            "(hand-written sm_100a CUDA; there is no CPU or PyTorch fallback)")
    lib = C.CDLL(str(p))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError here == header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int):
    if status != 0:
        raise PcpError(status, load().pcp_last_error().decode(errors="replace"))


def ptr(t):
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return C.c_void_p(t.ctypes.data)


def fptr(a):
    """float* of a C-contiguous float32 numpy array (None -> NULL)."""
    if a is None:
        return None
    return a.ctypes.data_as(_fp)


def stream_ptr(stream=None):
    """cudaStream_t of a torch stream (default: torch's current stream)."""
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
