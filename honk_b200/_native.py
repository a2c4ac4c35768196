"""ctypes binding of libhonk_b200.so (include/honk_b200.h) + thin torch custom ops.

Python reaches CUDA only through this module.  There is no CPU fallback: if the
shared library is missing, or a tensor is not on a CUDA device, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhonk_b200.so")

PCM_I16, PCM_F32, PCM_F64 = 0, 1, 2
FAMILY_RES, FAMILY_CNN = 0, 1

EXPORTS = [
    "honk_abi_version", "honk_last_error", "honk_device_check", "honk_frontend_create",
    "honk_frontend_destroy", "honk_frontend_num_frames", "honk_frontend_get_dct", "honk_frontend_get_mel",
    "honk_frontend_mfcc", "honk_frontend_workspace_bytes", "honk_frontend_path", "honk_frontend_mfcc_noalloc", "honk_model_create", "honk_model_destroy", "honk_model_workspace_bytes",
    "honk_model_forward", "honk_model_forward_pcm", "honk_model_workspace_bytes_pcm",
    "honk_model_forward_pcm_host", "honk_model_label_pcm_host", "honk_model_workspace_bytes_pcm_host",
    "honk_label_from_logits", "honk_eval_metrics", "honk_load_audio", "honk_listen", "honk_listen_num_windows", "honk_listen_workspace_bytes", "honk_launch_count", "honk_model_path", "honk_profile_enable",
    "honk_profile_reset", "honk_profile_get",
    "honk_trainer_create", "honk_trainer_destroy", "honk_trainer_workspace_bytes", "honk_trainer_step",
    "honk_trainer_reset_optimizer", "honk_trainer_steps", "honk_trainer_read", "honk_trainer_set_dropout",
    "honk_trainer_set_dropout_masks",
]


class FrontendConfig(C.Structure):
    _fields_ = [("sr", C.c_int32), ("n_fft", C.c_int32), ("hop_length", C.c_int32), ("n_mels", C.c_int32),
                ("n_dct_filters", C.c_int32), ("f_min", C.c_float), ("f_max", C.c_float)]


class ModelConfig(C.Structure):
    _fields_ = [("family", C.c_int32), ("n_labels", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
                ("n_feature_maps", C.c_int32), ("n_layers", C.c_int32), ("use_dilation", C.c_int32),
                ("res_pool", C.c_int32 * 2),
                ("n_feature_maps1", C.c_int32), ("conv1_size", C.c_int32 * 2), ("conv1_pool", C.c_int32 * 2),
                ("conv1_stride", C.c_int32 * 2), ("has_conv2", C.c_int32), ("n_feature_maps2", C.c_int32),
                ("conv2_size", C.c_int32 * 2), ("conv2_pool", C.c_int32 * 2), ("conv2_stride", C.c_int32 * 2),
                ("tf_variant", C.c_int32), ("dnn1_size", C.c_int32), ("dnn2_size", C.c_int32),
                ("precision", C.c_int32)]


class Sgd(C.Structure):
    _fields_ = [("lr", C.c_float), ("momentum", C.c_float), ("weight_decay", C.c_float), ("nesterov", C.c_int32)]


class Tensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("h_data", C.c_void_p), ("ndim", C.c_int32), ("dims", C.c_int64 * 4)]


_lib = None
_lock = threading.Lock()


def lib():
    """Loads the shared library once; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "honk_b200: %s is missing - run `python -m honk_b200.build` (needs nvcc). "
                "There is no CPU fallback." % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        L.honk_abi_version.restype = C.c_int
        L.honk_last_error.restype = C.c_char_p
        L.honk_device_check.argtypes = [C.c_int]
        L.honk_frontend_create.argtypes = [C.POINTER(FrontendConfig), C.c_int, C.POINTER(vp)]
        L.honk_frontend_destroy.argtypes = [vp]
        L.honk_frontend_num_frames.argtypes = [vp, i64]
        L.honk_frontend_num_frames.restype = i64
        L.honk_frontend_get_dct.argtypes = [vp, vp]
        L.honk_frontend_get_mel.argtypes = [vp, vp]
        L.honk_frontend_mfcc.argtypes = [vp, vp, C.c_int, i64, i64, i64, vp, vp, i64, vp]
        L.honk_frontend_workspace_bytes.argtypes = [vp, i64, i64, C.c_int]
        L.honk_frontend_workspace_bytes.restype = i64
        L.honk_frontend_mfcc_noalloc.argtypes = [vp, vp, i64, i64, i64, vp, vp]
        L.honk_frontend_path.argtypes = [vp]
        L.honk_frontend_path.restype = C.c_char_p
        L.honk_model_create.argtypes = [C.POINTER(ModelConfig), C.POINTER(Tensor), i32, C.c_int, C.POINTER(vp)]
        L.honk_model_destroy.argtypes = [vp]
        L.honk_model_workspace_bytes.argtypes = [vp, i64]
        L.honk_model_workspace_bytes.restype = i64
        L.honk_model_forward.argtypes = [vp, vp, i64, i32, i32, vp, vp, i64, vp]
        L.honk_model_forward_pcm.argtypes = [vp, vp, vp, C.c_int, i64, i64, i64, vp, vp, i64, vp]
        L.honk_model_workspace_bytes_pcm.argtypes = [vp, vp, i64, i64]
        L.honk_model_workspace_bytes_pcm.restype = i64
        L.honk_model_forward_pcm_host.argtypes = [vp, vp, vp, C.c_int, i64, i64, i64, vp, vp, i64, vp]
        L.honk_model_label_pcm_host.argtypes = [vp, vp, vp, C.c_int, i64, i64, i64, vp, vp, vp, i64, vp]
        L.honk_model_workspace_bytes_pcm_host.argtypes = [vp, vp, i64, i64, C.c_int]
        L.honk_model_workspace_bytes_pcm_host.restype = i64
        L.honk_label_from_logits.argtypes = [vp, i64, i32, vp, vp, vp]
        L.honk_eval_metrics.argtypes = [vp, vp, i64, i32, vp, vp, vp, vp]
        L.honk_load_audio.argtypes = [vp, i64, i64, vp, vp, vp, vp, vp, i32, vp, vp]
        L.honk_listen_num_windows.argtypes = [i64, i64, i64]
        L.honk_listen_num_windows.restype = i64
        L.honk_listen_workspace_bytes.argtypes = [vp, vp, i64, i64, i64]
        L.honk_listen_workspace_bytes.restype = i64
        L.honk_listen.argtypes = [vp, vp, vp, C.c_int, i64, i64, i64, i32, C.c_float, vp, vp, vp, vp, i64, vp]
        L.honk_launch_count.restype = i64
        L.honk_model_path.argtypes = [vp]
        L.honk_profile_enable.argtypes = [C.c_int]
        L.honk_profile_get.argtypes = [C.c_char_p, C.POINTER(C.c_double), C.POINTER(i64)]
        L.honk_model_path.restype = C.c_char_p
        L.honk_trainer_create.argtypes = [C.POINTER(ModelConfig), C.POINTER(Tensor), i32, C.c_int, C.POINTER(vp)]
        L.honk_trainer_destroy.argtypes = [vp]
        L.honk_trainer_workspace_bytes.argtypes = [vp, i64]
        L.honk_trainer_workspace_bytes.restype = i64
        L.honk_trainer_step.argtypes = [vp, vp, vp, i64, i32, i32, C.POINTER(Sgd), vp, vp, vp, i64, vp]
        L.honk_trainer_reset_optimizer.argtypes = [vp]
        L.honk_trainer_steps.argtypes = [vp]
        L.honk_trainer_steps.restype = i64
        L.honk_trainer_read.argtypes = [vp, C.c_char_p, vp, i64]
        L.honk_trainer_set_dropout.argtypes = [vp, C.c_float, C.c_uint64]
        L.honk_trainer_set_dropout_masks.argtypes = [vp, C.POINTER(vp), i32]
        if L.honk_abi_version() != 2:
            raise RuntimeError("honk_b200: ABI version mismatch")
        _lib = L
    return _lib


_ERRORS = {-1: ValueError, -2: RuntimeError, -3: RuntimeError, -4: RuntimeError, -5: RuntimeError,
           -6: NotImplementedError}


def check(rc):
    if rc != 0:
        msg = lib().honk_last_error().decode("utf-8", "replace")
        raise _ERRORS.get(rc, RuntimeError)("honk_b200: " + msg)


def require_cuda(t, what):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError("honk_b200: %s must be a CUDA tensor (there is no CPU fallback)" % what)


def stream_ptr(device):
    return torch.cuda.current_stream(device).cuda_stream


def launch_count():
    return int(lib().honk_launch_count())


def pcm_dtype_code(t):
    if t.dtype == torch.int16:
        return PCM_I16
    if t.dtype == torch.float32:
        return PCM_F32
    if t.dtype == torch.float64:
        return PCM_F64
    raise ValueError("honk_b200: PCM must be int16, float32 or float64, got %s" % t.dtype)


def profile_enable(on=True):
    lib().honk_profile_enable(1 if on else 0)


def profile_reset():
    lib().honk_profile_reset()


def profile_get(span):
    """(total milliseconds, span count) recorded for ``span`` since the last reset."""
    ms, n = C.c_double(0), C.c_int64(0)
    check(lib().honk_profile_get(span.encode(), C.byref(ms), C.byref(n)))
    return ms.value, n.value
