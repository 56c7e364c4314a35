"""ctypes binding of libitmo_b200.so (include/itmo_b200.h).

The library is the product: there is no CPU or PyTorch fallback. Importing this module without the
built library, or calling into it on a machine without an sm_100 GPU, raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_uint, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libitmo_b200.so')

PREC_TF32, PREC_BF16, PREC_FP32, PREC_FP16 = 0, 1, 2, 3
DT_U8, DT_F32 = 0, 1
OUT_U8, OUT_F32, OUT_HDR_INTENDED, OUT_HDR_AS_WRITTEN, OUT_HDR_INTENDED_F16 = 0, 1, 2, 3, 4
NUM_LAYERS = 19
TRAIN_NO_DROPOUT, TRAIN_FREEZE_MOVING, TRAIN_KEEP_GRADS = 1, 2, 4
OPT_RMSPROP, OPT_ADAM = 0, 1
IPC_HANDLE_BYTES = 64
REINHARD_STATIC_MAX_IMAGES = 256   # ITMO_REINHARD_STATIC_MAX_IMAGES


class LayerWeights(ctypes.Structure):
    _fields_ = [(n, POINTER(c_float)) for n in
                ('kernel', 'bias', 'gamma', 'beta', 'moving_mean', 'moving_variance')]


class CameraParams(ctypes.Structure):
    _fields_ = [('v', c_double), ('n', c_double), ('y', c_double)]


class AffineParams(ctypes.Structure):
    _fields_ = [(k, c_double) for k in ('m00', 'm01', 'm10', 'm11', 'off0', 'off1')] + [('flip_h', c_int), ('flip_v', c_int)]


class ItmoError(RuntimeError):
    pass


_lib = None

# every symbol include/itmo_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    'itmo_last_error': (c_char_p, []),
    'itmo_abi_version': (c_int, []),
    'itmo_create': (c_int, [POINTER(c_void_p), c_int, c_int, c_int]),
    'itmo_destroy': (None, [c_void_p]),
    'itmo_load_weights': (c_int, [c_void_p, POINTER(LayerWeights), c_int, c_float]),
    'itmo_forward': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_uint, c_void_p]),
    'itmo_predict_host': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_int, c_uint]),
    'itmo_launch_count': (c_longlong, [c_void_p]),
    'itmo_saturation_count': (c_int, [c_void_p, POINTER(ctypes.c_ulonglong), c_int]),
    'itmo_set_graph_capture': (c_int, [c_void_p, c_int]),
    'itmo_host_alloc': (c_int, [POINTER(c_void_p), c_size_t]),
    'itmo_host_free': (None, [c_void_p]),
    'itmo_set_profiling': (c_int, [c_void_p, c_int]),
    'itmo_layer_times': (c_int, [c_void_p, POINTER(ctypes.c_double), POINTER(c_longlong), c_int]),
    'itmo_read_activation': (c_int, [c_void_p, c_int, c_int, POINTER(c_float), c_size_t, POINTER(c_size_t)]),
    'itmo_tile_axis_plan': (c_int, [c_int, POINTER(c_int), POINTER(c_int), POINTER(c_int), c_int]),
    'itmo_predict_frames': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_double, c_void_p]),
    'itmo_inverse_reinhard': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_double, c_void_p, c_void_p]),
    'itmo_conv3x3_bn': (c_int, [c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    'itmo_convT2x2_bn': (c_int, [c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_int, c_void_p, c_void_p]),
    'itmo_conv3x3_bn_head': (c_int, [c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    'itmo_conv3x3_bn_rgb': (c_int, [c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_int, c_void_p, c_void_p]),
    # training step
    'itmo_train_param_count': (c_longlong, []),
    'itmo_train_bn_count': (c_longlong, []),
    'itmo_train_layout': (c_int, [c_int, c_int, POINTER(c_longlong), POINTER(c_longlong)]),
    'itmo_trainer_create': (c_int, [POINTER(c_void_p), c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    'itmo_trainer_destroy': (None, [c_void_p]),
    'itmo_trainer_sync_weights': (c_int, [c_void_p, c_void_p]),
    'itmo_train_zero_grad': (c_int, [c_void_p, c_void_p]),
    'itmo_train_backward': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_double, c_uint, c_uint,
                                    c_void_p]),
    'itmo_train_dropout_seed': (c_int, [c_uint, c_int, POINTER(c_uint)]),
    'itmo_train_metrics': (c_int, [c_void_p, POINTER(c_double), POINTER(c_double), c_int, c_void_p]),
    'itmo_train_update': (c_int, [c_void_p, c_int, c_float, c_float, c_float, c_float, c_float, c_void_p]),
    'itmo_train_set_step': (c_int, [c_void_p, c_longlong]),
    'itmo_train_get_step': (c_longlong, [c_void_p]),
    'itmo_train_read_activation': (c_int, [c_void_p, c_int, c_int, c_int, POINTER(c_float), c_size_t,
                                           POINTER(c_size_t)]),
    'itmo_train_set_profiling': (c_int, [c_void_p, c_int]),
    'itmo_train_phase_times': (c_int, [c_void_p, POINTER(c_double), POINTER(c_longlong), c_int]),
    'itmo_train_launch_count': (c_longlong, [c_void_p]),
    # training-data synthesis / augmentation
    'itmo_synth_hdr_field': (c_int, [c_void_p, c_int, c_int, c_int, c_uint, c_void_p]),
    'itmo_virtual_camera': (c_int, [c_void_p, c_int, c_int, c_int, POINTER(CameraParams), c_void_p, c_int, c_void_p]),
    'itmo_reinhard_target': (c_int, [c_void_p, c_int, c_int, c_int, c_double, c_void_p, c_int, c_void_p]),
    'itmo_augment': (c_int, [c_void_p, c_int, c_int, c_int, c_int, POINTER(AffineParams), c_void_p, c_float, c_void_p]),
    # allocation-free variant of itmo_reinhard_target (static device scratch; no counterpart in the reference)
    'itmo_reinhard_target_static': (c_int, [c_void_p, c_int, c_int, c_int, c_double, c_void_p, c_int, c_void_p]),
    # fused gradient exchange over NVLink peer memory (SURVEY §8 e, K8)
    'itmo_exchange_create': (c_int, [POINTER(c_void_p), c_int, c_void_p]),
    'itmo_exchange_grads': (c_void_p, [c_void_p]),
    'itmo_exchange_params': (c_void_p, [c_void_p]),
    'itmo_exchange_open': (c_int, [c_void_p, c_int, c_int, c_void_p]),
    'itmo_exchange_step': (c_int, [c_void_p, c_void_p, c_int, c_float, c_float, c_float, c_float, c_float, c_void_p]),
    'itmo_exchange_destroy': (None, [c_void_p]),
    'itmo_wgrad3x3': (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    'itmo_dgrad3x3': (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
}


def load():
    """Load libitmo_b200.so (once) and set the prototypes. Raises ItmoError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ItmoError(
            f'{LIB_PATH} is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
            '(nvcc, sm_100a). There is no CPU fallback.')
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        msg = load().itmo_last_error()
        raise ItmoError(f'libitmo_b200 error {rc}: {msg.decode() if msg else "?"}')
