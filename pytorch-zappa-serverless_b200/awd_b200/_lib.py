"""ctypes binding of libawd_b200.so (include/awd_b200.h).  Fails loudly when the library is
missing: there is no CPU or eager-PyTorch fallback behind it."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libawd_b200.so')

AWD_OK, AWD_EINVAL, AWD_ECUDA, AWD_ENODEV, AWD_ESTATE, AWD_ENOMEM, AWD_EMISSING, AWD_ETOKEN, AWD_ELIMIT = \
    0, -1, -2, -3, -4, -5, -6, -7, -8
PREC_FP32, PREC_BF16 = 0, 1


class AwdConfig(C.Structure):
    _fields_ = [('vocab_sz', C.c_int32), ('emb_sz', C.c_int32), ('n_hid', C.c_int32), ('n_layers', C.c_int32),
                ('tie_weights', C.c_int32), ('has_decoder_bias', C.c_int32), ('precision', C.c_int32),
                ('device', C.c_int32)]


class AwdSampling(C.Structure):
    _fields_ = [('temperature', C.c_float), ('greedy', C.c_int32), ('n_draws', C.c_int32), ('ban_zero', C.c_int32),
                ('n_banned', C.c_int32), ('_pad', C.c_int32), ('banned', C.POINTER(C.c_int64)), ('seed', C.c_uint64)]


# name -> (restype, argtypes); every symbol declared in include/awd_b200.h
SYMBOLS = {
    'awd_create': (C.c_int, [C.POINTER(AwdConfig), C.POINTER(C.c_void_p)]),
    'awd_destroy': (C.c_int, [C.c_void_p]),
    'awd_load_param': (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int32]),
    'awd_commit_weights': (C.c_int, [C.c_void_p]),
    'awd_reset': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p]),
    'awd_get_state': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'awd_set_state': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'awd_current_bs': (C.c_int32, [C.c_void_p]),
    'awd_gather': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    'awd_forward': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_void_p),
                              C.c_int32, C.c_void_p]),
    'awd_decode': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    'awd_tc_gemm_bf16': (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                  C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    'awd_generate': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(AwdSampling),
                               C.c_void_p, C.c_void_p]),
    'awd_generate_ragged': (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32,
                                     C.POINTER(AwdSampling), C.c_void_p, C.c_void_p]),
    'awd_generate_resume': (C.c_int, [C.c_void_p, C.POINTER(C.c_uint8), C.c_void_p, C.POINTER(C.c_int32), C.c_int32,
                                     C.c_int32, C.c_int32, C.POINTER(AwdSampling), C.c_int64, C.c_void_p, C.c_void_p]),
    'awd_decode_race_shard': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(AwdSampling), C.c_int64, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    'awd_merge_race': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(AwdSampling),
                                C.c_void_p, C.c_void_p]),
    'awd_decode_race_shard_packed': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(AwdSampling), C.c_int64, C.c_void_p, C.c_void_p]),
    'awd_merge_race_packed': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(AwdSampling), C.c_void_p, C.c_void_p]),
    'awd_sample': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(AwdSampling), C.c_int64, C.c_void_p,
                             C.c_void_p]),
    'awd_decode_sample_static': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(AwdSampling), C.c_int64,
                                          C.c_void_p, C.c_void_p]),
    'awd_uniform_from_bits': (C.c_float, [C.c_uint32]),
    'awd_gumbel_from_bits': (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    'awd_check_device_errors': (C.c_int, [C.c_void_p, C.c_void_p]),
    'awd_profile': (C.c_int, [C.c_void_p, C.c_int32]),
    'awd_profile_read': (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    'awd_launch_count': (C.c_int64, []),
    'awd_last_error': (C.c_char_p, []),
    'awd_version': (C.c_char_p, []),
}

_lib = None


class AwdError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def load():
    """dlopen the in-tree library and bind every declared symbol."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f'{LIB_PATH} is missing: build it with `python -m awd_b200.build` (nvcc, sm_100a). '
            'This engine has no CPU fallback.')
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)        # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int):
    if rc == AWD_OK:
        return
    msg = load().awd_last_error().decode('utf-8', 'replace')
    if rc == AWD_ETOKEN:
        raise IndexError(msg)
    if rc in (AWD_EINVAL, AWD_ELIMIT):
        raise ValueError(msg)
    raise AwdError(rc, msg)
