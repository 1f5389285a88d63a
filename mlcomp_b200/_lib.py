"""ctypes binding of libmlcomp_b200.so (C ABI: include/mlc.h).

There is NO CPU fallback: if the shared library is missing and cannot be
built (no nvcc), importing an op raises.  The library itself is torch-free;
torch only supplies device memory and the current stream.
"""
import ctypes
import os
from ctypes import (POINTER, c_char_p, c_double, c_float, c_int, c_int32,
                    c_int64, c_size_t, c_void_p)

from . import build as _build

MLC_OK = 0
MLC_ERR_ARG, MLC_ERR_CUDA, MLC_ERR_UNSUPPORTED, MLC_ERR_CAPACITY = -1, -2, -3, -4
MLC_F16, MLC_F32, MLC_U8 = 0, 1, 2
MLC_ACT_NONE, MLC_ACT_SIGMOID, MLC_ACT_SOFTMAX = 0, 1, 2
MLC_XF_HFLIP, MLC_XF_VFLIP, MLC_XF_TRANSPOSE = 1, 2, 4
MLC_MAX_SOURCES, MLC_MAX_THR, MLC_MAX_AREAS = 32, 255, 8
MLC_MAX_CLASSES_SEG = 64
MLC_LUT_BYTES = 69632

# every symbol include/mlc.h declares: (restype, argtypes)
SIGNATURES = {
    'mlc_version': (c_int, []),
    'mlc_last_error': (c_char_p, []),
    'mlc_confusion': (c_int, [c_void_p, c_int, c_int64, c_int, c_void_p,
                              c_void_p, c_void_p, c_void_p, c_void_p]),
    'mlc_seg_lut_build': (c_int, [POINTER(c_float), c_int, c_int, c_void_p,
                                  c_void_p]),
    'mlc_seg_hist': (c_int, [POINTER(c_void_p), POINTER(c_int32), c_int, c_int,
                             c_void_p, c_int64, c_int, c_int, c_int,
                             POINTER(c_float), c_int, c_int, c_void_p,
                             c_void_p, c_void_p]),
    'mlc_seg_hist_bins': (c_int, [POINTER(c_void_p), POINTER(c_int32), c_int,
                                  c_int, c_void_p, c_int64, c_int, c_int, c_int,
                                  POINTER(c_float), c_int, c_int, c_void_p,
                                  c_void_p, c_void_p]),
    'mlc_seg_prob': (c_int, [POINTER(c_void_p), POINTER(c_int32), c_int, c_int,
                             c_int64, c_int, c_int, c_int, c_int, c_void_p,
                             c_void_p]),
    'mlc_seg_argmax': (c_int, [POINTER(c_void_p), POINTER(c_int32), c_int,
                               c_int, c_int64, c_int, c_int, c_int, c_int,
                               c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_void_p]),
    'mlc_dice_search_workspace': (c_size_t, [c_int64, c_int, c_int, c_int]),
    'mlc_dice_search': (c_int, [c_void_p, c_int64, c_int, c_int,
                                POINTER(c_int32), c_int, c_int, c_double,
                                c_void_p, c_void_p, c_void_p, c_size_t,
                                c_void_p]),
    'mlc_dice_iou_search': (c_int, [c_void_p, c_int64, c_int, c_int,
                                    POINTER(c_int32), c_int, c_int, c_double,
                                    c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_size_t, c_void_p]),
    'mlc_dice_search_static': (c_int, [c_void_p, c_int64, c_int, c_int,
                                       POINTER(c_int32), c_int, c_int, c_double,
                                       c_void_p, c_void_p, c_void_p]),
    'mlc_combine_ranks': (c_int, [c_void_p, c_int, c_int64, c_int64, c_void_p,
                                  c_void_p]),
    'mlc_plane_dice': (c_int, [c_void_p, c_int64, c_int, c_int,
                               POINTER(c_int32), POINTER(c_int32), c_int,
                               c_double, c_void_p, c_void_p]),
    'mlc_rle_encode': (c_int, [POINTER(c_void_p), POINTER(c_int32), c_int,
                               c_int, c_int64, c_int, c_int, c_int,
                               POINTER(c_float), POINTER(c_int32), c_int,
                               c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p]),
    'mlc_rle_encode_known': (c_int, [POINTER(c_void_p), POINTER(c_int32), c_int,
                                     c_int, c_int64, c_int, c_int, c_int,
                                     POINTER(c_float), POINTER(c_int32), c_int,
                                     c_void_p, c_void_p, c_int64, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p]),
    'mlc_plane_area': (c_int, [c_void_p, c_int64, c_int, c_int,
                               POINTER(c_int32), c_void_p, c_void_p]),
    'mlc_rle_format': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p,
                               c_int64, c_void_p, c_void_p, c_void_p,
                               c_void_p]),
    'mlc_rle_decode': (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int,
                               c_int, c_void_p, c_void_p]),
    'mlc_mask_composite': (c_int, [c_void_p, c_int64, c_int, c_int, c_int,
                                   c_void_p, c_void_p, c_void_p]),
    'mlc_mask_outline': (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int,
                                 c_int, c_void_p, c_void_p]),
}

_lib = None


class MlcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f'libmlcomp_b200 error {code}: {message}')
        self.code = code


def lib_path():
    """The library to load: MLC_LIB (a variant built by
    ``python -m mlcomp_b200.build --bounds-check`` etc.) or the default."""
    return os.environ.get('MLC_LIB') or _build.LIB


def load():
    """Load the library; the default one is (re)built first when nvcc is
    available and a source is newer than it.  Without nvcc an existing
    library is used as it is (with a warning if it is older than the
    sources); a missing one is an ImportError - there is no CPU fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if path == _build.LIB:
        try:
            have_nvcc = bool(_build.nvcc_path())
        except RuntimeError:
            have_nvcc = False
        if have_nvcc:
            try:
                _build.build()      # returns at once when nothing is stale
            except Exception as e:
                if not os.path.exists(path):
                    raise ImportError(
                        f'{path} is missing and could not be built ({e}); '
                        'mlcomp_b200 has no CPU fallback - run '
                        '`python -m mlcomp_b200.build` where nvcc is available'
                    ) from e
                raise
        elif os.path.exists(path) and _build.needs_build():
            import warnings
            warnings.warn(f'{path} is older than its sources and nvcc is not '
                          'available to rebuild it')
    if not os.path.exists(path):
        raise ImportError(f'{path} is missing; mlcomp_b200 has no CPU '
                          'fallback - run `python -m mlcomp_b200.build`')
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if a symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != MLC_OK:
        msg = load().mlc_last_error()
        raise MlcError(rc, msg.decode() if msg else '')
