"""ctypes binding of the C ABI declared in ``include/afe.h``.

``load()`` returns the process-wide binding of the PRODUCT library
``audio_processor_b200/libafe.so`` (nvcc, sm_100a) and raises ``AfeLibraryError`` when
it is missing or is not the CUDA build: there is no CPU fallback.  ``Library(path)`` binds
an explicit path; only the test-suite uses it (for the host emulation of the kernels).
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libafe.so")

AFE_OK = 0
AFE_ERR_INVALID = -1
AFE_ERR_CUDA = -2
AFE_ERR_WORKSPACE = -3
AFE_ERR_UNSUPPORTED = -4
PRECISION_F64 = 0
PRECISION_F32 = 1
FBANK_RAW = 1
FBANK_STATIC_SCRATCH_BYTES = 8 << 20
FBANK_STATIC_MAX_BATCH = 256
K_COUNT = 9

c_i64p = C.POINTER(C.c_int64)
c_i32p = C.POINTER(C.c_int32)
c_f32p = C.POINTER(C.c_float)
c_f64p = C.POINTER(C.c_double)
c_intp = C.POINTER(C.c_int)


class AfeLibraryError(RuntimeError):
    """The native CUDA library is missing or unusable (the package has no fallback)."""


class AfeError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"afe error {code}: {message}")
        self.code = code
        self.message = message


class WavInfo(C.Structure):
    """``afe_wav_info`` of include/afe.h."""
    _fields_ = [("format", C.c_int32), ("channels", C.c_int32), ("sample_rate", C.c_int32),
                ("bits_per_sample", C.c_int32), ("block_align", C.c_int32), ("reserved", C.c_int32),
                ("data_offset", C.c_int64), ("data_bytes", C.c_int64), ("num_frames", C.c_int64)]


WAV_U8, WAV_S16, WAV_S24, WAV_S32, WAV_F32, WAV_F64 = 1, 2, 3, 4, 5, 6

# every symbol include/afe.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "afe_version": (C.c_char_p, []),
    "afe_backend": (C.c_int, []),
    "afe_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "afe_destroy": (None, [C.c_void_p]),
    "afe_last_error": (C.c_char_p, []),
    "afe_num_sms": (C.c_int, [C.c_void_p]),
    "afe_resampled_length": (C.c_int64, [C.c_int64, C.c_int, C.c_int]),
    "afe_fbank_num_frames": (C.c_int64, [C.c_int64]),
    "afe_fbank_padded_frames": (C.c_int64, [c_i64p, C.c_int]),
    "afe_chunk_table": (C.c_int, [C.c_int64, C.c_int64, C.c_int64, c_i64p, c_i64p, C.c_int, c_intp]),
    "afe_segment_table": (C.c_int, [C.c_int64, C.c_int, c_f64p, C.c_int, C.c_double, C.c_double, C.c_double,
                                    c_i64p, c_i64p, c_f64p, c_f64p, C.c_int, c_intp]),
    "afe_resample_taps": (C.c_int, [C.c_int, C.c_int, c_f32p, C.c_size_t, c_intp, c_intp, c_intp, c_intp]),
    "afe_resample": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_int,
                               C.c_void_p, C.c_void_p, C.c_int64]),
    "afe_resample_batch_workspace_bytes": (C.c_size_t, [C.c_int]),
    "afe_resample_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_size_t]),
    "afe_resample_batch_pcm16": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                           C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p,
                                           C.c_size_t]),
    "afe_pcm16_to_float": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "afe_wav_parse": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int64, C.c_void_p]),
    "afe_wav_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_void_p]),
    "afe_resample_batch_wav16": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                           C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p,
                                           C.c_size_t]),
    "afe_mix_stereo": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "afe_fbank_layout": (C.c_int, [c_i64p, C.c_int, C.c_int, c_i64p, c_i64p]),
    "afe_fbank_workspace_bytes": (C.c_size_t, [c_i64p, C.c_int]),
    "afe_fbank_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                  C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "afe_fbank_batch_static": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                         C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "afe_fbank_batch_placed": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, c_i64p, c_i64p,
                                         C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "afe_acoustic_batch_placed": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                            c_i32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "afe_acoustic_batch_scaled": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                            c_i32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                            C.c_void_p, C.c_size_t]),
    "afe_profile_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "afe_profile_reset": (C.c_int, [C.c_void_p]),
    "afe_profile_read": (C.c_int, [C.c_void_p, c_f64p, c_i64p]),
    "afe_kernel_name": (C.c_char_p, [C.c_int]),
    "afe_acoustic_workspace_bytes": (C.c_size_t, [c_i64p, C.c_int]),
    "afe_acoustic_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), c_i64p, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
}


class Library:
    """A bound libafe shared object."""

    def __init__(self, path: str):
        if not os.path.exists(path):
            raise AfeLibraryError(
                f"{path} not found: build it with `python -m audio_processor_b200.csrc.build` "
                "(nvcc, sm_100a). audio_processor_b200 has no CPU fallback.")
        self.path = path
        self.dll = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(self.dll, name)
            except AttributeError as e:
                raise AfeLibraryError(f"{path} does not export {name}") from e
            fn.restype = res
            fn.argtypes = args
            setattr(self, name, fn)
        self.backend = int(self.afe_backend())

    def check(self, code: int) -> None:
        if code != AFE_OK:
            msg = self.afe_last_error()
            raise AfeError(code, msg.decode() if msg else "")

    def create(self, device: int) -> C.c_void_p:
        ctx = C.c_void_p()
        self.check(self.afe_create(int(device), C.byref(ctx)))
        return ctx


_lock = threading.Lock()
_lib: Optional[Library] = None


def load() -> Library:
    """The product library; raises ``AfeLibraryError`` unless it is the CUDA build."""
    global _lib
    with _lock:
        if _lib is None:
            lib = Library(LIB_PATH)
            if lib.backend != 1:
                raise AfeLibraryError(f"{LIB_PATH} is not the CUDA sm_100a build (backend={lib.backend})")
            _lib = lib
        return _lib


def i64_array(values):
    """ctypes ``int64[n]`` over a numpy buffer (a ctypes splat of a Python list costs 4x as much; these arrays carry
    thousands of chunk lengths / rows per launch and sit on the host's critical path)."""
    import numpy as np
    buf = np.array(values, dtype=np.int64).reshape(-1) if len(values) else np.zeros(1, np.int64)
    return (C.c_int64 * max(len(values), 1)).from_buffer(buf)


def ptr_array(values):
    """ctypes ``void*[n]`` of device / host addresses given as Python ints (0 = NULL)."""
    import numpy as np
    buf = np.array(values, dtype=np.uint64).reshape(-1) if len(values) else np.zeros(1, np.uint64)
    return (C.c_void_p * max(len(values), 1)).from_buffer(buf)
