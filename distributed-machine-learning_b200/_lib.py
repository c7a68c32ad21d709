"""ctypes binding of libdml_b200.so (include/dml_b200.h) and its in-tree build.

The library is the product: there is no CPU or PyTorch fallback.  If the shared
object is missing or a call fails, this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libdml_b200.so")
SOURCES = ["dml_api.cu", "kernels.cu", "conv_tc.cu", "stem_tc.cu", "head_static.cu", "graph.cc"]
HEADERS = ["conv_tc.h", "conv_chain.inc", "conv_duo.inc", "conv_pool.inc", "graph.h", "kernels.h", os.path.join("..", "..", "include", "dml_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]

MODEL_IDS = {"ResNet50": 0, "InceptionV3": 1}
FLAG_FP32_CHECK, FLAG_NO_TENSOR_CORE, FLAG_KEEP_ALL = 1, 2, 4


class DmlError(RuntimeError):
    pass


class DmlCancelled(DmlError):
    """dml_cancel stopped the call between two micro-batches (DML_ECANCELLED)."""


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libdml_b200.so next to this file."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    # DML_B200_LEGACY_VARIANTS=1: also compile the superseded kernel variants (impl 1, 3, 4 of dml_conv2d_bf16), the A/B twins
    # the surviving kernels were checked against; the product library does not carry them
    extra = ["-DDML_LEGACY_VARIANTS"] if os.environ.get("DML_B200_LEGACY_VARIANTS") == "1" else []
    cmd = [nvcc] + NVCC_FLAGS + extra + ["-o", LIB_PATH] + [os.path.join(CSRC, f) for f in SOURCES] + ["-ldl"]
    if verbose:
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise DmlError("nvcc failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


_lib = None
_lock = threading.Lock()


def has_legacy_variants() -> bool:
    """True when the loaded library was built with -DDML_LEGACY_VARIANTS (dml_conv2d_bf16 impl 1, 3, 4)."""
    return b"legacy" in lib().dml_version()

_vp, _i, _i64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_size_t
_SIGS = {
    "dml_model_info": [_i, C.POINTER(_i), C.POINTER(_i)],
    "dml_model_num_weights": [_i, C.POINTER(_i), C.POINTER(_i64)],
    "dml_model_weight_info": [_i, _i, C.c_char_p, _i, C.POINTER(_i64), C.POINTER(_i)],
    "dml_preprocess_lut": [_i, C.POINTER(C.c_float)],
    "dml_create": [_i, _i, _i, _i, _i, C.POINTER(_vp)],
    "dml_workspace_bytes": [_vp, C.POINTER(_sz)],
    "dml_weights_bytes": [_vp, C.POINTER(_sz)],
    "dml_bind_workspace": [_vp, _vp, _sz],
    "dml_bind_weights": [_vp, _vp, _sz],
    "dml_load_weights": [_vp, _vp, _sz],
    "dml_weights_ready": [_vp],
    "dml_preprocess": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp],
    "dml_forward": [_vp, _i, _vp, _vp, _vp, _vp],
    "dml_infer": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp],
    "dml_cancel": [_vp],
    "dml_cancel_clear": [_vp],
    "dml_num_tensors": [_vp, C.POINTER(_i)],
    "dml_tensor_info": [_vp, _i, C.c_char_p, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_vp),
                        C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)],
    "dml_input_tensor": [_vp, C.POINTER(_vp), C.POINTER(_i)],
    "dml_input_layout": [_vp, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)],
    "dml_input_rows": [_vp, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)],
    "dml_logits": [_vp, C.POINTER(_vp)],
    "dml_launch_count": [_vp, C.POINTER(_i64)],
    "dml_num_ops": [_vp, C.POINTER(_i)],
    "dml_op_info": [_vp, _i, C.c_char_p, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i64), C.POINTER(_i64),
                    C.POINTER(_i64)],
    "dml_op_schedule": [_vp, _i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i)],
    "dml_profile_forward": [_vp, _i, _vp, _vp, _vp, _i, _vp],
    "dml_conv2d_bf16": [_i, _vp, _vp, _vp, _vp, _vp] + [_i] * 17 + [_vp],
    "dml_chain_bf16": [_vp] * 11 + [_i, _i, _i, _vp],
    "dml_duo_bf16": [_vp] * 9 + [_i] * 8 + [_vp],
    "dml_conv_pool_bf16": [_vp] * 4 + [_i] * 10 + [_vp],
    "dml_head_static": [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp],
    "dml_sync": [_vp],
}
EXPORTS = sorted(list(_SIGS) + ["dml_destroy", "dml_last_error", "dml_version"])


def lib():
    """The loaded library (built on first use if the sources are newer)."""
    global _lib
    with _lock:
        if _lib is None:
            if _stale():
                if os.path.exists("/usr/local/cuda/bin/nvcc") or os.environ.get("NVCC"):
                    build()
                elif not os.path.exists(LIB_PATH):
                    raise DmlError(f"{LIB_PATH} is missing and nvcc is not available to build it")
            l = C.CDLL(LIB_PATH)
            for name, args in _SIGS.items():
                fn = getattr(l, name)
                fn.argtypes = args
                fn.restype = _i
            l.dml_destroy.argtypes = [_vp]
            l.dml_destroy.restype = None
            l.dml_last_error.restype = C.c_char_p
            l.dml_version.restype = C.c_char_p
            _lib = l
    return _lib


def check(rc: int):
    if rc == -5:
        raise DmlCancelled(lib().dml_last_error().decode())
    if rc != 0:
        raise DmlError(f"dml error {rc}: {lib().dml_last_error().decode()}")


def model_id(model_name: str) -> int:
    if model_name not in MODEL_IDS:
        raise DmlError(f"Invalid model {model_name!r}")
    return MODEL_IDS[model_name]


def model_info(model_name: str):
    s, m = _i(), _i()
    check(lib().dml_model_info(model_id(model_name), C.byref(s), C.byref(m)))
    return s.value, ("caffe" if m.value == 0 else "tf")


def weight_spec(model_name: str):
    """[(name, shape)] of the Keras-layout blob, in blob order."""
    l, mid = lib(), model_id(model_name)
    n, tot = _i(), _i64()
    check(l.dml_model_num_weights(mid, C.byref(n), C.byref(tot)))
    out = []
    buf = C.create_string_buffer(128)
    shape = (_i64 * 4)()
    rank = _i()
    for i in range(n.value):
        check(l.dml_model_weight_info(mid, i, buf, 128, shape, C.byref(rank)))
        out.append((buf.value.decode(), tuple(int(shape[d]) for d in range(rank.value))))
    return out


def preprocess_lut(model_name: str):
    import numpy as np
    out = np.empty(768, dtype=np.float32)
    check(lib().dml_preprocess_lut(model_id(model_name), out.ctypes.data_as(C.POINTER(C.c_float))))
    return out.reshape(3, 256)
