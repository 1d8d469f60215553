"""ctypes binding of libkvedge_b200.so (the C ABI in include/kvedge_b200.h).

Loading never falls back to anything: a missing library raises, and every compute entry
returns KV_ENODEVICE on a machine without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkvedge_b200.so")
TOOLS_LIB_PATH = os.path.join(HERE, "libkvedge_b200_tools.so")

STATIC_SCRATCH_BYTES = 8 << 20  # KV_STATIC_SCRATCH_BYTES of include/kvedge_b200.h
KV_OK, KV_EINVAL, KV_ENODEVICE, KV_ECUDA, KV_ENOMEM, KV_ECORRUPT, KV_ETEMPLATE = 0, -1, -2, -3, -4, -5, -6

# name -> (restype, argtypes); every symbol declared in include/kvedge_b200.h
SIGNATURES = {
    "kv_version": (C.c_int, []),
    "kv_last_error": (C.c_char_p, []),
    "kv_device_count": (C.c_int, []),
    "kv_b64enc_len": (C.c_size_t, [C.c_size_t]),
    "kv_b64dec_maxlen": (C.c_size_t, [C.c_size_t]),
    "kv_b64_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64]),
    "kv_b64enc_batch_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "kv_b64enc_plan_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                     C.c_size_t, C.c_void_p]),
    "kv_b64enc_run_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                    C.c_void_p, C.c_size_t, C.c_void_p]),
    "kv_b64enc_batch_dev_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                             C.c_void_p, C.c_void_p]),
    "kv_b64_batch_status_dev": (C.c_int, [C.c_void_p, C.c_void_p]),
    "kv_b64dec_batch_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                      C.c_void_p]),
    "kv_b64enc": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p]),
    "kv_b64enc_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "kv_b64dec": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_size_t),
                            C.POINTER(C.c_size_t)]),
    "kv_b64dec_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p]),
    "kv_chart_load": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p)]),
    "kv_chart_free": (None, [C.c_void_p]),
    "kv_release_plan": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "kv_release_pending_count": (C.c_int, [C.c_void_p]),
    "kv_release_pending_input": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "kv_release_pending_fill": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_size_t]),
    "kv_release_encode_gpu": (C.c_int, [C.c_void_p]),
    "kv_release_manifest": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "kv_release_notes": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "kv_release_free": (None, [C.c_void_p]),
    "kv_helm_template": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_void_p),
                                   C.POINTER(C.c_size_t)]),
    "kv_fleet_template": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_char_p, C.c_void_p,
                                    C.POINTER(C.c_void_p), C.c_void_p]),
    "kv_helm_lint": (C.c_int, [C.c_char_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_void_p),
                               C.POINTER(C.c_size_t)]),
    "kv_free": (None, [C.c_void_p]),
    "kv_set_device": (C.c_int, [C.c_int]),
    "kv_host_pipeline_config": (C.c_int, [C.c_int64, C.c_int]),
    "kv_shutdown": (None, []),
    "kv_pinned_alloc": (C.c_void_p, [C.c_size_t]),
    "kv_pinned_free": (None, [C.c_void_p]),
}


class KvOverrides(C.Structure):
    """struct kv_overrides of include/kvedge_b200.h."""
    _fields_ = [("set", C.POINTER(C.c_char_p)), ("n_set", C.c_int),
                ("set_string", C.POINTER(C.c_char_p)), ("n_set_string", C.c_int),
                ("set_file", C.POINTER(C.c_char_p)), ("n_set_file", C.c_int),
                ("blob_keys", C.POINTER(C.c_char_p)), ("blob_data", C.POINTER(C.c_void_p)),
                ("blob_len", C.POINTER(C.c_size_t)), ("n_blob", C.c_int)]


class KvError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"kvedge_b200 error {code}: {msg}")
        self.code = code


_lib = None
_path = LIB_PATH


def use_tools_build() -> None:
    """tools/ only: load libkvedge_b200_tools.so (sweep shapes and KVEDGE_* knobs compiled in)
    instead of the product library.  Must be called before the first lib()."""
    global _path
    if _lib is not None:
        raise RuntimeError("the library is already loaded")
    from . import build
    build.build_native(tools=True)
    _path = TOOLS_LIB_PATH


def lib() -> C.CDLL:
    """The loaded library; raises if it has not been built (no silent fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_path):
            raise ImportError(
                f"{_path} is missing: run `python -m kvedge_b200.build` "
                "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(_path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != KV_OK:
        raise KvError(rc, lib().kv_last_error().decode("utf-8", "replace"))
