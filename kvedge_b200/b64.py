"""Host-side mirror of Sprig's ``b64enc`` / ``b64dec`` for the reference's two call sites
(``deployment/helm/templates/aziot-edge-runtime-config-secret.yaml:6`` and
``deployment/helm/templates/aziot-edge-vm-cloud-init-secret.yaml:6``), running on the GPU
through the C ABI of ``libkvedge_b200.so``.

Same names and argument meaning as the template functions: ``b64enc(s) -> s`` is total
(never fails on any byte string); ``b64dec`` reports malformed input the way Go's
``base64.CorruptInputError`` does.  torch is used only to own device memory and streams
for the device-resident entry points.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import KvError, check


class CorruptInputError(ValueError):
    """Go ``base64.CorruptInputError``: ``offset`` is the offending input byte."""

    def __init__(self, offset: int):
        super().__init__(f"illegal base64 data at input byte {offset}")
        self.offset = offset


def encoded_len(n: int) -> int:
    return (n + 2) // 3 * 4


def encode_offsets(in_off: np.ndarray, align: int = 16) -> np.ndarray:
    """Output offsets for a packed batch: unit i gets 4*ceil(len_i/3) bytes, every unit start
    rounded up to ``align`` (16 keeps the kernel on its 16-byte store path)."""
    lens = np.diff(np.asarray(in_off, dtype=np.int64))
    olen = (lens + 2) // 3 * 4
    if align > 1:
        olen = (olen + align - 1) // align * align
    out = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(olen, out=out[1:])
    return out


def decode_offsets(in_off: np.ndarray, align: int = 4) -> np.ndarray:
    lens = np.diff(np.asarray(in_off, dtype=np.int64))
    olen = lens // 4 * 3
    if align > 1:
        olen = (olen + align - 1) // align * align
    out = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(olen, out=out[1:])
    return out


# ------------------------------------------------------------------------- host buffers
def b64enc(data: bytes) -> bytes:
    """Sprig ``b64enc``: standard base64 of ``data`` (GPU; host buffers in and out)."""
    data = bytes(data)
    n = len(data)
    out = C.create_string_buffer(max(encoded_len(n), 1))
    check(_lib.lib().kv_b64enc(data, n, out))
    return out.raw[:encoded_len(n)]


def b64dec(text: bytes) -> bytes:
    """Go ``StdEncoding.DecodeString`` (what Kubernetes applies to ``Secret.data``)."""
    text = bytes(text)
    n = len(text)
    out = C.create_string_buffer(max(n // 4 * 3, 1) + 4)
    out_n = C.c_size_t(0)
    err = C.c_size_t(0)
    rc = _lib.lib().kv_b64dec(text, n, out, C.byref(out_n), C.byref(err))
    if rc == _lib.KV_ECORRUPT:
        raise CorruptInputError(err.value)
    check(rc)
    return out.raw[:out_n.value]


def b64enc_packed(blob: np.ndarray, in_off: np.ndarray, out_off: Optional[np.ndarray] = None,
                  out: Optional[np.ndarray] = None) -> Tuple[np.ndarray, np.ndarray]:
    """Encode a packed batch held in host memory; returns (out blob, out_off)."""
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    in_off = np.ascontiguousarray(in_off, dtype=np.int64)
    n = len(in_off) - 1
    if out_off is None:
        out_off = encode_offsets(in_off)
    out_off = np.ascontiguousarray(out_off, dtype=np.int64)
    if out is None:
        out = np.zeros(max(int(out_off[-1]), 1), dtype=np.uint8)
    check(_lib.lib().kv_b64enc_batch(blob.ctypes.data, in_off.ctypes.data, n, out.ctypes.data,
                                     out_off.ctypes.data))
    return out, out_off


def b64dec_packed(blob: np.ndarray, in_off: np.ndarray, out_off: Optional[np.ndarray] = None):
    """Decode a packed batch in host memory; returns (out blob, out_off, out_len, status) with
    status[i] == -1 for a clean unit, else Go's CorruptInputError offset."""
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    in_off = np.ascontiguousarray(in_off, dtype=np.int64)
    n = len(in_off) - 1
    if out_off is None:
        out_off = decode_offsets(in_off)
    out_off = np.ascontiguousarray(out_off, dtype=np.int64)
    out = np.zeros(max(int(out_off[-1]), 1) + 4, dtype=np.uint8)
    out_len = np.zeros(max(n, 1), dtype=np.int64)
    status = np.zeros(max(n, 1), dtype=np.int64)
    check(_lib.lib().kv_b64dec_batch(blob.ctypes.data, in_off.ctypes.data, n, out.ctypes.data,
                                     out_off.ctypes.data, out_len.ctypes.data, status.ctypes.data))
    return out, out_off, out_len[:n], status[:n]


def b64enc_many(units: Sequence[bytes]) -> List[bytes]:
    """``[b64enc(u) for u in units]`` in one batched GPU pass."""
    lens = np.fromiter((len(u) for u in units), dtype=np.int64, count=len(units))
    in_off = np.zeros(len(units) + 1, dtype=np.int64)
    np.cumsum(lens, out=in_off[1:])
    blob = np.frombuffer(b"".join(units), dtype=np.uint8) if in_off[-1] else np.zeros(0, np.uint8)
    out, out_off = b64enc_packed(blob, in_off)
    raw = out.tobytes()
    return [raw[int(out_off[i]):int(out_off[i]) + encoded_len(int(lens[i]))] for i in range(len(units))]


def b64dec_many(units: Sequence[bytes]) -> List[bytes]:
    lens = np.fromiter((len(u) for u in units), dtype=np.int64, count=len(units))
    in_off = np.zeros(len(units) + 1, dtype=np.int64)
    np.cumsum(lens, out=in_off[1:])
    blob = np.frombuffer(b"".join(units), dtype=np.uint8) if in_off[-1] else np.zeros(0, np.uint8)
    out, out_off, out_len, status = b64dec_packed(blob, in_off)
    bad = np.nonzero(status >= 0)[0]
    if len(bad):
        raise CorruptInputError(int(status[bad[0]]))
    raw = out.tobytes()
    return [raw[int(out_off[i]):int(out_off[i]) + int(out_len[i])] for i in range(len(units))]


# ------------------------------------------------------------------------- device resident
def scratch_bytes(n_units: int, total_in_bytes: int) -> int:
    return int(_lib.lib().kv_b64_scratch_bytes(n_units, total_in_bytes))


def _stream_handle(stream) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def b64enc_device(d_in, d_in_off, d_out, d_out_off, d_scratch, total_in_bytes: Optional[int] = None,
                  stream=None) -> None:
    """Encode a packed batch that already lives in HBM (torch uint8/int64 CUDA tensors).
    Asynchronous on ``stream`` (default: torch's current stream)."""
    n = d_in_off.numel() - 1
    if total_in_bytes is None:
        total_in_bytes = d_in.numel()
    check(_lib.lib().kv_b64enc_batch_dev(d_in.data_ptr(), d_in_off.data_ptr(), n, total_in_bytes,
                                         d_out.data_ptr(), d_out_off.data_ptr(),
                                         d_scratch.data_ptr(), d_scratch.numel() * d_scratch.element_size(),
                                         _stream_handle(stream)))


def b64enc_device_static(d_in, d_in_off, d_out, d_out_off, total_in_bytes: Optional[int] = None,
                         stream=None) -> None:
    """``b64enc_device`` without a scratch argument, for callers that cannot allocate device
    memory (HBM fully owned by another pool, allocation-free hot loops, pointer-only plugins).
    The tile table goes into a fixed ``_lib.STATIC_SCRATCH_BYTES`` array inside the library;
    batches whose table would not fit raise ``KvError`` (KV_ENOMEM).  Calls on one GPU share
    that array, so do not let two of them overlap; make the first call on a device outside
    any stream capture.  Asynchronous on ``stream``."""
    n = d_in_off.numel() - 1
    if total_in_bytes is None:
        total_in_bytes = d_in.numel()
    check(_lib.lib().kv_b64enc_batch_dev_static(d_in.data_ptr(), d_in_off.data_ptr(), n, total_in_bytes,
                                                d_out.data_ptr(), d_out_off.data_ptr(),
                                                _stream_handle(stream)))


def b64enc_plan_device(d_in_off, d_out_off, d_scratch, total_in_bytes: int, stream=None) -> None:
    """Build the tile table for a packed batch (reusable while the offsets do not change)."""
    n = d_in_off.numel() - 1
    check(_lib.lib().kv_b64enc_plan_dev(d_in_off.data_ptr(), n, total_in_bytes, d_out_off.data_ptr(),
                                        d_scratch.data_ptr(), d_scratch.numel() * d_scratch.element_size(),
                                        _stream_handle(stream)))


def b64enc_run_device(d_in, d_in_off, d_out, d_scratch, total_in_bytes: int, stream=None) -> None:
    """Run the encode kernel over a tile table built by ``b64enc_plan_device``."""
    n = d_in_off.numel() - 1
    check(_lib.lib().kv_b64enc_run_dev(d_in.data_ptr(), d_in_off.data_ptr(), n, total_in_bytes,
                                       d_out.data_ptr(), d_scratch.data_ptr(),
                                       d_scratch.numel() * d_scratch.element_size(),
                                       _stream_handle(stream)))


def b64dec_device(d_in, d_in_off, d_out, d_out_off, d_out_len, d_status, d_scratch,
                  total_in_bytes: Optional[int] = None, stream=None) -> None:
    """Decode a packed batch resident in HBM. Asynchronous on ``stream``."""
    n = d_in_off.numel() - 1
    if total_in_bytes is None:
        total_in_bytes = d_in.numel()
    check(_lib.lib().kv_b64dec_batch_dev(d_in.data_ptr(), d_in_off.data_ptr(), n, total_in_bytes,
                                         d_out.data_ptr(), d_out_off.data_ptr(),
                                         d_out_len.data_ptr(), d_status.data_ptr(),
                                         d_scratch.data_ptr(), d_scratch.numel() * d_scratch.element_size(),
                                         _stream_handle(stream)))


__all__ = ["b64enc", "b64dec", "b64enc_many", "b64dec_many", "b64enc_packed", "b64dec_packed",
           "b64enc_device", "b64enc_device_static", "b64dec_device", "b64enc_plan_device", "b64enc_run_device", "scratch_bytes", "encode_offsets", "decode_offsets",
           "encoded_len", "CorruptInputError", "KvError"]


def pinned_empty(n: int) -> np.ndarray:
    """``n`` bytes of page-locked host memory as a uint8 array (``kv_pinned_alloc``); buffers
    handed to the host-buffer entry points should come from here, see include/kvedge_b200.h."""
    import ctypes as C
    import weakref
    L = _lib.lib()
    p = L.kv_pinned_alloc(max(int(n), 1))
    if not p:
        raise MemoryError(f"kv_pinned_alloc({n}): {L.kv_last_error().decode('utf-8', 'replace')}")
    arr = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(int(n), 1),))[:int(n)]
    weakref.finalize(arr.base if arr.base is not None else arr, L.kv_pinned_free, p)
    return arr
