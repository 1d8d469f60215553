"""ctypes front-end of the CPU checker in ``b64_ref.c``.

TEST INFRASTRUCTURE ONLY: imported by ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs, never by
``kvedge_b200`` (the product path has no CPU fallback).

Follows the reference call sites
``deployment/helm/templates/aziot-edge-runtime-config-secret.yaml:6`` and
``deployment/helm/templates/aziot-edge-vm-cloud-init-secret.yaml:6`` (Sprig
``b64enc`` -> Go ``encoding/base64.StdEncoding``; third-party, not vendored in
the reference, version unpinned).  Parity unpinned by the reference's own
artefacts; pinned to RFC 4648 vectors, coreutils and binascii in
``tests/test_oracle_b64.py``.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_b64.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile ``b64_ref.c`` with gcc (``make -C oracle``)."""
    srcs = [os.path.join(_HERE, f) for f in ("b64_ref.c", "b64_pool.c", "fleet_cpu.c", "Makefile")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle_b64.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        c = ctypes
        L.oracle_b64_encoded_len.restype = c.c_size_t
        L.oracle_b64_encoded_len.argtypes = [c.c_size_t]
        L.oracle_b64_encode.restype = c.c_size_t
        L.oracle_b64_encode.argtypes = [c.c_void_p, c.c_size_t, c.c_void_p]
        L.oracle_b64_decode.restype = c.c_int
        L.oracle_b64_decode.argtypes = [c.c_void_p, c.c_size_t, c.c_void_p,
                                        c.POINTER(c.c_size_t), c.POINTER(c.c_size_t)]
        L.oracle_b64_encode_batch.restype = None
        L.oracle_b64_encode_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int64,
                                              c.c_void_p, c.c_void_p]
        L.oracle_b64_decode_batch.restype = None
        L.oracle_b64_decode_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_int64, c.c_void_p,
                                              c.c_void_p, c.c_void_p, c.c_void_p]
        L.oracle_pool_create.restype = c.c_void_p
        L.oracle_pool_create.argtypes = [c.c_int]
        L.oracle_pool_encode_batch.restype = None
        L.oracle_pool_encode_batch.argtypes = [c.c_void_p, c.c_void_p, c.c_void_p, c.c_int64,
                                               c.c_void_p, c.c_void_p]
        L.oracle_pool_destroy.restype = None
        L.oracle_pool_destroy.argtypes = [c.c_void_p]
        L.oracle_fleet_render_cpu.restype = c.c_int
        L.oracle_fleet_render_cpu.argtypes = [c.c_void_p, c.c_void_p, c.c_int, c.c_void_p, c.c_char_p, c.c_void_p,
                                              c.c_int, c.POINTER(c.c_void_p), c.c_void_p]
        L.oracle_free.restype = None
        L.oracle_free.argtypes = [c.c_void_p]
        _lib = L
    return _lib


class EncodePool:
    """Persistent C thread pool over the checker's encoder (b64_pool.c): what bench.py's CPU legs
    time.  Threads are created once; a step only wakes them; the batch is cut by input bytes."""

    def __init__(self, threads: int):
        self.threads = max(1, int(threads))
        self._h = lib().oracle_pool_create(self.threads)
        if not self._h:
            raise MemoryError("oracle_pool_create")

    def encode_batch(self, blob: np.ndarray, in_off: np.ndarray, out_off: np.ndarray, out: np.ndarray) -> None:
        assert blob.dtype == np.uint8 and in_off.dtype == np.int64 and out_off.dtype == np.int64
        assert blob.flags.c_contiguous and in_off.flags.c_contiguous and out_off.flags.c_contiguous
        lib().oracle_pool_encode_batch(self._h, blob.ctypes.data, in_off.ctypes.data, len(in_off) - 1,
                                       out.ctypes.data, out_off.ctypes.data)

    def close(self) -> None:
        if self._h:
            lib().oracle_pool_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def encoded_len(n: int) -> int:
    return (n + 2) // 3 * 4


def encode(data: bytes) -> bytes:
    """Go ``base64.StdEncoding.EncodeToString`` of ``data``."""
    data = bytes(data)
    out = ctypes.create_string_buffer(encoded_len(len(data)) or 1)
    n = lib().oracle_b64_encode(data, len(data), out)
    return out.raw[:n]


class CorruptInputError(ValueError):
    """Go ``base64.CorruptInputError``; ``offset`` is the offending input byte."""

    def __init__(self, offset: int, partial: bytes = b""):
        super().__init__(f"illegal base64 data at input byte {offset}")
        self.offset = offset
        self.partial = partial


def decode(data: bytes) -> bytes:
    """Go ``base64.StdEncoding.DecodeString`` of ``data`` (raises on corrupt input)."""
    data = bytes(data)
    out = ctypes.create_string_buffer(len(data) // 4 * 3 + 4)
    n = ctypes.c_size_t(0)
    eo = ctypes.c_size_t(0)
    rc = lib().oracle_b64_decode(data, len(data), out, ctypes.byref(n), ctypes.byref(eo))
    if rc:
        raise CorruptInputError(eo.value, out.raw[:n.value])
    return out.raw[:n.value]


def encode_offsets(in_off: np.ndarray, align: int = 1) -> np.ndarray:
    """Output offsets (n+1) of a packed batch: unit i holds 4*ceil(len_i/3) chars,
    each unit start rounded up to ``align`` bytes."""
    lens = np.diff(in_off.astype(np.int64))
    olen = (lens + 2) // 3 * 4
    if align > 1:
        olen = (olen + align - 1) // align * align
    out = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(olen, out=out[1:])
    return out


def encode_batch(blob: np.ndarray, in_off: np.ndarray, out_off: np.ndarray | None = None,
                 threads: int = 1, out: np.ndarray | None = None) -> np.ndarray:
    """Encode every unit of a packed batch; returns the packed output blob (uint8)."""
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    in_off = np.ascontiguousarray(in_off, dtype=np.int64)
    n = len(in_off) - 1
    if out_off is None:
        out_off = encode_offsets(in_off)
    out_off = np.ascontiguousarray(out_off, dtype=np.int64)
    assert len(out_off) == n + 1, "out_off holds n+1 entries (last = total capacity)"
    total = int(out_off[-1])
    if out is None:
        out = np.zeros(max(total, 1), dtype=np.uint8)
    L = lib()
    if threads <= 1 or n < 2 * threads:
        L.oracle_b64_encode_batch(blob.ctypes.data, in_off.ctypes.data, n,
                                  out.ctypes.data, out_off.ctypes.data)
    else:
        bounds = np.linspace(0, n, threads + 1).astype(np.int64)

        def work(k):
            lo, hi = int(bounds[k]), int(bounds[k + 1])
            if hi > lo:
                L.oracle_b64_encode_batch(blob.ctypes.data, in_off[lo:].ctypes.data, hi - lo,
                                          out.ctypes.data, out_off[lo:].ctypes.data)

        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(work, range(threads)))
    return out[:total]


def decode_batch(blob: np.ndarray, in_off: np.ndarray, out_off: np.ndarray):
    """Decode every unit; returns (out blob, out_len[n], status[n]) with status -1 = ok,
    else Go's CorruptInputError offset."""
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    in_off = np.ascontiguousarray(in_off, dtype=np.int64)
    out_off = np.ascontiguousarray(out_off, dtype=np.int64)
    n = len(in_off) - 1
    assert len(out_off) == n + 1
    cap = int(out_off[-1])
    out = np.zeros(max(cap, 1) + 4, dtype=np.uint8)
    out_len = np.zeros(max(n, 1), dtype=np.int64)
    status = np.zeros(max(n, 1), dtype=np.int64)
    lib().oracle_b64_decode_batch(blob.ctypes.data, in_off.ctypes.data, n, out.ctypes.data,
                                  out_off.ctypes.data, out_len.ctypes.data, status.ctypes.data)
    return out[:cap], out_len[:n], status[:n]


def fleet_render_cpu(product_lib, chart_handle, n: int, names_arr, namespace: bytes, overrides_arr, threads: int):
    """bench.py's CPU twin of kv_fleet_template (fleet_cpu.c): the product's host engine plans every
    install, this checker's scalar encoder fills the pending b64enc sites, all on `threads` host
    threads.  ``product_lib`` is the loaded libkvedge_b200.so (its kv_release_* entries are handed
    over as function pointers).  Returns (stream bytes object view, offsets)."""
    c = ctypes
    api = (c.c_void_p * 6)(*[c.cast(getattr(product_lib, nm), c.c_void_p).value for nm in (
        "kv_release_plan", "kv_release_pending_count", "kv_release_pending_input", "kv_release_pending_fill",
        "kv_release_manifest", "kv_release_free")])
    out = c.c_void_p()
    off = np.zeros(n + 1, dtype=np.int64)
    rc = lib().oracle_fleet_render_cpu(c.cast(api, c.c_void_p), chart_handle, n, c.cast(names_arr, c.c_void_p), namespace,
                                       c.cast(overrides_arr, c.c_void_p), threads, c.byref(out), off.ctypes.data)
    if rc:
        raise RuntimeError("oracle_fleet_render_cpu failed")
    return out, off
