"""Host-side mirror of the Helm operations the reference documents for its chart
(``README.md:56-61``: ``helm install --generate-name . --set ... --set-file
azIotEdgeConfig=config.toml``; BASELINE.json configs 1-2: ``helm lint`` / ``helm template``).

Same vocabulary as the Helm CLI -- ``template``, ``lint``, ``--set``, ``--set-string``,
``--set-file``, release name, namespace -- and the same error behaviour: a template error
(for example ``eq`` on a string against a bool, ``aziot-edge-vm-service.yaml:1``) raises
instead of rendering.  Templates execute in C++ behind the C ABI; every ``b64enc`` runs on
the GPU (there is no CPU fallback, so rendering a chart that reaches a non-empty ``b64enc``
without a CUDA device raises).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import KvError, KvOverrides, check

DEFAULT_RELEASE = "release-name"


class TemplateError(KvError):
    """A chart load / parse / execution error (Helm would exit non-zero)."""


def _raise(rc: int):
    msg = _lib.lib().kv_last_error().decode("utf-8", "replace")
    if rc == _lib.KV_ETEMPLATE:
        raise TemplateError(rc, msg)
    raise KvError(rc, msg)


class _Overrides:
    """Keeps the ctypes arrays behind a kv_overrides struct alive."""

    def __init__(self, set: Iterable[str] = (), set_string: Iterable[str] = (), set_file: Iterable[str] = (),
                 blobs: Optional[Dict[str, bytes]] = None):
        def arr(items):
            b = [s.encode() if isinstance(s, str) else bytes(s) for s in items]
            return (C.c_char_p * max(len(b), 1))(*b), len(b)

        self._set, n_set = arr(list(set))
        self._ss, n_ss = arr(list(set_string))
        self._sf, n_sf = arr(list(set_file))
        blobs = blobs or {}
        self._bk, n_b = arr(list(blobs.keys()))
        self._bufs = [C.create_string_buffer(bytes(v), max(len(v), 1)) for v in blobs.values()]
        self._bd = (C.c_void_p * max(n_b, 1))(*[C.cast(b, C.c_void_p) for b in self._bufs])
        self._bl = (C.c_size_t * max(n_b, 1))(*[len(v) for v in blobs.values()])
        self.struct = KvOverrides(C.cast(self._set, C.POINTER(C.c_char_p)), n_set,
                                  C.cast(self._ss, C.POINTER(C.c_char_p)), n_ss,
                                  C.cast(self._sf, C.POINTER(C.c_char_p)), n_sf,
                                  C.cast(self._bk, C.POINTER(C.c_char_p)),
                                  C.cast(self._bd, C.POINTER(C.c_void_p)),
                                  C.cast(self._bl, C.POINTER(C.c_size_t)), n_b)

    def ptr(self):
        return C.cast(C.pointer(self.struct), C.c_void_p)


class Chart:
    """A loaded chart directory (Chart.yaml, values.yaml, templates/, .helmignore)."""

    def __init__(self, chart_dir: str):
        self._h = C.c_void_p()
        rc = _lib.lib().kv_chart_load(chart_dir.encode(), C.byref(self._h))
        if rc:
            _raise(rc)
        self.dir = chart_dir

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:   # at interpreter exit the module may be gone
            _lib.lib().kv_chart_free(self._h)
            self._h = None

    def plan(self, release_name: str = DEFAULT_RELEASE, set: Iterable[str] = (), set_string: Iterable[str] = (),
             set_file: Iterable[str] = (), blobs: Optional[Dict[str, bytes]] = None,
             namespace: str = "default") -> "Release":
        ov = _Overrides(set, set_string, set_file, blobs)
        h = C.c_void_p()
        rc = _lib.lib().kv_release_plan(self._h, release_name.encode(), namespace.encode(), ov.ptr(), C.byref(h))
        if rc:
            _raise(rc)
        return Release(h)

    def template(self, release_name: str = DEFAULT_RELEASE, **kw) -> bytes:
        r = self.plan(release_name, **kw)
        r.encode_gpu()
        return r.manifest()

    def template_fleet(self, releases: Sequence[Tuple[str, dict]], namespace: str = "default") -> List[bytes]:
        """Render many installs of this chart; all their Secret payloads share one GPU batch.
        ``releases`` = [(release_name, {"set": [...], "set_file": [...], "blobs": {...}}), ...]."""
        n = len(releases)
        ovs = [_Overrides(**kw) for _, kw in releases]
        arr = (KvOverrides * max(n, 1))(*[o.struct for o in ovs])
        names = (C.c_char_p * max(n, 1))(*[name.encode() for name, _ in releases])
        out = C.c_void_p()
        off = np.zeros(n + 1, dtype=np.int64)
        rc = _lib.lib().kv_fleet_template(self._h, n, C.cast(names, C.c_void_p), namespace.encode(),
                                          C.cast(arr, C.c_void_p), C.byref(out), off.ctypes.data)
        if rc:
            _raise(rc)
        try:
            raw = C.string_at(out, int(off[-1]))
        finally:
            _lib.lib().kv_free(out)
        return [raw[int(off[i]):int(off[i + 1])] for i in range(n)]


class Release:
    """One planned release: rendered text with its ``b64enc`` sites still pending."""

    def __init__(self, handle):
        self._h = handle

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.lib().kv_release_free(self._h)
            self._h = None

    @property
    def pending(self) -> List[bytes]:
        """Input bytes of every pending ``b64enc`` site, in template order."""
        L = _lib.lib()
        out = []
        for i in range(L.kv_release_pending_count(self._h)):
            p, n = C.c_void_p(), C.c_size_t()
            check(L.kv_release_pending_input(self._h, i, C.byref(p), C.byref(n)))
            out.append(C.string_at(p, n.value))
        return out

    def fill(self, i: int, encoded: bytes) -> None:
        check(_lib.lib().kv_release_pending_fill(self._h, i, encoded, len(encoded)))

    def encode_gpu(self) -> None:
        check(_lib.lib().kv_release_encode_gpu(self._h))

    def manifest(self) -> bytes:
        p, n = C.c_void_p(), C.c_size_t()
        check(_lib.lib().kv_release_manifest(self._h, C.byref(p), C.byref(n)))
        return C.string_at(p, n.value)

    def notes(self) -> bytes:
        p, n = C.c_void_p(), C.c_size_t()
        check(_lib.lib().kv_release_notes(self._h, C.byref(p), C.byref(n)))
        return C.string_at(p, n.value)


def template(chart_dir: str, release_name: str = DEFAULT_RELEASE, set: Iterable[str] = (),
             set_string: Iterable[str] = (), set_file: Iterable[str] = (),
             blobs: Optional[Dict[str, bytes]] = None, namespace: str = "default") -> bytes:
    """``helm template <release> <chart> [--set ..] [--set-string ..] [--set-file ..]``."""
    ov = _Overrides(set, set_string, set_file, blobs)
    out, n = C.c_void_p(), C.c_size_t()
    rc = _lib.lib().kv_helm_template(chart_dir.encode(), release_name.encode(), namespace.encode(), ov.ptr(),
                                     C.byref(out), C.byref(n))
    if rc:
        _raise(rc)
    try:
        return C.string_at(out, n.value)
    finally:
        _lib.lib().kv_free(out)


def notes(chart_dir: str, release_name: str, **kw) -> bytes:
    """The NOTES block ``helm install`` prints for this release."""
    r = Chart(chart_dir).plan(release_name, **kw)
    r.encode_gpu()
    return r.notes()


def lint(chart_dir: str, set: Iterable[str] = (), set_file: Iterable[str] = ()) -> Tuple[bool, List[str]]:
    """``helm lint`` equivalent: (ok, messages tagged [INFO]/[WARNING]/[ERROR])."""
    ov = _Overrides(set, (), set_file)
    ok, msgs, n = C.c_int(0), C.c_void_p(), C.c_size_t()
    rc = _lib.lib().kv_helm_lint(chart_dir.encode(), ov.ptr(), C.byref(ok), C.byref(msgs), C.byref(n))
    if rc:
        _raise(rc)
    try:
        text = C.string_at(msgs, n.value).decode("utf-8", "replace")
    finally:
        _lib.lib().kv_free(msgs)
    return bool(ok.value), [m for m in text.split("\n") if m]


__all__ = ["Chart", "Release", "template", "notes", "lint", "TemplateError", "DEFAULT_RELEASE"]
