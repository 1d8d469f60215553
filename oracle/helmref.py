"""CPU restatement of what Helm v3 does around the reference chart: load the chart
directory, merge values, render every template, frame the result as
``helm template`` / ``helm install`` (NOTES) output, and lint.

TEST INFRASTRUCTURE ONLY (the checker); never imported by ``kvedge_b200``.

Follows (reference file:line, relative to /root/reference):
  * chart layout ............ ``deployment/helm/Chart.yaml:1-23``, ``values.yaml:1-17``,
                              ``templates/*``, ``.helmignore:1-25`` (l.25 disables the
                              HTTP-disk DataVolume and MUST be honoured)
  * value overrides ......... ``README.md:60`` (``--set k=v,k=v`` and ``--set-file k=path``),
                              ``deployment/az-iot-edge-k8s-kubevirt-ascii.cast:2280-2292``
  * install output .......... ``cast:2294`` (the only recorded real-Helm output; golden)

Helm itself is third-party and absent from the reference and from this image
([ext] in SURVEY.md): the ``helm template`` framing, manifest ordering and lint
messages below restate Helm 3's published behaviour (pkg/releaseutil
``SortManifests``/``InstallOrder``, pkg/action ``renderResources``,
pkg/strvals, pkg/chartutil ``ToRenderValues``, pkg/lint/rules) and are marked
UNVERIFIED against a real binary in DESIGN.md.
"""
from __future__ import annotations

import fnmatch
import os
import re
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple

import yaml

from . import gotmpl
from .gotmpl import GoStruct, TemplateError, TemplateSet


# --------------------------------------------------------------------------- YAML (go-yaml v2 typing)
_BOOL_TRUE = {"y", "Y", "yes", "Yes", "YES", "on", "On", "ON", "true", "True", "TRUE"}
_BOOL_FALSE = {"n", "N", "no", "No", "NO", "off", "Off", "OFF", "false", "False", "FALSE"}
_NULLS = {"", "~", "null", "Null", "NULL"}
_FLOAT_RE = re.compile(r"^[-+]?(\.[0-9]+|[0-9]+(\.[0-9]*)?)([eE][-+]?[0-9]+)?$")


def _resolve_plain(s: str, json_numbers: bool) -> Any:
    """Type a plain (unquoted) YAML scalar the way gopkg.in/yaml.v2 ``resolve`` does; with
    ``json_numbers`` every number becomes float64 (sigs.k8s.io/yaml round-trips via JSON,
    which is how Helm reads values.yaml)."""
    if s in _NULLS:
        return None
    if s in _BOOL_TRUE:
        return True
    if s in _BOOL_FALSE:
        return False
    plain = s.replace("_", "")
    m = re.match(r"^([-+]?)(0[xX][0-9a-fA-F]+|0[oO][0-7]+|0[bB][01]+|0[0-7]*|[1-9][0-9]*)$", plain)
    if m:
        body = m.group(2)
        if len(body) > 1 and body[0] == "0" and body[1] in "01234567":
            v = int(body, 8)
        else:
            v = int(body, 0)
        if m.group(1) == "-":
            v = -v
        if -(1 << 63) <= v < (1 << 64):
            return float(v) if json_numbers else v
    if _FLOAT_RE.match(plain):
        return float(plain)
    if plain in (".inf", ".Inf", ".INF", "+.inf", "+.Inf", "+.INF"):
        return float("inf")
    if plain in ("-.inf", "-.Inf", "-.INF"):
        return float("-inf")
    if plain in (".nan", ".NaN", ".NAN"):
        return float("nan")
    return s.encode("utf-8")


def _construct(node: yaml.Node, json_numbers: bool) -> Any:
    if isinstance(node, yaml.ScalarNode):
        if node.style is None and node.tag in ("tag:yaml.org,2002:str", "tag:yaml.org,2002:null",
                                               "tag:yaml.org,2002:bool", "tag:yaml.org,2002:int",
                                               "tag:yaml.org,2002:float", "tag:yaml.org,2002:timestamp",
                                               "tag:yaml.org,2002:value", "tag:yaml.org,2002:merge"):
            return _resolve_plain(node.value, json_numbers)
        return node.value.encode("utf-8")
    if isinstance(node, yaml.SequenceNode):
        return [_construct(c, json_numbers) for c in node.value]
    if isinstance(node, yaml.MappingNode):
        out: Dict[bytes, Any] = {}
        for k, v in node.value:
            kk = _construct(k, json_numbers)
            if kk is None:
                raise ValueError("yaml: unsupported map key of type <nil>")  # sigs.k8s.io/yaml refuses it
            out[gotmpl.go_print(kk) if not isinstance(kk, bytes) else kk] = _construct(v, json_numbers)
        return out
    raise ValueError("unsupported YAML node")


def load_yaml(data: bytes, json_numbers: bool = True) -> Any:
    """Parse one YAML document into template values (bytes strings, float64 numbers)."""
    node = yaml.compose(data.decode("utf-8"), Loader=yaml.SafeLoader)
    if node is None:
        return None
    return _construct(node, json_numbers)


# --------------------------------------------------------------------------- strvals (--set)
def _set_typed(val: str, force_string: bool) -> Any:
    """pkg/strvals ``typedVal``."""
    if force_string:
        return val.encode("utf-8")
    if val.lower() == "true":
        return True
    if val.lower() == "false":
        return False
    if val.lower() == "null":
        return None
    if val == "0":
        return 0
    if val and val[0] != "0" and re.match(r"^[-+]?[0-9]+$", val):
        try:
            v = int(val)
            if -(1 << 63) <= v < (1 << 63):
                return v
        except ValueError:
            pass
    return val.encode("utf-8")


def _split_set(s: str) -> List[Tuple[str, str]]:
    """Split ``k=v,k2=v2`` on unescaped commas (``\\,`` is a literal comma); a value may be a
    ``{a,b}`` list."""
    pairs: List[Tuple[str, str]] = []
    i = 0
    n = len(s)
    while i < n:
        key = []
        while i < n and s[i] != "=":
            if s[i] == "\\" and i + 1 < n:
                key.append(s[i:i + 2])
                i += 2
                continue
            if s[i] == ",":
                raise ValueError(f"key {''.join(key)!r} has no value")
            key.append(s[i])
            i += 1
        if i >= n:
            raise ValueError(f"key {''.join(key)!r} has no value")
        i += 1
        val = []
        if i < n and s[i] == "{":
            j = s.find("}", i)
            if j < 0:
                raise ValueError("list must terminate with '}'")
            val.append(s[i:j + 1])
            i = j + 1
            if i < n and s[i] == ",":
                i += 1
        else:
            while i < n:
                if s[i] == "\\" and i + 1 < n:
                    val.append(s[i + 1])
                    i += 2
                    continue
                if s[i] == ",":
                    i += 1
                    break
                val.append(s[i])
                i += 1
        pairs.append(("".join(key), "".join(val)))
    return pairs


def _split_key(key: str) -> List[Any]:
    """``a.b[2].c`` -> ['a','b',2,'c'] with ``\\.`` as a literal dot."""
    parts: List[Any] = []
    cur = []
    i = 0
    while i < len(key):
        c = key[i]
        if c == "\\" and i + 1 < len(key):
            cur.append(key[i + 1])
            i += 2
        elif c == ".":
            if cur:
                parts.append("".join(cur))
            cur = []
            i += 1
        elif c == "[":
            if cur:
                parts.append("".join(cur))
            cur = []
            j = key.index("]", i)
            parts.append(int(key[i + 1:j]))
            i = j + 1
        else:
            cur.append(c)
            i += 1
    if cur:
        parts.append("".join(cur))
    return parts


def _assign(root: Dict[bytes, Any], path: List[Any], value: Any) -> None:
    cur: Any = root
    for idx, p in enumerate(path):
        last = idx == len(path) - 1
        nxt_is_index = (not last) and isinstance(path[idx + 1], int)
        if isinstance(p, int):
            if not isinstance(cur, list):
                raise ValueError(f"--set: cannot index into a map with [{p}]")
            while len(cur) <= p:
                cur.append(None)
            if last:
                cur[p] = value
            else:
                if not isinstance(cur[p], (dict, list)):
                    cur[p] = [] if nxt_is_index else {}
                cur = cur[p]
        else:
            if not isinstance(cur, dict):
                # Helm's strvals type-asserts here and panics; both engines refuse the expression
                raise ValueError(f"--set: key {p!r} addresses a list, not a map")
            k = p.encode("utf-8")
            if last:
                cur[k] = value
            else:
                if not isinstance(cur.get(k), (dict, list)):
                    cur[k] = [] if nxt_is_index else {}
                cur = cur[k]


def apply_set(values: Dict[bytes, Any], expr: str, force_string: bool = False) -> None:
    """``--set`` / ``--set-string`` (README.md:60)."""
    for key, val in _split_set(expr):
        if val.startswith("{") and val.endswith("}"):
            items = [x for x in val[1:-1].split(",")] if val != "{}" else []
            v: Any = [_set_typed(x.strip(), force_string) for x in items]
        else:
            v = _set_typed(val, force_string)
        _assign(values, _split_key(key), v)


def apply_set_file(values: Dict[bytes, Any], expr: str) -> None:
    """``--set-file key=path`` (README.md:60): the file's bytes, verbatim, as a string."""
    for key, path in _split_set(expr):
        with open(path, "rb") as f:
            _assign(values, _split_key(key), f.read())


def deep_merge(dst: Dict[bytes, Any], src: Dict[bytes, Any]) -> Dict[bytes, Any]:
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            deep_merge(dst[k], v)
        else:
            dst[k] = v
    return dst


# --------------------------------------------------------------------------- chart
def _ignore_rules(text: str) -> List[Tuple[str, bool, bool, bool]]:
    rules = []
    for line in text.splitlines():
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        neg = line.startswith("!")
        if neg:
            line = line[1:]
        dir_only = line.endswith("/")
        if dir_only:
            line = line[:-1]
        anchored = line.startswith("/")
        if anchored:
            line = line[1:]
        rules.append((line, neg, dir_only, anchored))
    return rules


def _ignored(rel: str, is_dir: bool, rules) -> bool:
    """pkg/ignore: a pattern without '/' matches the base name; with '/' the relative path."""
    ig = False
    for pat, neg, dir_only, anchored in rules:
        if dir_only and not is_dir:
            continue
        if "/" in pat or anchored:
            hit = fnmatch.fnmatchcase(rel, pat)
        else:
            hit = fnmatch.fnmatchcase(os.path.basename(rel), pat)
        if hit:
            ig = not neg
    return ig


@dataclass
class Chart:
    path: str
    metadata: Dict[bytes, Any]
    values: Dict[bytes, Any]
    templates: Dict[str, bytes]          # "templates/<file>" -> source
    raw_files: Dict[str, bytes] = field(default_factory=dict)

    @property
    def name(self) -> str:
        return gotmpl.go_print(self.metadata.get(b"name", b"")).decode()


def load_chart(path: str) -> Chart:
    with open(os.path.join(path, "Chart.yaml"), "rb") as f:
        meta = load_yaml(f.read(), json_numbers=False) or {}
    vals: Dict[bytes, Any] = {}
    vp = os.path.join(path, "values.yaml")
    if os.path.exists(vp):
        with open(vp, "rb") as f:
            vals = load_yaml(f.read()) or {}
    rules = []
    ip = os.path.join(path, ".helmignore")
    if os.path.exists(ip):
        with open(ip, "r") as f:
            rules = _ignore_rules(f.read())
    templates: Dict[str, bytes] = {}
    tdir = os.path.join(path, "templates")
    for root, dirs, files in os.walk(tdir):
        dirs[:] = sorted(d for d in dirs
                         if not _ignored(os.path.relpath(os.path.join(root, d), path), True, rules))
        for fn in sorted(files):
            rel = os.path.relpath(os.path.join(root, fn), path).replace(os.sep, "/")
            if _ignored(rel, False, rules):
                continue
            with open(os.path.join(root, fn), "rb") as f:
                templates[rel] = f.read()
    return Chart(path, meta, vals, templates)


def _chart_struct(meta: Dict[bytes, Any]) -> GoStruct:
    """chart.Metadata as the template sees it (``.Chart.Name`` ...)."""
    def g(k, d=b""):
        v = meta.get(k, d)
        return v if v is not None else d
    return GoStruct("chart.Metadata", {
        "Name": g(b"name"), "Home": g(b"home"), "Sources": g(b"sources", []),
        "Version": gotmpl.strval(g(b"version")), "Description": g(b"description"),
        "Keywords": g(b"keywords", []), "Maintainers": g(b"maintainers", []),
        "Icon": g(b"icon"), "APIVersion": g(b"apiVersion"), "Condition": g(b"condition"),
        "Tags": g(b"tags"), "AppVersion": gotmpl.strval(g(b"appVersion")),
        "Deprecated": bool(g(b"deprecated", False)), "Annotations": g(b"annotations", {}),
        "KubeVersion": g(b"kubeVersion"), "Dependencies": g(b"dependencies", []),
        "Type": g(b"type"),
    })


DEFAULT_RELEASE = "release-name"


def render_values(chart: Chart, set_args=(), set_string_args=(), set_file_args=(),
                  values_override: Optional[Dict[bytes, Any]] = None) -> Dict[bytes, Any]:
    import copy
    # Helm: the flags build a map of their own (pkg/cli/values MergeValues), then
    # chartutil.CoalesceValues slides the chart's defaults under it.
    user: Dict[bytes, Any] = {}
    if values_override:
        deep_merge(user, copy.deepcopy(values_override))
    for e in set_args:
        apply_set(user, e)
    for e in set_string_args:
        apply_set(user, e, force_string=True)
    for e in set_file_args:
        apply_set_file(user, e)
    return coalesce_values(chart.values, user)


def coalesce_values(defaults: Dict[bytes, Any], user: Dict[bytes, Any]) -> Dict[bytes, Any]:
    """chartutil.coalesceValues / coalesceTablesFullKey: a user null deletes a key the chart
    defines; tables merge key by key; the user's lists and scalars win whole."""
    import copy
    for k, dv in defaults.items():
        if k not in user:
            user[k] = copy.deepcopy(dv)
        elif user[k] is None:
            del user[k]
        elif isinstance(dv, dict) and isinstance(user[k], dict):
            coalesce_values(dv, user[k])
    return user


def render_files(chart: Chart, values: Dict[bytes, Any], release_name: str = DEFAULT_RELEASE,
                 namespace: str = "default", is_install: bool = True) -> Dict[str, bytes]:
    """pkg/engine ``Render``: every non-partial file under templates/ -> rendered bytes, keyed
    ``<chart>/templates/<file>``; "<no value>" replaced by "" (missingkey=zero, non-strict)."""
    ts = TemplateSet()
    prefix = chart.name + "/"
    for rel, src in chart.templates.items():
        ts.parse(prefix + rel, src)
    out: Dict[str, bytes] = {}
    for rel in chart.templates:
        base = os.path.basename(rel)
        if base.startswith("_"):
            continue
        top = {
            "Values": values,
            "Chart": _chart_struct(chart.metadata),
            "Release": {"Name": release_name.encode(), "Namespace": namespace.encode(),
                        "IsUpgrade": not is_install, "IsInstall": is_install, "Revision": 1,
                        "Service": b"Helm"},
            "Template": {"Name": (prefix + rel).encode(),
                         "BasePath": (prefix + "templates").encode()},
            "Capabilities": {"KubeVersion": {"Version": b"v1.20.0", "Major": b"1", "Minor": b"20"},
                             "HelmVersion": {"Version": b"v3.7.0"}},
        }
        try:
            rendered = ts.execute(prefix + rel, top)
        except TemplateError as e:
            raise TemplateError(f"template: {prefix + rel}: {e}") from None
        out[prefix + rel] = rendered.replace(b"<no value>", b"")
    return out


# pkg/releaseutil InstallOrder
INSTALL_ORDER = [
    "Namespace", "NetworkPolicy", "ResourceQuota", "LimitRange", "PodSecurityPolicy",
    "PodDisruptionBudget", "ServiceAccount", "Secret", "SecretList", "ConfigMap", "StorageClass",
    "PersistentVolume", "PersistentVolumeClaim", "CustomResourceDefinition", "ClusterRole",
    "ClusterRoleList", "ClusterRoleBinding", "ClusterRoleBindingList", "Role", "RoleList",
    "RoleBinding", "RoleBindingList", "Service", "DaemonSet", "Pod", "ReplicationController",
    "ReplicaSet", "Deployment", "HorizontalPodAutoscaler", "StatefulSet", "Job", "CronJob",
    "IngressClass", "Ingress", "APIService",
]
# releaseutil.SplitManifests' separator, verbatim; Go's \s is [\t\n\f\r ] (no \v)
_SEP = re.compile(rb"(?:^|[\t\n\f\r ]*\n)---[\t\n\f\r ]*")
_KIND = re.compile(rb"(?m)^kind:[ \t]*([^\s#]+)")


@dataclass
class Manifest:
    name: str      # "<chart>/templates/<file>"
    kind: str
    content: bytes


def split_manifests(files: Dict[str, bytes]) -> Tuple[List[Manifest], bytes]:
    """pkg/releaseutil ``SortManifests``: split files into documents, drop empties, take the
    kind from the document head, stable-sort by install order; NOTES.txt is returned apart."""
    notes = b""
    manifests: List[Manifest] = []
    for name in sorted(files):
        content = files[name]
        if os.path.basename(name) == "NOTES.txt":
            notes = content
            continue
        if not content.strip():
            continue
        for doc in _SEP.split(content.strip()):
            doc = doc.lstrip()
            if not doc:
                continue
            m = _KIND.search(doc)
            kind = m.group(1).decode() if m else ""
            manifests.append(Manifest(name, kind.strip("\"'"), doc))
    order = {k: i for i, k in enumerate(INSTALL_ORDER)}

    def key(m: Manifest):
        if m.kind in order:
            return (0, order[m.kind], "")
        return (1, 0, m.kind)

    manifests.sort(key=key)  # Python's sort is stable, like sort.SliceStable
    return manifests, notes


def helm_template(chart_dir: str, release_name: str = DEFAULT_RELEASE, set_args=(),
                  set_string_args=(), set_file_args=(), namespace: str = "default") -> bytes:
    """``helm template <release> <chart> [--set ...] [--set-file ...]`` output stream."""
    chart = load_chart(chart_dir)
    vals = render_values(chart, set_args, set_string_args, set_file_args)
    files = render_files(chart, vals, release_name, namespace)
    manifests, _ = split_manifests(files)
    out = bytearray()
    for m in manifests:
        out += b"---\n# Source: " + m.name.encode() + b"\n" + m.content + b"\n"
    return bytes(out)


def helm_notes(chart_dir: str, release_name: str, set_args=(), set_string_args=(),
               set_file_args=(), namespace: str = "default") -> bytes:
    """The NOTES block ``helm install`` prints (cast:2294 is the golden)."""
    chart = load_chart(chart_dir)
    vals = render_values(chart, set_args, set_string_args, set_file_args)
    files = render_files(chart, vals, release_name, namespace)
    _, notes = split_manifests(files)
    return notes.strip()


def secret_userdata(stream: bytes, secret_name: str) -> Optional[bytes]:
    """Byte-level extraction of ``data.userdata`` from the Secret called ``secret_name`` in a
    ``helm template`` stream (by bytes, not through a YAML parser: SURVEY.md 7.1 step 4)."""
    for doc in re.split(rb"(?m)^---\n", stream):
        if re.search(rb"(?m)^kind: Secret\s*$", doc) and \
                re.search(rb"(?m)^  name: " + re.escape(secret_name.encode()) + rb"\s*$", doc):
            m = re.search(rb"(?m)^  userdata:[ ]?(.*)$", doc)
            return m.group(1) if m else None
    return None


# --------------------------------------------------------------------------- lint
_DNS1123 = re.compile(r"^[a-z0-9]([-a-z0-9]*[a-z0-9])?(\.[a-z0-9]([-a-z0-9]*[a-z0-9])?)*$")
_SEMVER = re.compile(r"^v?\d+\.\d+\.\d+(-[0-9A-Za-z.-]+)?(\+[0-9A-Za-z.-]+)?$")


def helm_lint(chart_dir: str, set_args=(), set_file_args=()) -> Tuple[bool, List[str]]:
    """``helm lint`` equivalent (BASELINE.json config 1): Chart.yaml fields, values parse,
    every template renders, every document is YAML with apiVersion/kind/metadata.name and a
    DNS-1123 name. Messages carry Helm's severity tags; ok == no [ERROR]."""
    msgs: List[str] = []
    try:
        chart = load_chart(chart_dir)
    except Exception as e:  # noqa: BLE001
        return False, [f"[ERROR] Chart.yaml: {e}"]
    meta = chart.metadata
    if not meta.get(b"name"):
        msgs.append("[ERROR] Chart.yaml: name is required")
    if not meta.get(b"apiVersion"):
        msgs.append("[ERROR] Chart.yaml: apiVersion is required. The value must be either \"v1\" or \"v2\"")
    elif meta[b"apiVersion"] not in (b"v1", b"v2"):
        msgs.append("[ERROR] Chart.yaml: apiVersion '%s' is not valid. The value must be either \"v1\" or \"v2\""
                    % gotmpl.go_print(meta[b"apiVersion"]).decode())
    ver = gotmpl.go_print(meta.get(b"version", b"")).decode() if meta.get(b"version") is not None else ""
    if not ver:
        msgs.append("[ERROR] Chart.yaml: version is required")
    elif not _SEMVER.match(ver):
        msgs.append(f"[ERROR] Chart.yaml: version '{ver}' is not a valid SemVer")
    if meta.get(b"type") not in (None, b"application", b"library"):
        msgs.append("[ERROR] Chart.yaml: chart type is not valid in apiVersion 'v2'. It must be 'application' or 'library'")
    if not meta.get(b"icon"):
        msgs.append("[INFO] Chart.yaml: icon is recommended")
    if not os.path.isdir(os.path.join(chart_dir, "templates")):
        msgs.append("[WARNING] templates/: directory not found")
    try:
        vals = render_values(chart, set_args, (), set_file_args)
        files = render_files(chart, vals, "test-release")
    except (TemplateError, ValueError, OSError) as e:
        msgs.append(f"[ERROR] templates/: {e}")
        return False, msgs
    for name in sorted(files):
        if os.path.basename(name) == "NOTES.txt" or not files[name].strip():
            continue
        short = name.split("/", 1)[1]
        for doc in _SEP.split(files[name].strip()):
            if not doc.strip():
                continue
            try:
                obj = yaml.compose(doc.decode("utf-8"), Loader=yaml.SafeLoader)
            except Exception as e:  # noqa: BLE001
                msgs.append(f"[ERROR] {short}: unable to parse YAML: {str(e).splitlines()[0]}")
                continue
            if not isinstance(obj, yaml.MappingNode):
                msgs.append(f"[ERROR] {short}: document is not a mapping")
                continue
            top = {k.value: v for k, v in obj.value}
            for req in ("apiVersion", "kind"):
                if req not in top:
                    msgs.append(f"[ERROR] {short}: {req} is required")
            md = top.get("metadata")
            nm = None
            if isinstance(md, yaml.MappingNode):
                for k, v in md.value:
                    if k.value == "name":
                        nm = v.value
            if not nm:
                msgs.append(f"[ERROR] {short}: object name not present in metadata")
            elif len(nm) > 253 or not _DNS1123.match(nm):
                msgs.append(f"[ERROR] {short}: object name does not conform to Kubernetes naming "
                            f"requirements: \"{nm}\"")
    ok = not any(m.startswith("[ERROR]") for m in msgs)
    return ok, msgs
