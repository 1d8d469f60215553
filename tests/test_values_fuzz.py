"""Differential test of the values front end: random values.yaml files and
--set expressions go through the C++ engine (kvedge_b200) and through the Python restatement
(oracle/helmref.py, PyYAML-based with go-yaml typing); a template that prints .Values with Go's
fmt rules must come out identical."""
import os

import pytest

from hypothesis import given, settings, strategies as st

from kvedge_b200 import helm
from oracle import helmref

KEY = st.text(alphabet="abcdefghijklmnopqrstuvwxyzABCDEF_", min_size=1, max_size=8)
PLAIN = st.one_of(
    st.sampled_from(["true", "false", "yes", "No", "on", "null", "~", "4Gi", "fe:7e:48:a0:7d:22", "52:54:00:12:34:56",
                     "0x1F", "017", "1_000", "1e3", "-2.5", ".5", "3.", "hello world", "a-b_c", "1.2.3", "12abc",
                     "http://x.y/z", "0", "-0", "007", "+5"]),
    st.integers(-10**9, 10**9).map(str),
    st.floats(allow_nan=False, allow_infinity=False, width=32).map(lambda f: repr(float(f))),
)
QUOTED = st.text(alphabet="abc XYZ019:#-_/'", max_size=12)


def scalar_yaml(draw):
    kind = draw(st.integers(0, 3))
    if kind == 0:
        return draw(PLAIN)
    if kind == 1:
        return '"' + draw(QUOTED).replace('"', "") + '"'
    if kind == 2:
        return "'" + draw(QUOTED).replace("'", "''") + "'"
    return ""


@st.composite
def yaml_doc(draw, depth=0, indent=0):
    """A block-style YAML mapping as text."""
    lines = []
    keys = draw(st.lists(KEY, min_size=1, max_size=4, unique=True))
    for k in keys:
        shape = draw(st.integers(0, 5 if depth < 2 else 2))
        pad = " " * indent
        if shape <= 2:
            lines.append(f"{pad}{k}: {scalar_yaml(draw)}".rstrip())
        elif shape == 3:
            lines.append(f"{pad}{k}:")
            lines.append(draw(yaml_doc(depth=depth + 1, indent=indent + 2)))
        elif shape == 4:
            lines.append(f"{pad}{k}:")
            for _ in range(draw(st.integers(1, 3))):
                lines.append(f"{pad}  - {scalar_yaml(draw) or 'x'}")
        else:
            lines.append(f"{pad}{k}: [{', '.join(draw(st.lists(st.sampled_from(['a', '1', 'true', '\"q r\"', '2.5']), max_size=3)))}]")
        if draw(st.integers(0, 6)) == 0:
            lines.append(f"{pad}# a comment")
    return "\n".join(lines)


def _chart(tmp, values_text):
    os.makedirs(os.path.join(tmp, "templates"), exist_ok=True)
    open(os.path.join(tmp, "Chart.yaml"), "w").write("apiVersion: v2\nname: fz\nversion: 0.1.0\n")
    open(os.path.join(tmp, "values.yaml"), "w").write(values_text + "\n")
    open(os.path.join(tmp, "templates", "v.yaml"), "w").write(
        "v: {{ .Values }}\n{{ range $k, $v := .Values }}k {{ $k }} = {{ $v }} / {{ $v | quote }} / {{ if $v }}T{{ else }}F{{ end }}\n{{ end }}")
    return tmp


@settings(max_examples=150, deadline=None)
@given(yaml_doc())
def test_values_yaml_parsers_agree(tmp_path_factory, text):
    d = _chart(str(tmp_path_factory.mktemp("fz")), text)
    try:
        ref = helmref.helm_template(d, "r")
    except Exception as e:  # noqa: BLE001  e.g. a nil map key: both sides must refuse it
        with pytest.raises(helm.KvError):
            helm.Chart(d).plan("r")
        return
    got = helm.Chart(d).plan("r").manifest()
    assert got == ref, text


SETVAL = st.one_of(st.sampled_from(["true", "False", "null", "0", "012", "12", "-7", "4Gi", "a b", "x\\,y", "{a,b,1}", "1.5", ""]),
                   st.text(alphabet="abcXYZ019 :/_-+=@.", max_size=10))


@settings(max_examples=150, deadline=None)
@given(st.lists(st.tuples(st.lists(KEY, min_size=1, max_size=3), st.one_of(st.none(), st.integers(0, 2)), SETVAL),
                min_size=1, max_size=4), st.booleans())
def test_set_expressions_agree(tmp_path_factory, items, as_string):
    parts = []
    for path, idx, val in items:
        key = ".".join(path) + (f"[{idx}]" if idx is not None else "")
        parts.append(f"{key}={val}")
    expr = ",".join(parts)
    d = _chart(str(tmp_path_factory.mktemp("fs")), "base: 1")
    kw_ref = {"set_string_args": [expr]} if as_string else {"set_args": [expr]}
    kw_got = {"set_string": [expr]} if as_string else {"set": [expr]}
    try:
        ref = helmref.helm_template(d, "r", **kw_ref)
    except Exception as e:  # noqa: BLE001  both sides must refuse the same inputs
        try:
            helm.Chart(d).plan("r", **kw_got)
        except helm.KvError:
            return
        raise AssertionError(f"oracle refused {expr!r} ({e}) but the engine accepted it")
    got = helm.Chart(d).plan("r", **kw_got).manifest()
    assert got == ref, expr
