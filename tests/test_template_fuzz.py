"""Differential fuzz of the template executor (the helpers of
deployment/helm/templates/_helper.tpl are pipelines of exactly these functions): random templates
built from the Go-template / Sprig subset go through the C++ engine and through the Python
restatement; both must print the same bytes or both must refuse the template.  The two engines
share an author but no code, so this catches slips of either, not a shared misreading of Helm."""
import os

import pytest
from hypothesis import assume, given, settings, strategies as st

from kvedge_b200 import helm
from oracle import b64 as ob64
from oracle import helmref

VALUES = """\
s: "Hello, World-"
t: "  padded\\tline\\n  two  "
u: "aziot-edge-kubevirt-vm-0123456789-0123456789-"
e: ""
n: 42
f: 2.5
big: 1234567
yes: true
no: false
nothing: null
lst: [a, "b c", 3]
m:
  k1: v1
  k2: 7
  inner:
    deep: "x-y"
"""

# rooted at $ so that they still resolve inside with / range bodies (a few stay relative to dot)
OPERAND = st.sampled_from(["$.Values.s", "$.Values.t", "$.Values.u", "$.Values.e", "$.Values.n", "$.Values.f", "$.Values.big",
                           "$.Values.yes", "$.Values.no", "$.Values.nothing", "$.Values.missing", "$.Values.lst", "$.Values.m",
                           "$.Values.m.k1", "$.Values.m.inner.deep", "$.Release.Name", "$.Chart.Name", "$.Chart.AppVersion",
                           '"lit"', '"-"', '""', "7", "-3", "true", "nil", ".", ".Values.s", '(include "h.name" $)'])
SMALL = st.sampled_from(["0", "1", "2", "5", "8", "40", "-4"])
STR = st.sampled_from(['"-"', '"l"', '"o, W"', '""', '"\\n"', '"aziot"', '"é"'])


@st.composite
def stage(draw):
    fn = draw(st.sampled_from(["quote", "squote", "upper", "lower", "trim", "trunc", "indent", "nindent", "replace",
                               "trimSuffix", "trimPrefix", "default", "printf", "len", "empty", "not", "toString",
                               "contains", "hasPrefix", "hasSuffix", "repeat", "cat", "eq", "ne", "ternary", "print",
                               "required", "hasKey"]))
    if fn in ("trunc", "indent", "nindent", "repeat"):
        return f"{fn} {draw(SMALL)}"
    if fn == "replace":
        return f"replace {draw(STR)} {draw(STR)}"
    if fn in ("trimSuffix", "trimPrefix", "contains", "hasPrefix", "hasSuffix", "default", "cat", "eq", "ne", "print"):
        return f"{fn} {draw(st.one_of(STR, OPERAND))}"
    if fn == "printf":
        fmt = draw(st.sampled_from(['"%s|%v"', '"%d-%q"', '"%t"', '"%5s"']))
        return "printf " + fmt + " " + draw(OPERAND)
    if fn == "ternary":
        return f"ternary {draw(STR)} {draw(OPERAND)}"
    if fn == "required":
        return 'required "need it"'
    if fn == "hasKey":
        return "hasKey $.Values.m"      # piped value becomes the key
    return fn


STRING_OPERAND = st.sampled_from(["$.Values.s", "$.Values.t", "$.Values.u", "$.Values.e", "$.Values.m.k1",
                                  "$.Values.m.inner.deep", "$.Release.Name", "$.Chart.Name", '"lit"', '(include "h.name" $)'])
STRING_STAGE = st.sampled_from(["quote", "upper", "lower", "trim", "trunc 5", "trunc -4", "indent 2", "nindent 1",
                                'replace "l" "L"', 'replace "" "."', 'trimSuffix "-"', 'trimPrefix "H"', 'default "d"',
                                'printf "%s|%q"', "repeat 2", 'cat "x"', "toString", "squote", "b64enc"])


@st.composite
def pipeline(draw):
    if draw(st.integers(0, 2)):    # mostly: string in, string functions (they do not refuse their input)
        return " | ".join([draw(STRING_OPERAND)] + draw(st.lists(STRING_STAGE, max_size=4)))
    return " | ".join([draw(OPERAND)] + draw(st.lists(stage(), max_size=4)))


@st.composite
def node(draw, depth=0):
    kind = draw(st.integers(0, 9 if depth < 2 else 4))
    if kind <= 4:
        trim_l, trim_r = draw(st.sampled_from(["", "- "])), draw(st.sampled_from(["", " -"]))
        return "{{" + trim_l + draw(pipeline()) + trim_r + "}}"
    if kind == 5:
        body, other = draw(node(depth + 1)), draw(node(depth + 1))
        return "{{ if " + draw(pipeline()) + " }}" + body + "{{ else }}" + other + "{{ end }}"
    if kind == 6:
        return "{{ with " + draw(pipeline()) + " }}[{{ . }}]" + draw(node(depth + 1)) + "{{ end }}"
    if kind == 7:
        src = draw(st.sampled_from([".Values.lst", ".Values.m", ".Values.nothing", "3", ".Values.s"]))
        return "{{ range $i, $x := " + src + " }}<{{ $i }}={{ $x }}>" + draw(node(depth + 1)) + "{{ else }}none{{ end }}"
    if kind == 8:
        return "{{ $v := " + draw(pipeline()) + " }}({{ $v }}){{ $v = " + draw(pipeline()) + " }}({{ $v }})"
    return draw(st.sampled_from([" text \n", "\n\n", "  ", "a: b\n", "{{/* note {{ x }} */}}", "{{- /* t */ -}}"]))


def _chart(tmp, body):
    os.makedirs(os.path.join(tmp, "templates"), exist_ok=True)
    open(os.path.join(tmp, "Chart.yaml"), "w").write("apiVersion: v2\nname: fuzz\nversion: 0.1.0\nappVersion: \"1.2\"\n")
    open(os.path.join(tmp, "values.yaml"), "w").write(VALUES)
    open(os.path.join(tmp, "templates", "_h.tpl"), "w").write(
        '{{- define "h.name" -}}{{- default .Chart.Name .Values.u | trunc 40 | trimSuffix "-" -}}{{- end -}}\n')
    open(os.path.join(tmp, "templates", "t.yaml"), "w").write("out: |\n" + body + "\n")
    return tmp


def _check(tmp_path_factory, nodes, have_gpu):
    body = " ".join(nodes)
    d = _chart(str(tmp_path_factory.mktemp("tf")), body)

    def engine():
        r = helm.Chart(d).plan("rel")
        if have_gpu:
            r.encode_gpu()
        else:
            for i, p in enumerate(r.pending):
                r.fill(i, ob64.encode(p))
        return r.manifest()

    try:
        ref = helmref.helm_template(d, "rel")
    except Exception:  # noqa: BLE001  parse or execution error: the engine must refuse it too
        with pytest.raises(helm.KvError):
            engine()
        return
    try:
        got = engine()
    except helm.TemplateError as e:
        # a b64enc whose result feeds another function is evaluated at once, on the GPU: without
        # one the engine refuses loudly (there is no CPU fallback), which is right, not a mismatch
        if not have_gpu and "no CPU fallback" in str(e):
            assume(False)
        raise
    assert got == ref, body


@settings(max_examples=400, deadline=None)
@given(st.lists(node(), min_size=1, max_size=5))
def test_engines_agree_on_random_templates(tmp_path_factory, nodes):
    from kvedge_b200 import _lib
    _check(tmp_path_factory, nodes, have_gpu=_lib.lib().kv_device_count() > 0)


@pytest.mark.gpu
@settings(max_examples=300, deadline=None)
@given(st.lists(node(), min_size=1, max_size=5))
def test_engines_agree_on_random_templates_gpu(tmp_path_factory, nodes):
    """The same property with the b64enc sites encoded by the CUDA kernels, deferred and immediate."""
    _check(tmp_path_factory, nodes, have_gpu=True)
