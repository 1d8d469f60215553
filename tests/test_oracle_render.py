"""Pins the render restatement (oracle/gotmpl.py, oracle/helmref.py).

* cast:2294 of the reference's own asciinema recording holds the only real-Helm output in
  the repo (the NOTES block); the golden copy is in tests/golden/reference_chart.json.
* The chart itself is materialised from tests/golden/reference_chart_files.json (conftest.py),
  so every render test runs on every machine; the two tests marked needs_reference check the
  frozen copies (chart files, cast line) against /root/reference in the build container.
"""
import hashlib
import json
import os
import re

import pytest

from kvedge_b200 import workload
from oracle import b64 as ob64
from oracle import gotmpl, helmref

from conftest import GOLDEN, REF_CHART, REF_CHART_LIVE, REFERENCE

CAST = os.path.join(REFERENCE, "deployment", "az-iot-edge-k8s-kubevirt-ascii.cast")


@pytest.mark.needs_reference
def test_golden_notes_are_the_casts_text(golden_chart):
    with open(CAST) as f:
        ev = json.loads(f.read().splitlines()[2293])
    recorded = ev[2].replace("\r\n", "\n").split("NOTES:\n", 1)[1].strip()
    assert recorded == golden_chart["notes_cast2294"]  # the committed golden is the cast's text


@pytest.mark.needs_reference
def test_frozen_chart_files_are_the_references_bytes():
    with open(os.path.join(GOLDEN, "reference_chart_files.json")) as f:
        frozen = {e["path"]: e for e in json.load(f)["files"]}
    live = {}
    for base, _d, names in os.walk(REF_CHART_LIVE):
        for nm in names:
            full = os.path.join(base, nm)
            live[os.path.relpath(full, REF_CHART_LIVE)] = open(full, "rb").read()
    assert sorted(frozen) == sorted(live) and len(live) == 11
    for path, raw in live.items():
        assert hashlib.sha256(raw).hexdigest() == frozen[path]["sha256"], path
        assert open(os.path.join(REF_CHART, path), "rb").read() == raw, path


def test_notes_equals_real_helm_output_cast2294(golden_chart):
    recorded = golden_chart["notes_cast2294"]
    notes = helmref.helm_notes(REF_CHART, golden_chart["release_name_cast2294"],
                               set_args=["nameOverride=aziot-edge-kubevirt"])
    assert notes.decode() == recorded


def test_object_names_seen_in_cast(golden_chart):
    stream = helmref.helm_template(REF_CHART).decode()
    assert "name: aziot-edge-kubevirt-linux-dv\n" in stream
    assert "name: aziot-edge-kubevirt-linux\n" in stream
    assert "aziot-edge-kubevirt-linux-dv" in golden_chart["cast_object_lines"]["cast:2418"]
    assert "aziot-edge-kubevirt-linux" in golden_chart["cast_object_lines"]["cast:2579"]


def test_config1_stream_matches_golden(golden_chart):
    s = helmref.helm_template(REF_CHART)
    g = golden_chart["config1"]
    assert len(s) == g["stream_len"] and hashlib.sha256(s).hexdigest() == g["stream_sha256"]
    ud = helmref.secret_userdata(s, "aziot-edge-kubevirt-vm-cloudconfig")
    assert ud.decode() == g["cloudinit_default_b64"]
    assert helmref.secret_userdata(s, "aziot-edge-kubevirt-vm-aziotedgeconfig") == b""


def test_config2_secret_is_byte_exact(golden_chart, tmp_path):
    """BASELINE.json config 2: helm template --set-file azIotEdgeConfig=config.toml."""
    g = golden_chart["config2"]
    cfg = workload.config_toml(g["config_toml_seed"])
    p = tmp_path / "config.toml"
    p.write_bytes(cfg)
    s = helmref.helm_template(REF_CHART, set_args=[g["set"]], set_file_args=[f"azIotEdgeConfig={p}"])
    assert hashlib.sha256(s).hexdigest() == g["stream_sha256"]
    ud = helmref.secret_userdata(s, "aziot-edge-kubevirt-vm-aziotedgeconfig")
    assert ud == ob64.encode(cfg)
    assert ob64.decode(ud) == cfg                       # consumer side: _helper.tpl:73
    assert re.search(rb"(?m)^  userdata: " + re.escape(ud) + rb"$", s)  # 2-space indent, one line
    assert hashlib.sha256(ud).hexdigest() == g["runtime_secret_userdata_sha256"]
    ci = ob64.decode(helmref.secret_userdata(s, "aziot-edge-kubevirt-vm-cloudconfig"))
    assert b"\n  - " + workload.ssh_key(g["ssh_key_seed"]) + b"\napt:" in ci   # _helper.tpl:35


def test_helmignore_drops_http_disk_datavolume():
    chart = helmref.load_chart(REF_CHART)
    assert "templates/aziot-edge-data-volume-disk.yaml" not in chart.templates   # .helmignore:25
    assert "templates/aziot-edge-data-volume-container.yaml" in chart.templates
    assert helmref.helm_template(REF_CHART).count(b"kind: DataVolume") == 1


def test_manifest_order_and_framing():
    s = helmref.helm_template(REF_CHART).decode()
    srcs = re.findall(r"(?m)^# Source: (.*)$", s)
    assert srcs == ["aziot-edge-kubevirt/templates/" + f for f in (
        "aziot-edge-runtime-config-secret.yaml", "aziot-edge-vm-cloud-init-secret.yaml",
        "aziot-edge-vm-service.yaml", "aziot-edge-data-volume-container.yaml", "aziot-edge-vm.yaml")]
    assert s.startswith("---\n# Source: ")
    # the labels helper begins with a newline, so `indent` leaves a spaces-only line
    assert "  labels:\n      \n    app.kubernetes.io/version: \"0.1.0\"\n" in s
    # an action inside a YAML comment is still executed (aziot-edge-vm.yaml:57)
    assert "# TODO: we should be using aziot-edge-kubevirt here instead" in s


def test_service_toggle_and_eq_type_error():
    off = helmref.helm_template(REF_CHART, set_args=["aziotEdgeVmEnableExternalSsh=false"])
    assert b"kind: Service" not in off
    with pytest.raises(gotmpl.TemplateError, match="incompatible types"):
        helmref.helm_template(REF_CHART, set_string_args=["aziotEdgeVmEnableExternalSsh=true"])


def test_name_helper_vs_raw_name_override():
    long = "a" * 39 + "-bcdef"
    s = helmref.helm_template(REF_CHART, set_args=[f"nameOverride={long}"]).decode()
    assert f"name: {'a' * 39}-vm-aziotedgeconfig" in s          # trunc 40 | trimSuffix "-"
    assert f"name: {long}-vm-cloudconfig" in s                   # raw .Values.nameOverride
    s = helmref.helm_template(REF_CHART, set_args=["nameOverride="]).decode()
    assert "name: aziot-edge-kubevirt-vm-aziotedgeconfig" in s   # default .Chart.Name
    assert "\n  name: -vm-cloudconfig" in s


def test_lint_passes_with_icon_info(golden_chart):
    ok, msgs = helmref.helm_lint(REF_CHART)
    assert ok and msgs == ["[INFO] Chart.yaml: icon is recommended"]
    assert golden_chart["lint"] == {"ok": ok, "messages": msgs}


# ---- engine semantics that need no reference checkout -------------------------------------
def _render(src: bytes, data=None, defines: bytes = b"") -> bytes:
    ts = gotmpl.TemplateSet()
    if defines:
        ts.parse("defs", defines)
    ts.parse("t", src)
    return ts.execute("t", data if data is not None else {})


def test_trim_markers_and_comment_with_braces():
    assert _render(b"a  {{- 1 }}  b") == b"a1  b"
    assert _render(b"a  {{ 1 -}}  b") == b"a  1b"
    assert _render(b"a {{-3}} b") == b"a -3 b"
    assert _render(b"x{{/* c {{ include \"n\" . }}*/}}y") == b"xy"
    assert _render(b"x \n{{- /* c */ -}}\n y") == b"xy"


def test_sprig_functions_used_by_the_chart():
    v = {"Values": {b"n": b"", b"s": b"abc-", b"k": True}, "Chart": {"Name": b"chartname"}}
    assert _render(b'{{ default .Chart.Name .Values.n | trunc 5 | trimSuffix "-" }}', v) == b"chart"
    assert _render(b'{{ .Values.s | trunc 4 | trimSuffix "-" }}', v) == b"abc"
    assert _render(b'{{ "0.1.0" | quote }}', v) == b'"0.1.0"'
    assert _render(b'{{ "a\\nb\\"" | quote }}') == b'"a\\nb\\""'
    assert _render(b'{{ "a\\nb" | indent 2 }}') == b"  a\n  b"
    assert _render(b'{{ printf "%s-%s" "a+b" "1" | replace "+" "_" }}') == b"a_b-1"
    assert _render(b'{{ "foobar" | b64enc }}') == b"Zm9vYmFy"
    assert _render(b'{{ if eq .Values.k true }}y{{ else }}n{{ end }}', v) == b"y"
    assert _render(b"{{ .Values.missing }}", v) == b"<no value>"


def test_number_printing_follows_go():
    assert _render(b"{{ .a }} {{ .b }} {{ .c }}", {"a": 4.0, "b": 1234567.0, "c": 7}) == b"4 1.234567e+06 7"


def test_values_yaml_typing_like_go_yaml():
    v = helmref.load_yaml(b"a: 4Gi\nb: true\nc: \"true\"\nd: 12\ne: fe:7e:48:a0:7d:22\nf: 52:54:00:12:34:56\ng: ''\nh: yes\n")
    assert v[b"a"] == b"4Gi" and v[b"b"] is True and v[b"c"] == b"true" and v[b"d"] == 12.0
    assert v[b"e"] == b"fe:7e:48:a0:7d:22"
    assert v[b"f"] == b"52:54:00:12:34:56"      # NOT a sexagesimal integer (SURVEY 7.1 hazard)
    assert v[b"g"] == b"" and v[b"h"] is True


def test_strvals_set_parsing():
    vals = {}
    helmref.apply_set(vals, "publicSshKey=ssh-rsa AAAA+/= u@h,aziotEdgeVmEnableExternalSsh=true,"
                            "aziotEdgeVmDiskSize=4Gi,nameOverride=aziot-edge-kubevirt,n=12,z=012,a.b=c\\,d")
    assert vals[b"publicSshKey"] == b"ssh-rsa AAAA+/= u@h"
    assert vals[b"aziotEdgeVmEnableExternalSsh"] is True
    assert vals[b"aziotEdgeVmDiskSize"] == b"4Gi" and vals[b"n"] == 12 and vals[b"z"] == b"012"
    assert vals[b"a"][b"b"] == b"c,d"
