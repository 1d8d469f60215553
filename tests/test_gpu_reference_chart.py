"""BASELINE.json configs[0] and configs[1] on the GPU, on the reference's OWN chart.

The chart (deployment/helm, eleven files) is materialised from
tests/golden/reference_chart_files.json (conftest.REF_CHART), so these run on the B200 box.
Everything here goes through the product path -- C++ engine -> pending `b64enc` sites encoded
by the CUDA kernels -> splice -> `helm template` framing -- and is compared with the committed
goldens (stream SHA-256s, the real-Helm NOTES text of cast:2294, the lint verdict) and with
the oracle's render; nothing is filled in from the oracle.

Quirks of the reference chart these exercise (paths relative to deployment/helm/templates):
  _helper.tpl:20-26     labels helper: comment containing {{ }}, leading newline, `indent`
  aziot-edge-vm.yaml:57 an action evaluated inside a YAML comment
  aziot-edge-vm-cloud-init-secret.yaml:4  raw .Values.nameOverride, not the name helper
  ../.helmignore:25     drops aziot-edge-data-volume-disk.yaml
"""
import hashlib

import pytest

from kvedge_b200 import helm, workload
from oracle import b64 as ob64
from oracle import helmref

from conftest import REF_CHART

pytestmark = pytest.mark.gpu


def _sha(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()


def test_config0_lint_and_template_with_defaults(golden_chart):
    """configs[0]: helm lint + helm template of deployment/helm with default values."""
    assert helm.lint(REF_CHART) == (golden_chart["lint"]["ok"], golden_chart["lint"]["messages"])
    assert helm.lint(REF_CHART) == (True, ["[INFO] Chart.yaml: icon is recommended"])
    got = helm.template(REF_CHART)                      # kv_helm_template: GPU encode inside
    g = golden_chart["config1"]
    assert len(got) == g["stream_len"] == 5930
    assert _sha(got) == g["stream_sha256"]
    assert got == helmref.helm_template(REF_CHART)
    ud = helmref.secret_userdata(got, "aziot-edge-kubevirt-vm-cloudconfig")
    assert ud.decode() == g["cloudinit_default_b64"] and _sha(ud) == g["cloudinit_default_b64_sha256"]
    assert helmref.secret_userdata(got, "aziot-edge-kubevirt-vm-aziotedgeconfig") == b""   # values.yaml:14
    assert got.count(b"kind: DataVolume") == 1          # .helmignore:25
    assert b"aziot-edge-data-volume-disk.yaml" not in got


def test_config1_set_file_secret_is_byte_exact(golden_chart, tmp_path):
    """configs[1]: helm template --set-file azIotEdgeConfig=config.toml, byte-exact Secret."""
    g = golden_chart["config2"]
    cfg = workload.config_toml(g["config_toml_seed"])
    p = tmp_path / "config.toml"
    p.write_bytes(cfg)
    got = helm.template(REF_CHART, set=[g["set"]], set_file=[f"azIotEdgeConfig={p}"])
    assert len(got) == g["stream_len"] == 20279
    assert _sha(got) == g["stream_sha256"]
    ud = helmref.secret_userdata(got, "aziot-edge-kubevirt-vm-aziotedgeconfig")
    assert len(ud) == g["runtime_secret_userdata_len"] and _sha(ud) == g["runtime_secret_userdata_sha256"]
    assert ud == ob64.encode(cfg) and ob64.decode(ud) == cfg
    udc = helmref.secret_userdata(got, "aziot-edge-kubevirt-vm-cloudconfig")
    assert len(udc) == g["cloudinit_b64_len"] and _sha(udc) == g["cloudinit_b64_sha256"]
    # the same through plan -> encode_gpu -> manifest (the split form of the call)
    r = helm.Chart(REF_CHART).plan(set=[g["set"]], set_file=[f"azIotEdgeConfig={p}"])
    assert r.pending[0] == cfg
    r.encode_gpu()
    assert r.manifest() == got


def test_notes_equal_cast2294(golden_chart):
    got = helm.notes(REF_CHART, golden_chart["release_name_cast2294"], set=["nameOverride=aziot-edge-kubevirt"])
    assert got.decode() == golden_chart["notes_cast2294"]


def test_reference_chart_quirks_survive_the_gpu_path():
    got = helm.template(REF_CHART, set=["nameOverride=edge-q"]).decode()
    assert "  name: edge-q-vm-cloudconfig  \n" in got   # cloud-init-secret.yaml:4: raw nameOverride, trailing blanks
    # aziot-edge-vm.yaml:57: the action inside the trailing YAML comment is evaluated like any other
    assert ("name: edge-q-vm-cloudconfig # TODO: we should be using edge-q here instead but that fails, "
            "needs further investigation.") in got
    # _helper.tpl:20-26 through `indent`: the label block keeps its leading newline
    assert "app.kubernetes.io/managed-by: Helm" in got
    assert got == helmref.helm_template(REF_CHART, set_args=["nameOverride=edge-q"]).decode()
    for override in ["a" * 39 + "-bcdef", "", "x-", "-" * 45]:
        sets = [f"nameOverride={override}"]
        assert helm.template(REF_CHART, set=sets) == helmref.helm_template(REF_CHART, set_args=sets), override


def test_fleet_of_the_reference_chart_equals_per_install_renders(golden_chart):
    """kv_fleet_template over 320 installs (README.md:72: one install per nameOverride): every
    stream equals the single-install render of the same values, and installs with default
    values equal the config-1 golden except for their names."""
    n = 320
    releases = []
    for i in range(n):
        sets = [f"nameOverride=site-{i:04d}"]
        blobs = {}
        if i % 4:
            sets[0] += f",publicSshKey={workload.ssh_key(i).decode()}"
        if i % 3:
            blobs["azIotEdgeConfig"] = workload.config_toml(i)
        if i % 7 == 0:
            sets[0] += ",aziotEdgeVmEnableExternalSsh=false"
        releases.append((f"chart-{1634218954 + i}", {"set": sets, "blobs": blobs}))
    chart = helm.Chart(REF_CHART)
    got = chart.template_fleet(releases)
    assert len(got) == n
    ch = helmref.load_chart(REF_CHART)
    for i, ((name, kw), g) in enumerate(zip(releases, got)):
        if i % 16 == 0:         # oracle render (slow, Python) on a sample ...
            vals = helmref.render_values(ch, kw["set"])
            for k, v in kw["blobs"].items():
                helmref._assign(vals, [k], v)
            ms, _ = helmref.split_manifests(helmref.render_files(ch, vals, name))
            ref = b"".join(b"---\n# Source: " + m.name.encode() + b"\n" + m.content + b"\n" for m in ms)
            assert g == ref, name
        # ... and the product's own single-install call on every install
        assert g == chart.template(name, **kw), name
        if kw["blobs"]:
            ud = helmref.secret_userdata(g, f"site-{i:04d}-vm-aziotedgeconfig")
            assert ob64.decode(ud) == kw["blobs"]["azIotEdgeConfig"]
    # an all-defaults fleet: each stream is the config-1 golden
    got = chart.template_fleet([("release-name", {})] * 256)
    assert {_sha(g) for g in got} == {golden_chart["config1"]["stream_sha256"]}
