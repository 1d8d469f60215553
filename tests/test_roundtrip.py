"""SURVEY.md 8(f) rows 3 and 4, CPU side: what sits on the other side of the path.

f3  consumer round trip: Kubernetes decodes Secret.data.userdata and KubeVirt exposes it as the
    file `userdata` on the disk whose serial the guest greps for (aziot-edge-vm.yaml:25-28,49-51;
    _helper.tpl:64,73) -- the decoded bytes must equal the input file, and the names/serials
    that tie the objects together must agree.
f4  cloud-config sanity: the text behind the second b64enc site parses as YAML and keeps its
    bootcmd/runcmd shape and the apt key block.
The reference chart is only present in the build container (needs_reference); the byte-level
half also runs on the fixture chart everywhere."""
import os
import re

import pytest
import yaml

from kvedge_b200 import helm, workload
from oracle import b64 as ob64
from oracle import helmref

from conftest import REF_CHART, ROOT

FIXTURE = os.path.join(ROOT, "tests", "fixtures", "edgevm")


def _render(chart_dir, **kw):
    r = helm.Chart(chart_dir).plan("rt", **kw)
    for i, p in enumerate(r.pending):
        r.fill(i, ob64.encode(p))
    return r.manifest()


def _docs(stream):
    return [d for d in re.split(rb"(?m)^---\n", stream) if d.strip()]


@pytest.mark.parametrize("name", ["f2", "f3", "f5", "f6", "f7"])
def test_fixture_secret_round_trips_to_the_input_file(name):
    raw = bytes(c for c in workload.named_inputs(0)[name] if c != 1 and not 0x10 <= c < 0x20)
    stream = _render(FIXTURE, blobs={"runtimeConfig": raw})
    ud = helmref.secret_userdata(stream, "edgevm-config")
    assert ob64.decode(ud) == raw
    assert b"\n" not in ud and b" " not in ud and len(ud) == (len(raw) + 2) // 3 * 4


def test_reference_consumer_side_round_trip_and_cross_references(tmp_path):
    cfg = workload.config_toml(0, crlf=True)          # CRLF file: must survive byte for byte
    p = tmp_path / "config.toml"
    p.write_bytes(cfg)
    stream = _render(REF_CHART, set=["nameOverride=edge-a"], set_file=[f"azIotEdgeConfig={p}"])
    s = stream.decode()
    # what the guest copies to /etc/aziot/config.toml (_helper.tpl:73)
    assert ob64.decode(helmref.secret_userdata(stream, "edge-a-vm-aziotedgeconfig")) == cfg
    # Secret names referenced by the VM == names the Secret templates produce (vm.yaml:51,57)
    assert "secretName: edge-a-vm-aziotedgeconfig" in s and "  name: edge-a-vm-aziotedgeconfig" in s
    assert re.search(r"secretRef:\n\s+name: edge-a-vm-cloudconfig ", s) and "  name: edge-a-vm-cloudconfig" in s
    # dataVolume name (vm.yaml:47) == DataVolume object (dv-container.yaml:6)
    assert "         name: edge-a-linux-dv" in s and "  name: edge-a-linux-dv" in s
    # Service selector (service.yaml:11) == VM label (vm.yaml:14)
    assert s.count("kubevirt.io/domain: edge-a-vm") == 2
    # disk serial (vm.yaml:28) == what bootcmd greps for (_helper.tpl:64)
    ci = ob64.decode(helmref.secret_userdata(stream, "edge-a-vm-cloudconfig")).decode()
    serial = re.search(r"serial: (\S+)", s).group(1)
    assert f"grep {serial} " in ci and "cp /mnt/app-secret/userdata /etc/aziot/config.toml" in ci


def test_reference_cloud_config_is_sane_yaml():
    key = workload.ssh_key(0).decode()
    stream = _render(REF_CHART, set=[f"publicSshKey={key}"])
    text = ob64.decode(helmref.secret_userdata(stream, "aziot-edge-kubevirt-vm-cloudconfig"))
    assert text.startswith(b"#cloud-config\n") and not text.endswith(b"\n")
    doc = yaml.safe_load(text)
    assert doc["hostname"] == "iotedgevm" and doc["ssh_authorized_keys"] == [key]
    assert [c.split()[0] for c in doc["bootcmd"]] == ["mkdir", "mount"]
    assert len(doc["runcmd"]) == 6 and doc["runcmd"][-1] == "sudo iotedge config apply"
    src = doc["apt"]["sources"]["microsoft-prod.list"]
    assert src["source"].startswith("deb [arch=amd64,arm64,armhf] https://packages.microsoft.com/")
    k = src["key"]
    assert k.startswith("-----BEGIN PGP PUBLIC KEY BLOCK-----") and k.rstrip().endswith("-----END PGP PUBLIC KEY BLOCK-----")
    assert all(d.strip() for d in _docs(stream))


def test_fixture_manifest_documents_parse_and_reference_each_other():
    stream = _render(FIXTURE, set=["nameOverride=site-9,runtimeConfig=x"])
    docs = [yaml.safe_load(d) for d in _docs(stream)]
    kinds = [d["kind"] for d in docs]
    assert kinds == ["Secret", "ConfigMap", "Service", "VirtualMachine"]      # Helm install order
    vm = docs[3]
    assert vm["spec"]["template"]["spec"]["volumes"][0]["secret"]["secretName"] == docs[0]["metadata"]["name"]
