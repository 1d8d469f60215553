"""GPU parity of the whole drop-in path: chart -> C++ engine -> pending b64enc sites encoded
on the GPU -> `helm template` stream, against the oracle's render of the same chart.
These run on the repo's own fixture chart (range/with/maps/lists, immediate b64enc sites); the
reference's chart itself runs through the same path in tests/test_gpu_reference_chart.py."""
import os

import pytest

from kvedge_b200 import helm, workload
from oracle import b64 as ob64
from oracle import helmref

from conftest import ROOT

pytestmark = pytest.mark.gpu
FIXTURE = os.path.join(ROOT, "tests", "fixtures", "edgevm")


def test_template_matches_oracle_config2_shape(tmp_path):
    cfg = workload.config_toml(0)
    p = tmp_path / "config.toml"
    p.write_bytes(cfg)
    sets = [f"sshKey={workload.ssh_key(0).decode()},exposeSsh=true,diskSize=4Gi,nameOverride=aziot-edge-kubevirt"]
    got = helm.template(FIXTURE, "chart-1634218954", set=sets, set_file=[f"runtimeConfig={p}"])
    ref = helmref.helm_template(FIXTURE, "chart-1634218954", set_args=sets, set_file_args=[f"runtimeConfig={p}"])
    assert got == ref
    ud = helmref.secret_userdata(got, "aziot-edge-kubevirt-config")
    assert ud == ob64.encode(cfg) and ob64.decode(ud) == cfg          # byte-exact Secret, both ways


def test_immediate_b64enc_sites_run_on_the_gpu_too():
    got = helm.template(FIXTURE, "r", set=["immediate=true,runtimeConfig=abc"])
    ref = helmref.helm_template(FIXTURE, "r", set_args=["immediate=true,runtimeConfig=abc"])
    assert got == ref
    assert b'quoted: "ZWRnZXZt"' in got and b"roundtrip: edgevm" in got


def test_placeholder_colliding_values_fall_back_to_immediate_mode():
    payload = b"\x01\x10\x11\x1f binary \x00 payload"
    r = helm.Chart(FIXTURE).plan("r", blobs={"runtimeConfig": payload})
    assert r.pending == []                              # nothing deferred ...
    ud = helmref.secret_userdata(r.manifest(), "edgevm-config")
    assert ud == ob64.encode(payload)                   # ... yet encoded correctly


def test_notes_and_release_object():
    r = helm.Chart(FIXTURE).plan("rel9", set=["runtimeConfig=x"], namespace="edge")
    assert len(r.pending) == 2
    r.encode_gpu()
    assert r.notes() == helmref.helm_notes(FIXTURE, "rel9", set_args=["runtimeConfig=x"], namespace="edge")


def test_fleet_template_one_batch_many_installs():
    n = 96
    releases = []
    for i in range(n):
        cfg = workload.config_toml(i) if i % 5 else b""          # some installs with an empty config
        releases.append((f"edge-{i:03d}", {"set": [f"nameOverride=site-{i:03d},sshKey={workload.ssh_key(i).decode()}"],
                                           "blobs": {"runtimeConfig": cfg}}))
    got = helm.Chart(FIXTURE).template_fleet(releases)
    assert len(got) == n
    ch = helmref.load_chart(FIXTURE)
    for (name, kw), g in zip(releases, got):
        vals = helmref.render_values(ch, kw["set"])
        helmref._assign(vals, ["runtimeConfig"], kw["blobs"]["runtimeConfig"])
        ms, _ = helmref.split_manifests(helmref.render_files(ch, vals, name))
        ref = b"".join(b"---\n# Source: " + m.name.encode() + b"\n" + m.content + b"\n" for m in ms)
        assert g == ref, name
