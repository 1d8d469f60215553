"""examples/secret_userdata.c: the C ABI used from plain C (compiled with gcc against the header
and the in-tree library), checked against coreutils `base64 -w0` -- BASELINE.json config 2's
byte-exact Secret regression, end to end without Python in the product path."""
import os
import shutil
import subprocess

import pytest

from kvedge_b200 import workload
from kvedge_b200.build import build_native

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_program_matches_coreutils_base64(tmp_path):
    lib = build_native()
    exe = tmp_path / "secret_userdata"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "secret_userdata.c"), "-o", str(exe),
                    "-L", os.path.dirname(lib), "-lkvedge_b200", f"-Wl,-rpath,{os.path.dirname(lib)}"], check=True)
    files = []
    for name, data in workload.named_inputs(0).items():
        p = tmp_path / f"{name}.bin"
        p.write_bytes(data)
        files.append(str(p))
    res = subprocess.run([str(exe)] + files, capture_output=True, timeout=120)
    assert res.returncode == 0, res.stderr.decode()
    lines = res.stdout.decode().splitlines()
    assert len(lines) == len(files)
    assert shutil.which("base64"), "coreutils base64 is the independent encoder of this test"
    for path, line in zip(files, lines):
        ref = subprocess.run(["base64", "-w0", path], capture_output=True, check=True).stdout.decode()
        assert line == "  userdata: " + ref, path


def test_c_fleet_template_matches_single_renders_and_coreutils(tmp_path):
    """examples/fleet_template.c: kv_fleet_template from plain C on the reference's chart; every
    install's stream equals the single-install render, and the Secret payload equals coreutils
    `base64 -w0` of the config file (BASELINE.json configs[1] through the C ABI alone)."""
    from kvedge_b200 import helm
    from oracle import helmref
    from conftest import REF_CHART
    lib = build_native()
    exe = tmp_path / "fleet_template"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "examples", "fleet_template.c"), "-o", str(exe),
                    "-L", os.path.dirname(lib), "-lkvedge_b200", f"-Wl,-rpath,{os.path.dirname(lib)}"], check=True)
    cfg = workload.config_toml(7)
    p = tmp_path / "config.toml"
    p.write_bytes(cfg)
    n = 40
    res = subprocess.run([str(exe), REF_CHART, str(p), str(n)], capture_output=True, timeout=120)
    assert res.returncode == 0, res.stderr.decode()
    ref_b64 = subprocess.run(["base64", "-w0", str(p)], capture_output=True, check=True).stdout
    out, pos = res.stdout, 0
    chart = helm.Chart(REF_CHART)
    for i in range(n):
        head, _, rest = out[pos:].partition(b"\n")
        tag, idx, size = head.split()[1:]
        assert tag == b"install" and int(idx) == i
        start = pos + len(head) + 1
        stream = out[start:start + int(size)]
        pos = start + int(size)
        sets = [f"nameOverride=site-{i},aziotEdgeVmEnableExternalSsh={'true' if i % 2 else 'false'}"]
        assert stream == chart.template(f"chart-{1634218954 + i}", set=sets, blobs={"azIotEdgeConfig": cfg}), i
        assert helmref.secret_userdata(stream, f"site-{i}-vm-aziotedgeconfig") == ref_b64
        assert (b"kind: Service" in stream) == bool(i % 2)       # aziot-edge-vm-service.yaml:1
    assert pos == len(out)
