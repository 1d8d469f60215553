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
