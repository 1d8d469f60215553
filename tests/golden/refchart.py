"""The reference's chart (deployment/helm, eleven files) rebuilt in a temporary directory from
tests/golden/reference_chart_files.json (written by make_golden.py), so that tests and
bench.py can render it on machines where /root/reference does not exist (the GPU box)."""
import atexit
import base64
import hashlib
import json
import os
import shutil
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))


def materialise() -> str:
    with open(os.path.join(HERE, "reference_chart_files.json")) as f:
        frozen = json.load(f)["files"]
    top = tempfile.mkdtemp(prefix="kvedge_refchart_")
    atexit.register(shutil.rmtree, top, ignore_errors=True)
    # the directory keeps the reference's own base name: it is not part of any output (Helm uses
    # Chart.yaml's name), but error messages quote it
    root = os.path.join(top, "helm")
    for ent in frozen:
        raw = base64.b64decode(ent["b64"])
        assert hashlib.sha256(raw).hexdigest() == ent["sha256"] and len(raw) == ent["len"], ent["path"]
        dst = os.path.join(root, ent["path"])
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(dst, "wb") as out:
            out.write(raw)
    return root
