import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = "/root/reference"
GOLDEN = os.path.join(ROOT, "tests", "golden")


sys.path.insert(0, GOLDEN)
import refchart  # noqa: E402  tests/golden/refchart.py

REF_CHART = refchart.materialise()
REF_CHART_LIVE = os.path.join(REFERENCE, "deployment", "helm")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "needs_reference: reads /root/reference (build container only)")


def pytest_collection_modifyitems(config, items):
    have_ref = os.path.isdir(REF_CHART_LIVE)
    for it in items:
        if "needs_reference" in it.keywords and not have_ref:
            it.add_marker(pytest.mark.skip(reason="/root/reference is not present on this machine"))


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure both the product library and the CPU checker exist before any test runs."""
    from kvedge_b200 import build as kbuild
    from oracle import b64 as ob64
    import shutil
    if shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc"):
        kbuild.build_native()
    ob64.build()
    yield


@pytest.fixture(scope="session")
def golden_vectors():
    import json
    with open(os.path.join(GOLDEN, "b64_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_chart():
    import json
    with open(os.path.join(GOLDEN, "reference_chart.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def go_tables():
    """Go's own encoding/base64 test tables (tests/golden/go_base64_tables.json)."""
    import json
    with open(os.path.join(GOLDEN, "go_base64_tables.json")) as f:
        return json.load(f)
