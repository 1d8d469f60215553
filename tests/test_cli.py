"""`python -m kvedge_b200 lint|template`: the two Helm verbs of BASELINE.json configs 1-2."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

FIXTURE = os.path.join(ROOT, "tests", "fixtures", "edgevm")


def _run(*args):
    return subprocess.run([sys.executable, "-m", "kvedge_b200", *args], cwd=ROOT, capture_output=True, timeout=300)


def test_lint_cli_passes_on_the_fixture_chart():
    r = _run("lint", FIXTURE)
    assert r.returncode == 0, r.stderr
    assert b"1 chart(s) linted, 0 chart(s) failed" in r.stdout


def test_lint_cli_fails_on_a_broken_chart(tmp_path):
    d = tmp_path / "bad"
    (d / "templates").mkdir(parents=True)
    (d / "Chart.yaml").write_text("apiVersion: v2\nname: bad\n")
    r = _run("lint", str(d))
    assert r.returncode == 1 and b"[ERROR] Chart.yaml: version is required" in r.stdout


@pytest.mark.gpu
def test_template_cli_matches_the_oracle(tmp_path):
    from kvedge_b200 import workload
    from oracle import helmref
    cfg = tmp_path / "config.toml"
    cfg.write_bytes(workload.config_toml(1))
    args = ["--set", "nameOverride=cli-test,exposeSsh=false", "--set-file", f"runtimeConfig={cfg}"]
    r = _run("template", "rel-cli", FIXTURE, *args)
    assert r.returncode == 0, r.stderr
    assert r.stdout == helmref.helm_template(FIXTURE, "rel-cli", set_args=["nameOverride=cli-test,exposeSsh=false"],
                                             set_file_args=[f"runtimeConfig={cfg}"])
