"""Host logic of the render path, no GPU needed: the C++ engine behind the C ABI plans a
release (b64enc sites pending), the test fills the pending sites with the CPU checker's
encodings, and the spliced stream must equal the oracle's render byte for byte."""
import hashlib
import os

import pytest
from hypothesis import given, settings, strategies as st

from kvedge_b200 import helm, workload
from oracle import b64 as ob64
from oracle import helmref

from conftest import REF_CHART, ROOT

FIXTURE = os.path.join(ROOT, "tests", "fixtures", "edgevm")


def _plan_and_fill(chart_dir, release="release-name", **kw):
    r = helm.Chart(chart_dir).plan(release, **kw)
    for i, p in enumerate(r.pending):
        r.fill(i, ob64.encode(p))
    return r


def test_fixture_chart_matches_oracle():
    sets = ["runtimeConfig=hello world,sshKey=ssh-rsa AAAA+/= u@h", "extraLabels.team=iot,ports[0].port=2222"]
    r = _plan_and_fill(FIXTURE, "rel1", set=sets)
    ref = helmref.helm_template(FIXTURE, "rel1", set_args=sets)
    assert r.manifest() == ref
    assert r.notes() == helmref.helm_notes(FIXTURE, "rel1", set_args=sets)
    assert b'big: "1.234567e+06"' in ref          # values.yaml ints are float64 in Helm
    assert b"disabled-extra" not in ref           # .helmignore glob


def test_fixture_pending_sites_are_the_b64enc_inputs():
    cfg = workload.config_toml(3)
    r = helm.Chart(FIXTURE).plan("rel2", blobs={"runtimeConfig": cfg})
    pend = r.pending
    assert pend[0] == cfg                                   # --set-file content passes verbatim
    assert pend[1].startswith(b"#cloud-config\nhostname: edgevm\n")
    with pytest.raises(helm.KvError, match="still pending"):
        r.manifest()
    with pytest.raises(helm.KvError, match="encoded length"):
        r.fill(0, b"AAAA")


def test_empty_b64enc_input_is_not_deferred():
    r = helm.Chart(FIXTURE).plan("rel3")
    assert len(r.pending) == 1          # runtimeConfig == "" encodes to "" on the spot
    r.fill(0, ob64.encode(r.pending[0]))
    assert b"  userdata: \n" in r.manifest()


def test_set_file_from_disk_and_binary_payload(tmp_path):
    raw = bytes(c for c in workload.named_inputs(0)["f4"] if c != 1 and not 0x10 <= c < 0x20)
    p = tmp_path / "config.bin"
    p.write_bytes(raw)
    r = _plan_and_fill(FIXTURE, "rel4", set_file=[f"runtimeConfig={p}"])
    ref = helmref.helm_template(FIXTURE, "rel4", set_file_args=[f"runtimeConfig={p}"])
    assert r.manifest() == ref
    ud = helmref.secret_userdata(r.manifest(), "edgevm-config")
    assert ob64.decode(ud) == raw


def test_values_containing_placeholder_bytes_disable_deferral():
    # control bytes that collide with the placeholder protocol force immediate evaluation,
    # which needs the GPU: without one the engine must fail loudly, never guess
    from kvedge_b200 import _lib
    if _lib.lib().kv_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(helm.TemplateError, match="no CPU fallback"):
        helm.Chart(FIXTURE).plan("rel5", blobs={"runtimeConfig": b"\x01\x10binary"})


def test_template_errors_surface_like_helm(tmp_path):
    d = tmp_path / "t"
    (d / "templates").mkdir(parents=True)
    (d / "Chart.yaml").write_text("apiVersion: v2\nname: t\nversion: 0.1.0\n")
    (d / "values.yaml").write_text("flag: true\n")
    (d / "templates" / "a.yaml").write_text("{{ if eq .Values.flag true }}x: 1{{ end }}\n")
    assert helm.Chart(str(d)).plan("r").manifest() == b"---\n# Source: t/templates/a.yaml\nx: 1\n"
    with pytest.raises(helm.TemplateError, match="incompatible types"):
        helm.Chart(str(d)).plan("r", set_string=["flag=true"])     # eq bool vs string, as in Helm
    (d / "templates" / "b.yaml").write_text("{{ .Values.a.b }}\n")
    with pytest.raises(helm.TemplateError, match="nil pointer"):
        helm.Chart(str(d)).plan("r")
    (d / "templates" / "b.yaml").write_text("{{ if }}\n")
    with pytest.raises(helm.TemplateError):
        helm.Chart(str(d))


def test_lint_fixture_and_broken_chart(tmp_path):
    assert helm.lint(FIXTURE) == helmref.helm_lint(FIXTURE) == (True, [])
    d = tmp_path / "bad"
    (d / "templates").mkdir(parents=True)
    (d / "Chart.yaml").write_text("apiVersion: v9\nname: bad\nversion: one\n")
    (d / "templates" / "a.yaml").write_text("apiVersion: v1\nkind: ConfigMap\nmetadata:\n  name: Not_Valid\n")
    ok, msgs = helm.lint(str(d))
    assert (ok, msgs) == helmref.helm_lint(str(d))
    assert not ok and any("SemVer" in m for m in msgs) and any("naming" in m for m in msgs)


@settings(max_examples=60, deadline=None)
@given(st.text(alphabet=st.characters(min_codepoint=32, max_codepoint=126, blacklist_characters=",\\{}[]="), max_size=60),
       st.binary(max_size=300))
def test_property_engines_agree_on_random_overrides(name, payload):
    payload = bytes(c for c in payload if c != 1 and not (0x10 <= c < 0x20))
    sets = [f"nameOverride={name}"]
    # oracle side: inject the same blob through its values API
    ch = helmref.load_chart(FIXTURE)
    vals = helmref.render_values(ch, sets)
    helmref._assign(vals, ["sshKey"], payload)
    try:
        files = helmref.render_files(ch, vals, "p")
    except helmref.TemplateError as e:
        # `--set nameOverride=1` is an int64 and `true` a bool: `| trunc 24` refuses them in Helm
        assert "expected string" in str(e)
        with pytest.raises(helm.TemplateError, match="expected string"):
            helm.Chart(FIXTURE).plan("p", set=sets, blobs={"sshKey": payload})
        return
    r = helm.Chart(FIXTURE).plan("p", set=sets, blobs={"sshKey": payload})
    ms, _ = helmref.split_manifests(files)
    ref = b"".join(b"---\n# Source: " + m.name.encode() + b"\n" + m.content + b"\n" for m in ms)
    for i, p in enumerate(r.pending):
        r.fill(i, ob64.encode(p))
    assert r.manifest() == ref


# ---- the reference chart itself (build container only) -------------------------------------
def test_reference_chart_config1_matches_oracle_and_golden(golden_chart):
    r = _plan_and_fill(REF_CHART)
    m = r.manifest()
    assert m == helmref.helm_template(REF_CHART)
    g = golden_chart["config1"]
    assert len(m) == g["stream_len"] and hashlib.sha256(m).hexdigest() == g["stream_sha256"]
    assert len(r.pending) == 1 and hashlib.sha256(r.pending[0]).hexdigest() == g["cloudinit_default_payload_sha256"]
    assert helm.lint(REF_CHART) == (golden_chart["lint"]["ok"], golden_chart["lint"]["messages"])


def test_reference_chart_config2_matches_golden(golden_chart, tmp_path):
    g = golden_chart["config2"]
    cfg = workload.config_toml(g["config_toml_seed"])
    p = tmp_path / "config.toml"
    p.write_bytes(cfg)
    r = _plan_and_fill(REF_CHART, set=[g["set"]], set_file=[f"azIotEdgeConfig={p}"])
    m = r.manifest()
    assert hashlib.sha256(m).hexdigest() == g["stream_sha256"] and len(m) == g["stream_len"]
    assert r.pending[0] == cfg                      # a1: the file's bytes, verbatim (values.yaml:14)
    assert r.pending[1].startswith(b"#cloud-config\nhostname: iotedgevm\n")   # a3: _helper.tpl:31-75
    assert r.pending[1].endswith(b"sudo iotedge config apply")


def test_reference_notes_equal_cast2294(golden_chart):
    r = _plan_and_fill(REF_CHART, golden_chart["release_name_cast2294"], set=["nameOverride=aziot-edge-kubevirt"])
    assert r.notes().decode() == golden_chart["notes_cast2294"]


def test_reference_name_helper_edge_cases():
    for override in ["a" * 39 + "-bcdef", "", "x-", "-" * 45, "é" * 25]:
        sets = [f"nameOverride={override}"]
        r = _plan_and_fill(REF_CHART, set=sets)
        assert r.manifest() == helmref.helm_template(REF_CHART, set_args=sets), override


@pytest.mark.parametrize("byte", [0x01] + list(range(0x10, 0x20)))
def test_placeholder_byte_anywhere_in_a_value_disables_deferral(byte):
    """The scan for placeholder bytes reads eight bytes per step: every such byte must be found
    at every position of a word and in the ragged tail, or a user payload could be mistaken for
    a pending site."""
    from kvedge_b200 import _lib
    have_gpu = _lib.lib().kv_device_count() > 0
    for length in (1, 7, 8, 9, 15, 16, 17, 63, 64, 65, 1000):
        for pos in sorted({0, 1, 6, 7, 8, 9, length // 2, length - 2, length - 1} & set(range(length))):
            payload = bytearray(b"x" * length)
            payload[pos] = byte
            if not have_gpu:
                # immediate evaluation needs the GPU: the loud failure IS the detection
                with pytest.raises(helm.TemplateError, match="no CPU fallback"):
                    helm.Chart(FIXTURE).plan("rel", blobs={"runtimeConfig": bytes(payload)})
                continue
            r = helm.Chart(FIXTURE).plan("rel", blobs={"runtimeConfig": bytes(payload)})
            # the payload's own site was encoded on the spot; only the cloud-init site may still pend
            assert bytes(payload) not in r.pending, (byte, length, pos)
    # and the neighbours of the range stay deferrable
    for ok in (0x00, 0x02, 0x0F, 0x20, 0x7F, 0x80, 0x90, 0xFF):
        payload = bytes([ok]) * 24
        r = helm.Chart(FIXTURE).plan("rel", blobs={"runtimeConfig": payload})
        assert payload in r.pending, ok


# ---- string-typed Sprig functions refuse non-strings; null deletes a chart default ----------
@pytest.mark.parametrize("flag,needle", [
    ("azIotEdgeConfig=12345", "wrong type for value; expected string; got int"),
    ("azIotEdgeConfig=true", "wrong type for value; expected string; got bool"),
])
def test_b64enc_of_a_non_string_aborts_like_helm(flag, needle):
    """Sprig's b64enc is func(string) string: text/template aborts the render instead of
    encoding the printed form (aziot-edge-runtime-config-secret.yaml:6)."""
    with pytest.raises(helm.TemplateError, match=needle):
        helm.Chart(REF_CHART).plan(set=[flag])
    with pytest.raises(helmref.TemplateError, match=needle):
        helmref.helm_template(REF_CHART, set_args=[flag])
    # --set-string keeps it a string: renders, and the pending site holds the digits
    r = helm.Chart(REF_CHART).plan(set_string=[flag])
    assert r.pending[0] == flag.split("=", 1)[1].encode()


def test_null_override_deletes_the_charts_default():
    """chartutil.CoalesceValues: `--set key=null` removes a key the chart defines."""
    # nameOverride gone: the raw reference at aziot-edge-vm-cloud-init-secret.yaml:4 prints nothing
    sets = ["nameOverride=null"]
    r = _plan_and_fill(REF_CHART, set=sets)
    assert r.manifest() == helmref.helm_template(REF_CHART, set_args=sets)
    assert b"  name: -vm-cloudconfig  \n" in r.manifest()
    assert b"<nil>" not in r.manifest() and b"<no value>" not in r.manifest()
    # azIotEdgeConfig gone: b64enc receives no value and the render aborts
    with pytest.raises(helm.TemplateError, match="invalid value; expected string"):
        helm.Chart(REF_CHART).plan(set=["azIotEdgeConfig=null"])
    with pytest.raises(helmref.TemplateError, match="invalid value; expected string"):
        helmref.helm_template(REF_CHART, set_args=["azIotEdgeConfig=null"])


def test_user_lists_replace_chart_lists_and_tables_merge():
    sets = ["ports[0].port=2222,extraLabels.team=iot,extraLabels.zone=null"]
    r = _plan_and_fill(FIXTURE, "rel6", set=sets)
    m = r.manifest()
    assert m == helmref.helm_template(FIXTURE, "rel6", set_args=sets)
    assert b"9100" not in m and b"2222" in m          # the list was replaced, not patched
    assert b'team: "iot"' in m and b'tier: "edge"' in m and b"zone:" not in m


def test_multi_document_template_files_split_like_helm(tmp_path):
    """releaseutil.SplitManifests splits on `(?:^|\\s*\\n)---\\s*`: white space in front of a
    separator leaves the document before it, text behind the dashes opens the next one, and a
    leading `---` is swallowed.  IngressClass sorts in front of Ingress (Helm 3.7 InstallOrder)."""
    d = tmp_path / "multi"
    (d / "templates").mkdir(parents=True)
    (d / "Chart.yaml").write_text("apiVersion: v2\nname: multi\nversion: 0.1.0\n")
    (d / "values.yaml").write_text("v: hello\n")
    (d / "templates" / "a.yaml").write_text(
        "---\nkind: Secret\nk: {{ .Values.v | b64enc }}\n\n---\nkind: ConfigMap\nname: one\n  \n\n--- # next\n"
        "kind: Ingress\n---\nkind: IngressClass\n---   \n\n---\nkind: Pod\n")
    r = _plan_and_fill(str(d), "r")
    got = r.manifest()
    assert got == helmref.helm_template(str(d), "r")
    src = b"---\n# Source: multi/templates/a.yaml\n"
    assert got == (src + b"kind: Secret\nk: aGVsbG8=\n" + src + b"kind: ConfigMap\nname: one\n" +
                   # "---   \n\n---": the first separator eats the newlines, so the second run of dashes has no
                   # newline in front of it any more and stays in the document (Go regexp semantics)
                   src + b"---\nkind: Pod\n" +
                   src + b"kind: IngressClass\n" + src + b"# next\nkind: Ingress\n")


def test_cpu_twin_of_the_fleet_render_matches_the_oracle():
    """oracle/fleet_cpu.c (bench.py's `render_e2e` comparator: the product's engine plans, the
    checker's scalar encoder fills the b64enc sites, on host threads) yields, per install, exactly
    the oracle's `helm template` stream of the reference chart."""
    import ctypes as C
    from kvedge_b200 import _lib
    n = 24
    rel = [(f"chart-{i}", {"blobs": {"azIotEdgeConfig": workload.config_toml(i)} if i % 3 else {},
                           "set": [f"nameOverride=site-{i:03d},publicSshKey={workload.ssh_key(i).decode()}"]})
           for i in range(n)]
    chart = helm.Chart(REF_CHART)
    ovs = [helm._Overrides(**kw) for _, kw in rel]
    arr = (_lib.KvOverrides * n)(*[o.struct for o in ovs])
    names = (C.c_char_p * n)(*[nm.encode() for nm, _ in rel])
    out, off = ob64.fleet_render_cpu(_lib.lib(), chart._h, n, names, b"default", arr, 4)
    try:
        stream = C.string_at(out, int(off[-1]))
    finally:
        ob64.lib().oracle_free(out)
    ch = helmref.load_chart(REF_CHART)
    for i, (name, kw) in enumerate(rel):
        vals = helmref.render_values(ch, kw["set"])
        for k, v in kw["blobs"].items():
            helmref._assign(vals, [k], v)
        ms, _ = helmref.split_manifests(helmref.render_files(ch, vals, name))
        ref = b"".join(b"---\n# Source: " + m.name.encode() + b"\n" + m.content + b"\n" for m in ms)
        assert stream[int(off[i]):int(off[i + 1])] == ref, name


def test_fleet_render_fails_loudly_and_cleanly(tmp_path):
    """kv_fleet_template's error paths end the call (no hang, no partial output): a template
    error in one install of the fleet surfaces as TemplateError; on a machine without a CUDA
    device the call plans on the host and then refuses to go on (no CPU fallback)."""
    from kvedge_b200 import _lib
    chart = helm.Chart(REF_CHART)
    ok = [(f"r{i}", {"set": [f"nameOverride=site-{i}"], "blobs": {"azIotEdgeConfig": b"x = %d\n" % i}}) for i in range(40)]
    bad = list(ok)
    bad[17] = ("r17", {"set": ["azIotEdgeConfig=12345"]})          # b64enc of an int64: Helm aborts
    with pytest.raises(helm.TemplateError, match="expected string"):
        chart.template_fleet(bad)
    if _lib.lib().kv_device_count() == 0:
        with pytest.raises(helm.KvError, match="no CPU fallback|no CUDA device"):
            chart.template_fleet(ok)
    else:
        got = chart.template_fleet(ok)
        assert len(got) == 40 and got[3] == chart.template("r3", **ok[3][1])
    assert chart.template_fleet([]) == []
