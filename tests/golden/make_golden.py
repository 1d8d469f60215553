"""Regenerates the committed golden fixtures from the reference checkout (run in the build
container, where /root/reference exists):  python tests/golden/make_golden.py

What is frozen, and where it comes from:
  * notes_cast2294 ........ REAL Helm output recorded by the reference itself:
                            deployment/az-iot-edge-k8s-kubevirt-ascii.cast:2294 (the NOTES
                            block `helm install` printed on 2021-10-14).  The only real-Helm
                            vector in the reference; it pins the renderer.
  * object names .......... cast:2418, 2518, 2579 (DataVolume / VMI names seen in kubectl output).
  * cloudinit_default_b64 . the oracle's rendering of `include "aziotedgevm.cloudinit" . | b64enc`
                            (aziot-edge-vm-cloud-init-secret.yaml:6) with default values; its
                            SHA-256 and that of the decoded 2006-byte payload equal the
                            fingerprints in SURVEY.md 8c.  Restatement-derived: to be confirmed
                            against a real Helm binary.
  * stream hashes ......... SHA-256 of the oracle's `helm template` stream for BASELINE.json
                            configs 1 and 2 (restatement-derived).
  * reference_chart_files  the eleven files of deployment/helm (Chart.yaml, values.yaml, .helmignore,
                            templates/*), byte for byte, base64-wrapped with their paths and
                            SHA-256.  They are the INPUT of the path (the chart is data, not
                            program): frozen so that tests/conftest.py can materialise the
                            chart in a temporary directory on the GPU box, where
                            /root/reference does not exist, and BASELINE.json configs[0] and
                            configs[1] can run through the product path there.
  * b64_vectors ........... RFC 4648 section 10 vectors plus seeded vectors whose expected
                            outputs come from two encoders independent of this repo:
                            coreutils `base64 -w0` and CPython binascii.
Apart from the chart's data files named above, only outputs and hashes are stored.
"""
import hashlib
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import binascii  # noqa: E402

import numpy as np  # noqa: E402

from kvedge_b200 import workload  # noqa: E402
from oracle import b64 as ob64  # noqa: E402
from oracle import helmref  # noqa: E402

REF = "/root/reference"
CHART = os.path.join(REF, "deployment/helm")
CAST = os.path.join(REF, "deployment/az-iot-edge-k8s-kubevirt-ascii.cast")


def sha(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()


def coreutils_b64(data: bytes) -> bytes:
    return subprocess.run(["base64", "-w0"], input=data, capture_output=True, check=True).stdout


def main():
    cast_lines = open(CAST, "r").read().splitlines()
    ev = json.loads(cast_lines[2293])  # cast:2294 (1-based)
    install_out = ev[2].replace("\r\n", "\n")
    notes = install_out.split("NOTES:\n", 1)[1].strip()
    release = install_out.split("NAME: ", 1)[1].split("\n", 1)[0]

    names = {}
    for ln, key in ((2418, "datavolume"), (2579, "vmi")):
        names[f"cast:{ln}"] = json.loads(cast_lines[ln - 1])[2]

    # config 1: defaults
    s1 = helmref.helm_template(CHART)
    ud = helmref.secret_userdata(s1, "aziot-edge-kubevirt-vm-cloudconfig")
    payload = ob64.decode(ud)
    # config 2: --set as in README.md:60 / cast:2280-2292, --set-file config.toml (synthetic f2)
    cfg = workload.config_toml(0)
    key = workload.ssh_key(0).decode()
    tmp = os.path.join(HERE, "_config.toml")
    with open(tmp, "wb") as f:
        f.write(cfg)
    set_arg = (f"publicSshKey={key},aziotEdgeVmEnableExternalSsh=true,aziotEdgeVmDiskSize=4Gi,"
               "nameOverride=aziot-edge-kubevirt")
    s2 = helmref.helm_template(CHART, set_args=[set_arg], set_file_args=[f"azIotEdgeConfig={tmp}"])
    os.remove(tmp)
    ud2 = helmref.secret_userdata(s2, "aziot-edge-kubevirt-vm-aziotedgeconfig")
    assert ud2 == coreutils_b64(cfg) == binascii.b2a_base64(cfg, newline=False)
    ud2c = helmref.secret_userdata(s2, "aziot-edge-kubevirt-vm-cloudconfig")

    ok, lint_msgs = helmref.helm_lint(CHART)

    ref = {
        "source": "levi106/kvedge @ /root/reference (see make_golden.py docstring)",
        "release_name_cast2294": release,
        "notes_cast2294": notes,
        "cast_object_lines": names,
        "config1": {
            "stream_sha256": sha(s1), "stream_len": len(s1),
            "cloudinit_default_b64": ud.decode(),
            "cloudinit_default_b64_sha256": sha(ud), "cloudinit_default_b64_len": len(ud),
            "cloudinit_default_payload_sha256": sha(payload), "cloudinit_default_payload_len": len(payload),
            "runtime_secret_userdata": helmref.secret_userdata(s1, "aziot-edge-kubevirt-vm-aziotedgeconfig").decode(),
        },
        "config2": {
            "set": set_arg, "config_toml_seed": 0, "ssh_key_seed": 0,
            "stream_sha256": sha(s2), "stream_len": len(s2),
            "runtime_secret_userdata_sha256": sha(ud2), "runtime_secret_userdata_len": len(ud2),
            "cloudinit_b64_sha256": sha(ud2c), "cloudinit_b64_len": len(ud2c),
        },
        "lint": {"ok": ok, "messages": lint_msgs},
    }
    with open(os.path.join(HERE, "reference_chart.json"), "w") as f:
        json.dump(ref, f, indent=1, sort_keys=True)
        f.write("\n")

    # the chart itself, as test input for machines without /root/reference
    files = []
    for base, _dirs, names in sorted(os.walk(CHART)):
        for nm in sorted(names):
            full = os.path.join(base, nm)
            raw = open(full, "rb").read()
            files.append({"path": os.path.relpath(full, CHART), "sha256": sha(raw), "len": len(raw),
                          "b64": binascii.b2a_base64(raw, newline=False).decode()})
    with open(os.path.join(HERE, "reference_chart_files.json"), "w") as f:
        json.dump({"source": "levi106/kvedge deployment/helm (input data of the path; see make_golden.py)",
                   "files": files}, f, indent=1, sort_keys=True)
        f.write("\n")

    # base64 vectors, expected values from coreutils AND binascii (must agree)
    vecs = []
    for raw, enc in workload.RFC4648_VECTORS:
        vecs.append({"name": "rfc4648:" + raw.decode(), "in_hex": raw.hex(), "b64": enc.decode()})
    rng = np.random.default_rng(4648)
    for n in list(range(0, 50)) + [63, 64, 65, 255, 256, 257, 767, 768, 769, 3071, 3072, 3073, 4097]:
        raw = rng.bytes(n)
        e1, e2 = coreutils_b64(raw), binascii.b2a_base64(raw, newline=False)
        assert e1 == e2 == ob64.encode(raw)
        vecs.append({"name": f"rand:{n}", "in_hex": raw.hex(), "b64": e1.decode()})
    # every byte value, and all-ones / all-zero sextet patterns
    for name, raw in (("allbytes", bytes(range(256))), ("ff", b"\xff" * 48), ("00", b"\x00" * 48),
                      ("sextets", bytes.fromhex("00108310518720928b30d38f41149351559761969b71d79f8218a39259a7a29aabb2dbafc31cb3d35db7e39ebbf3dfbf"))):
        e1 = coreutils_b64(raw)
        assert e1 == binascii.b2a_base64(raw, newline=False) == ob64.encode(raw)
        vecs.append({"name": name, "in_hex": raw.hex(), "b64": e1.decode()})
    # named inputs f2..f8: hashes only (they are pure functions of the seed)
    named = {}
    for k, raw in workload.named_inputs(0).items():
        e1 = coreutils_b64(raw)
        assert e1 == binascii.b2a_base64(raw, newline=False) == ob64.encode(raw)
        named[k] = {"len": len(raw), "in_sha256": sha(raw), "b64_sha256": sha(e1), "b64_len": len(e1)}
    # decode error cases: offsets per Go's encoding/base64 decodeQuantum (restated in oracle/b64_ref.c)
    dec = []
    for text in [b"Zg=", b"Z", b"Zg==Zg==", b"Zm9v\nYg==", b"Zm9*", b"=Zg=", b"Zg=a", b"Zm8=\n", b"Zm8=x",
                 b"Zm9v\r\nYmFy", b"Zm9vYg", b"Zm9vY", b"Z===", b"Zm9vYmE=\n\n", b"Zm 9v", b"", b"\n", b"Zg=\n=",
                 b"Zh==", b"Zm9=", b"\xff\xfe\xfd\xfc"]:
        try:
            out = ob64.decode(text)
            dec.append({"text_hex": text.hex(), "ok": True, "out_hex": out.hex()})
        except ob64.CorruptInputError as e:
            dec.append({"text_hex": text.hex(), "ok": False, "offset": e.offset, "partial_hex": e.partial.hex()})
    with open(os.path.join(HERE, "b64_vectors.json"), "w") as f:
        json.dump({"encode": vecs, "named": named, "decode": dec}, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote reference_chart.json, b64_vectors.json")


if __name__ == "__main__":
    main()
