"""Pins the CPU checker (oracle/b64_ref.c) before anything is compared against it.

The reference holds no test or golden vector for its two b64enc sites (SURVEY.md 8c:
"parity unpinned"), so the pins are: RFC 4648 section 10, coreutils `base64`, CPython
binascii, the committed golden vectors, and the SURVEY.md 8c fingerprints.
"""
import binascii
import hashlib
import shutil
import subprocess

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from kvedge_b200 import workload
from oracle import b64 as ob64


def test_rfc4648_section10():
    for raw, enc in workload.RFC4648_VECTORS:
        assert ob64.encode(raw) == enc
        assert ob64.decode(enc) == raw


def test_golden_encode_vectors(golden_vectors):
    for v in golden_vectors["encode"]:
        raw = bytes.fromhex(v["in_hex"])
        assert ob64.encode(raw) == v["b64"].encode(), v["name"]
        assert ob64.decode(v["b64"].encode()) == raw, v["name"]


def test_golden_named_inputs(golden_vectors):
    named = workload.named_inputs(0)
    for k, meta in golden_vectors["named"].items():
        raw = named[k]
        assert len(raw) == meta["len"] and hashlib.sha256(raw).hexdigest() == meta["in_sha256"], \
            f"synthetic input {k} drifted from the seed"
        enc = ob64.encode(raw)
        assert len(enc) == meta["b64_len"] and hashlib.sha256(enc).hexdigest() == meta["b64_sha256"]


def test_golden_decode_cases(golden_vectors):
    for c in golden_vectors["decode"]:
        text = bytes.fromhex(c["text_hex"])
        if c["ok"]:
            assert ob64.decode(text) == bytes.fromhex(c["out_hex"])
        else:
            with pytest.raises(ob64.CorruptInputError) as ei:
                ob64.decode(text)
            assert ei.value.offset == c["offset"]
            assert ei.value.partial == bytes.fromhex(c["partial_hex"])


@pytest.mark.skipif(shutil.which("base64") is None, reason="coreutils base64 not installed")
def test_against_coreutils():
    rng = np.random.default_rng(1)
    for n in [0, 1, 2, 3, 4, 5, 57, 58, 76, 77, 1000, 12288, 12289, 12290]:
        raw = rng.bytes(n)
        enc = subprocess.run(["base64", "-w0"], input=raw, capture_output=True, check=True).stdout
        assert ob64.encode(raw) == enc
        dec = subprocess.run(["base64", "-d"], input=enc, capture_output=True, check=True).stdout
        assert dec == raw == ob64.decode(enc)


@settings(max_examples=300, deadline=None)
@given(st.binary(min_size=0, max_size=4096))
def test_property_matches_binascii_and_round_trips(raw):
    enc = ob64.encode(raw)
    assert enc == binascii.b2a_base64(raw, newline=False)
    assert len(enc) == ob64.encoded_len(len(raw))
    assert b"\n" not in enc and b" " not in enc
    assert ob64.decode(enc) == raw


@settings(max_examples=200, deadline=None)
@given(st.binary(min_size=0, max_size=600), st.lists(st.integers(0, 800), max_size=6))
def test_property_newlines_are_ignored_on_decode(raw, cuts):
    enc = bytearray(ob64.encode(raw))
    for c in sorted(cuts, reverse=True):
        c = min(c, len(enc))
        enc[c:c] = b"\r\n" if c % 2 else b"\n"
    assert ob64.decode(bytes(enc)) == raw


def test_batch_matches_single_and_threads_agree():
    rng = np.random.default_rng(3)
    units = [rng.bytes(int(n)) for n in rng.integers(0, 5000, 200)]
    blob, off = workload.pack(units)
    out_off = ob64.encode_offsets(off)
    out1 = ob64.encode_batch(blob, off, out_off, threads=1)
    out4 = ob64.encode_batch(blob, off, out_off, threads=4)
    assert np.array_equal(out1, out4)
    raw = out1.tobytes()
    for i, u in enumerate(units):
        assert raw[out_off[i]:out_off[i + 1]] == binascii.b2a_base64(u, newline=False)
    dec_off = np.zeros(len(units) + 1, dtype=np.int64)
    np.cumsum(np.diff(out_off) // 4 * 3, out=dec_off[1:])
    dec, dlen, status = ob64.decode_batch(out1, out_off, dec_off)
    assert (status == -1).all()
    for i, u in enumerate(units):
        assert dec.tobytes()[dec_off[i]:dec_off[i] + dlen[i]] == u


def test_survey_fingerprints_of_default_cloudinit_payload(golden_chart):
    """payload 2006 B / base64 2676 B with the stated SHA-256 values."""
    b64_text = golden_chart["config1"]["cloudinit_default_b64"].encode()
    assert len(b64_text) == 2676
    assert hashlib.sha256(b64_text).hexdigest() == \
        "38c73a6250c78e18f34a580cc2bc06fdf4808a4a08d55f79f4918a5298a5f628"
    payload = ob64.decode(b64_text)
    assert len(payload) == 2006
    assert hashlib.sha256(payload).hexdigest() == \
        "ef7bf5cbe95aa7e902597eff8c05371ede2610db779ff872da9c1f090c3cb37b"
    assert payload.startswith(b"#cloud-config\nhostname: iotedgevm\n")
    assert payload.endswith(b"sudo iotedge config apply")
    assert ob64.encode(payload) == b64_text


# ---- pin (vi): Go's own encoding/base64 test tables (not produced by this repo's oracle) --------
def _coreutils_decode(text: bytes):
    """coreutils `base64 -d`: (ok, bytes).  It skips '\\n' on its own; '\\r' needs -i, which would
    also hide garbage, so texts with '\\r' are not sent here."""
    res = subprocess.run(["base64", "-d"], input=text, capture_output=True)
    return res.returncode == 0, res.stdout


def test_go_pairs_and_bigtest(go_tables):
    rows = [(bytes.fromhex(p["decoded_hex"]), p["encoded"].encode()) for p in go_tables["pairs"]]
    rows.append((go_tables["bigtest"]["decoded"].encode(), go_tables["bigtest"]["encoded"].encode()))
    for raw, enc in rows:
        assert ob64.encode(raw) == enc
        assert ob64.decode(enc) == raw
        assert ob64.encoded_len(len(raw)) == len(enc)


def test_go_decode_corrupt_table(go_tables):
    """TestDecodeCorrupt: Go's published input -> CorruptInputError offset pairs (-1 = accepted).
    The consumer side of the path (aziot-edge-vm.yaml:49-51, _helper.tpl:64,73) decodes with
    exactly these semantics."""
    assert len(go_tables["decode_corrupt"]) == 24
    for c in go_tables["decode_corrupt"]:
        text = c["input"].encode()
        if c["offset"] < 0:
            out = ob64.decode(text)
            assert out == binascii.a2b_base64(text), text
            if shutil.which("base64") and b"\r" not in text:
                assert _coreutils_decode(text) == (True, out), text
        else:
            with pytest.raises(ob64.CorruptInputError) as ei:
                ob64.decode(text)
            assert ei.value.offset == c["offset"], text


def test_go_newline_characters(go_tables):
    want = go_tables["newline_characters"]["expected"].encode()
    for ex in go_tables["newline_characters"]["examples"]:
        text = ex.encode()
        assert ob64.decode(text) == want, text
        assert binascii.a2b_base64(text) == want
        if shutil.which("base64") and b"\r" not in text:
            assert _coreutils_decode(text) == (True, want), text
