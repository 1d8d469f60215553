"""Seeded synthetic inputs for the render -> b64enc -> Secret path.

There is no network and the reference ships no sample ``config.toml`` (its demo file
is only visible as "299 lines" in ``deployment/az-iot-edge-k8s-kubevirt-ascii.cast:1803,2016``),
so tests and ``bench.py`` build inputs of that shape here.  Everything is a pure
function of the seed.

f0  empty                                   f4  12 KiB uniform random bytes (NUL, invalid UTF-8)
f1  RFC 4648 section 10 vectors             f5..f7  lengths 3k, 3k+1, 3k+2 around 12 KiB
f2  299-line TOML-like text, LF endings     f8  just under the 1 MiB Secret ceiling
f3  the same with CRLF endings
"""
from __future__ import annotations

from typing import Dict, List, Tuple

import numpy as np

RFC4648_VECTORS: List[Tuple[bytes, bytes]] = [
    (b"", b""), (b"f", b"Zg=="), (b"fo", b"Zm8="), (b"foo", b"Zm9v"),
    (b"foob", b"Zm9vYg=="), (b"fooba", b"Zm9vYmE="), (b"foobar", b"Zm9vYmFy"),
]

SECRET_CEILING = 1 << 20  # Kubernetes Secret size limit [ext]


_ALNUM = np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/", np.uint8)


def _key(rng: np.random.Generator, nbytes: int = 32) -> str:
    """A key-looking token (random characters; nothing is encoded here)."""
    n = (nbytes + 2) // 3 * 4
    s = _ALNUM[rng.integers(0, 64, n)].tobytes().decode()
    pad = (3 - nbytes % 3) % 3
    return s[:n - pad] + "=" * pad


def config_toml(seed: int = 0, lines: int = 299, crlf: bool = False) -> bytes:
    """A TOML-shaped IoT Edge ``config.toml`` stand-in: ``lines`` lines, ~12 KiB at 299."""
    rng = np.random.default_rng(seed)
    out: List[str] = []
    dev = f"edge-{int(rng.integers(0, 1 << 30)):08x}"
    out.append("# Azure IoT Edge runtime configuration (synthetic stand-in, seed %d)" % seed)
    out.append("")
    out.append("[provisioning]")
    out.append('source = "manual"')
    out.append('connection_string = "HostName=hub-%04d.azure-devices.net;DeviceId=%s;SharedAccessKey=%s"'
               % (int(rng.integers(0, 10000)), dev, _key(rng)))
    out.append("")
    sect = 0
    while len(out) < lines:
        sect += 1
        out.append("[agent.env.section_%03d]" % sect)
        for _ in range(int(rng.integers(3, 9))):
            if len(out) >= lines:
                break
            kind = int(rng.integers(0, 4))
            k = "key_%05d" % int(rng.integers(0, 100000))
            if kind == 0:
                out.append('%s = "%s"' % (k, _key(rng, int(rng.integers(24, 72)))))
            elif kind == 1:
                out.append("%s = %d" % (k, int(rng.integers(0, 1 << 31))))
            elif kind == 2:
                out.append("%s = %s" % (k, "true" if rng.integers(0, 2) else "false"))
            else:
                out.append("# %s" % " ".join("w%04x" % int(x) for x in rng.integers(0, 65536, 11)))
        if len(out) < lines:
            out.append("")
    eol = "\r\n" if crlf else "\n"
    return (eol.join(out[:lines]) + eol).encode("utf-8")


def ssh_key(seed: int = 0) -> bytes:
    """A ~390-character ``ssh-rsa AAAA... user@host`` line (cast:2280-2292 passes one in)."""
    rng = np.random.default_rng(seed + 7919)
    body = "AAAAB3NzaC1yc2E" + _key(rng, 268)[:357]
    return ("ssh-rsa %s ops@edge-%03d" % (body, seed % 1000)).encode()


def named_inputs(seed: int = 0) -> Dict[str, bytes]:
    """The f0..f8 set of SURVEY.md 8(d)."""
    rng = np.random.default_rng(seed)
    k = 12 * 1024 // 3 * 3
    d: Dict[str, bytes] = {"f0": b""}
    d["f2"] = config_toml(seed)
    d["f3"] = config_toml(seed, crlf=True)
    d["f4"] = rng.bytes(12 * 1024)
    d["f5"] = rng.bytes(k)
    d["f6"] = rng.bytes(k + 1)
    d["f7"] = rng.bytes(k + 2)
    d["f8"] = rng.bytes(SECRET_CEILING // 4 * 3 - 1024)
    return d


def pack(units: List[bytes]) -> Tuple[np.ndarray, np.ndarray]:
    """Pack byte strings into one blob + (n+1) int64 offsets (unit i = [off[i], off[i+1]))."""
    off = np.zeros(len(units) + 1, dtype=np.int64)
    np.cumsum([len(u) for u in units], out=off[1:])
    blob = np.frombuffer(b"".join(units), dtype=np.uint8).copy() if off[-1] else np.zeros(0, np.uint8)
    return blob, off


def fleet_batch(n_installs: int, seed: int = 0, distinct: int = 64):
    """The bench workload (BASELINE.json config 2 at fleet scale): ``n_installs`` installs, each
    contributing its two Secret inputs -- a ~12 KiB ``config.toml`` (runtime-config Secret)
    and the ~2.4 KB cloud-init text with a 390-char SSH key (cloud-config Secret is produced
    from the chart at render time; here only its size matters so the caller supplies it).
    To keep host set-up bounded the configs cycle through ``distinct`` seeded files whose
    order is then permuted per install.  Returns (list of config bytes per install)."""
    pool = [config_toml(seed * 1000 + i, lines=299) for i in range(distinct)]
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, distinct, n_installs)
    return [pool[int(i)] for i in pick]
