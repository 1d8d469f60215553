"""The host-buffer entry points (kv_b64enc_batch / kv_b64dec_batch) cut a batch into chunks and
chain them through upload / kernel / download streams.  These cases force many chunks through few
slots (so every slot is reused several times) and feed host pointers at every alignment the
upload's head/body split has to deal with.  Run in a subprocess so that the small chunk size set
through kv_host_pipeline_config does not leak into the other tests."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r'''
import ctypes, sys
import numpy as np
sys.path.insert(0, %(root)r)
from kvedge_b200 import b64, _lib, workload
from oracle import b64 as ob64

rng = np.random.default_rng(5)
sizes = []
while sum(sizes) < 24 << 20:
    k = int(rng.integers(0, 4))
    sizes += [int(rng.integers(0, 3000))] * 7 if k == 0 else [int(rng.integers(3000, 200000))] if k == 1 else [int(rng.integers(1, 40))] * 20 if k == 2 else [0]
units = [rng.bytes(s) for s in sizes]
blob, in_off = workload.pack(units)
L = _lib.lib()
_lib.check(L.kv_host_pipeline_config(%(chunk_mb)d << 20, %(slots)d))
for skew_in, skew_out, align in ((0, 0, 16), (1, 0, 16), (255, 3, 4), (17, 1, 1), (256, 16, 16)):
    out_off = b64.encode_offsets(in_off, align=align)
    hin = np.zeros(len(blob) + 512, dtype=np.uint8)
    base_in = (-hin.ctypes.data) %% 256 + skew_in          # host address = 256-aligned + skew
    hin[base_in:base_in + len(blob)] = blob
    hout = np.full(int(out_off[-1]) + 512, 0x7E, dtype=np.uint8)
    base_out = (-hout.ctypes.data) %% 256 + skew_out
    _lib.check(L.kv_b64enc_batch(hin.ctypes.data + base_in, in_off.ctypes.data, len(units),
                                 hout.ctypes.data + base_out, out_off.ctypes.data))
    raw = hout[base_out:base_out + int(out_off[-1])].tobytes()
    for i, u in enumerate(units):
        a = int(out_off[i])
        assert raw[a:a + b64.encoded_len(len(u))] == ob64.encode(u), (skew_in, skew_out, align, i, len(u))
    assert (hout[:base_out] == 0x7E).all() and (hout[base_out + int(out_off[-1]):] == 0x7E).all()
# decode the last layout back through the host API, chunked the same way
enc_units = [ob64.encode(u) for u in units]
eblob, eoff = workload.pack(enc_units)
doff = b64.decode_offsets(eoff)
out, _, out_len, status = b64.b64dec_packed(eblob, eoff, doff)
assert (status == -1).all()
for i, u in enumerate(units):
    assert out[int(doff[i]):int(doff[i]) + int(out_len[i])].tobytes() == u, i
print("ok", len(units))
'''


@pytest.mark.parametrize("chunk_mb,slots", [(1, 2), (2, 3)])
def test_many_chunks_through_few_slots(chunk_mb, slots):
    script = SCRIPT % {"root": ROOT, "chunk_mb": chunk_mb, "slots": slots}
    res = subprocess.run([sys.executable, "-c", script], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout + res.stderr
    assert res.stdout.startswith("ok")
