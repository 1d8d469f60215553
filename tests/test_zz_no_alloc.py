"""kv_b64enc_batch_dev_static: the encode entry point for callers that cannot allocate.

Some embedders have no way to give the library scratch memory -- all HBM already belongs to
another pool, or they sit on a path where cudaMalloc is forbidden -- so this variant keeps
its tile table in a fixed array inside the library.  The test calls it through the C ABI
twice in a row (the second batch has a different shape, so anything left over from the
first call would show) and requires byte-identical output to kv_b64enc_batch_dev, which is
itself checked against the CPU oracle elsewhere.  It lives in a file that pytest collects
last so that a run stopping at the first failure has already run everything else.
"""
import numpy as np
import pytest

from kvedge_b200 import _lib, b64, workload

pytestmark = pytest.mark.gpu


def _batch(seed, n_units, max_len):
    rng = np.random.default_rng(seed)
    units = [rng.bytes(int(x)) for x in rng.integers(0, max_len, n_units)]
    units[n_units // 2] = b""                       # an empty unit in the middle
    units[-1] = workload.config_toml(seed)          # and a config-sized one at the end
    return workload.pack(units)


def _encode_both(torch, blob, in_off):
    out_off = b64.encode_offsets(in_off, align=16)
    d_in = torch.from_numpy(blob).cuda()
    d_in_off, d_out_off = torch.from_numpy(in_off).cuda(), torch.from_numpy(out_off).cuda()
    n = len(in_off) - 1
    ref = torch.full((int(out_off[-1]),), 0x7E, dtype=torch.uint8, device="cuda")
    scratch = torch.empty(b64.scratch_bytes(n, len(blob)), dtype=torch.uint8, device="cuda")
    b64.b64enc_device(d_in, d_in_off, ref, d_out_off, scratch, total_in_bytes=len(blob))
    got = torch.full((int(out_off[-1]),), 0x7E, dtype=torch.uint8, device="cuda")
    b64.b64enc_device_static(d_in, d_in_off, got, d_out_off, total_in_bytes=len(blob))
    torch.cuda.synchronize()
    return ref, got


def test_static_scratch_entry_matches_the_allocating_one_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    # first call: many small units; second call, straight after: fewer, larger units
    for seed, n_units, max_len in ((1, 700, 6000), (2, 90, 40000)):
        blob, in_off = _batch(seed, n_units, max_len)
        ref, got = _encode_both(torch, blob, in_off)
        assert torch.equal(ref, got)          # byte work: exact, no tolerance
        assert int((got != 0x7E).sum()) > 0


def test_static_scratch_entry_refuses_batches_beyond_its_limit():
    # 400k units need a ~12.8 MB tile table, more than the 8 MiB the library carries
    L = _lib.lib()
    rc = L.kv_b64enc_batch_dev_static(None, 1, 400_000, 0, None, 1, None)
    assert rc == _lib.KV_ENOMEM
    msg = L.kv_last_error().decode()
    assert "too large" in msg and str(_lib.STATIC_SCRATCH_BYTES) in msg
