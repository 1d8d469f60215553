"""kv_b64enc_batch_dev_static: the encode entry point for callers that cannot allocate.

Some embedders have no way to give the library scratch memory -- all HBM already belongs to
another pool, or they sit on a path where cudaMalloc is forbidden -- so this variant keeps
its tile table in a fixed array inside the library.  The test calls it through the C ABI
twice in a row (the second batch has a different shape, so anything left over from the
first call would show) and requires byte-identical output to kv_b64enc_batch_dev and to the
CPU oracle, and that the free device memory is the same after the calls as before them (the
first call on a device is made beforehand: that is where the driver loads the library's
device code and the array itself).  It lives in a file that pytest collects
last so that a run stopping at the first failure has already run everything else.
"""
import numpy as np
import pytest

from kvedge_b200 import _lib, b64, workload
from oracle import b64 as ob64

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
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    b64.b64enc_device_static(d_in, d_in_off, got, d_out_off, total_in_bytes=len(blob))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    return ref, got, out_off, free_before - free_after


def _warm_up(torch):
    """What an embedder does at start-up: one call with a single empty unit.  The driver loads the
    library's device image (code and the 8 MiB array) here, and the kernels' launch set-up is done."""
    off = torch.zeros(2, dtype=torch.int64, device="cuda")
    dummy = torch.zeros(16, dtype=torch.uint8, device="cuda")
    b64.b64enc_device_static(dummy, off, dummy, off, total_in_bytes=0)
    torch.cuda.synchronize()


def test_static_scratch_entry_matches_the_allocating_one_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    _warm_up(torch)
    # first call: many small units; second call, straight after: fewer, larger units; then a tiny
    # batch, which takes the other encode kernel
    for seed, n_units, max_len in ((1, 700, 6000), (2, 90, 40000), (3, 4, 900)):
        blob, in_off = _batch(seed, n_units, max_len)
        ref, got, out_off, hbm_taken = _encode_both(torch, blob, in_off)
        assert torch.equal(ref, got)          # byte work: exact, no tolerance
        assert int((got != 0x7E).sum()) > 0
        assert hbm_taken == 0, f"free device memory shrank by {hbm_taken} bytes across the call"
        want = ob64.encode_batch(blob, in_off, out_off)
        raw = got.cpu().numpy()
        for i in range(len(in_off) - 1):
            a, b = int(out_off[i]), int(out_off[i]) + b64.encoded_len(int(in_off[i + 1] - in_off[i]))
            assert np.array_equal(raw[a:b], want[a:b]), (seed, i)


def test_static_scratch_entry_refuses_batches_beyond_its_limit():
    # 400k units need a ~12.8 MB tile table, more than the 8 MiB the library carries
    L = _lib.lib()
    rc = L.kv_b64enc_batch_dev_static(None, 1, 400_000, 0, None, 1, None)
    assert rc == _lib.KV_ENOMEM
    msg = L.kv_last_error().decode()
    assert "too large" in msg and str(_lib.STATIC_SCRATCH_BYTES) in msg
