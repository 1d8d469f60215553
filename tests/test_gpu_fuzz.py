"""Seeded randomised parity: many random packed batches (ragged lengths, empty units, units
longer than a tile, arbitrary input shift and output alignment) through the device-resident
C ABI, bit-exact against the CPU checker; decode gets random corruption as well."""
import numpy as np
import pytest

from kvedge_b200 import b64, workload
from oracle import b64 as ob64

pytestmark = pytest.mark.gpu


def _lens(rng, n):
    kind = rng.integers(0, 4, n)
    small = rng.integers(0, 64, n)
    mid = rng.integers(64, 5000, n)
    big = rng.integers(5000, 40000, n)
    edge = rng.choice([12287, 12288, 12289, 24575, 24576, 24577, 8191, 8192, 8193, 384, 383, 385], n)
    return np.where(kind == 0, small, np.where(kind == 1, mid, np.where(kind == 2, big, edge)))


@pytest.mark.parametrize("seed", range(12))
def test_random_batches_encode(seed):
    import torch
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(1, 300))
    units = [rng.bytes(int(x)) for x in _lens(rng, n)]
    blob, in_off = workload.pack(units)
    in_shift = int(rng.integers(0, 20))
    align = int(rng.choice([1, 4, 16]))
    out_shift = int(rng.integers(0, 9)) if align == 1 else int(rng.choice([0, 4, 16]))
    in_off = in_off + in_shift
    out_off = b64.encode_offsets(in_off, align=align) + out_shift
    d_in = torch.zeros(int(in_off[-1]), dtype=torch.uint8, device="cuda")
    d_in[in_shift:] = torch.from_numpy(blob).cuda()
    d_out = torch.full((int(out_off[-1]) + 16,), 0x7E, dtype=torch.uint8, device="cuda")
    scratch = torch.empty(b64.scratch_bytes(n, int(in_off[-1])), dtype=torch.uint8, device="cuda")
    b64.b64enc_device(d_in, torch.from_numpy(in_off).cuda(), d_out, torch.from_numpy(out_off).cuda(), scratch,
                      total_in_bytes=int(in_off[-1]))
    torch.cuda.synchronize()
    raw = d_out.cpu().numpy()
    expect = np.full(len(raw), 0x7E, dtype=np.uint8)
    for i, u in enumerate(units):
        e = np.frombuffer(ob64.encode(u), np.uint8)
        expect[int(out_off[i]):int(out_off[i]) + len(e)] = e
    assert np.array_equal(raw, expect)          # every unit exact, and not one byte written elsewhere


@pytest.mark.parametrize("seed", range(12))
def test_random_batches_decode(seed):
    import torch
    rng = np.random.default_rng(2000 + seed)
    n = int(rng.integers(1, 300))
    texts = []
    for x in _lens(rng, n):
        t = bytearray(ob64.encode(rng.bytes(int(x))))
        r = rng.integers(0, 10)
        if r == 0 and len(t) > 4:
            t[int(rng.integers(0, len(t)))] = int(rng.choice([ord("*"), 0x80, 0, ord("="), ord(" ")]))
        elif r == 1 and len(t) > 4:
            p = int(rng.integers(0, len(t) + 1))
            t[p:p] = b"\n" if rng.integers(0, 2) else b"\r\n"
        elif r == 2 and len(t) > 4:
            del t[int(rng.integers(1, 4)) * -1:]
        texts.append(bytes(t))
    blob, in_off = workload.pack(texts)
    in_shift = int(rng.integers(0, 20))
    in_off = in_off + in_shift
    out_off = b64.decode_offsets(in_off, align=int(rng.choice([1, 4, 16])))
    d_in = torch.zeros(int(in_off[-1]), dtype=torch.uint8, device="cuda")
    d_in[in_shift:] = torch.from_numpy(blob).cuda()
    d_out = torch.zeros(int(out_off[-1]) + 16, dtype=torch.uint8, device="cuda")
    d_len = torch.zeros(n, dtype=torch.int64, device="cuda")
    d_st = torch.zeros(n, dtype=torch.int64, device="cuda")
    scratch = torch.empty(b64.scratch_bytes(n, int(in_off[-1])), dtype=torch.uint8, device="cuda")
    b64.b64dec_device(d_in, torch.from_numpy(in_off).cuda(), d_out, torch.from_numpy(out_off).cuda(), d_len, d_st,
                      scratch, total_in_bytes=int(in_off[-1]))
    torch.cuda.synchronize()
    full = np.zeros(int(in_off[-1]), np.uint8)
    full[in_shift:] = blob
    ref_out, ref_len, ref_st = ob64.decode_batch(full, in_off, out_off)
    assert np.array_equal(d_st.cpu().numpy(), ref_st)          # Go's CorruptInputError offsets
    assert np.array_equal(d_len.cpu().numpy(), ref_len)
    got = d_out.cpu().numpy()
    for i in range(n):
        if ref_st[i] == -1:
            a = int(out_off[i])
            assert np.array_equal(got[a:a + int(ref_len[i])], ref_out[a:a + int(ref_len[i])]), i
