#!/usr/bin/env python
"""A longer randomized parity run than the test suite affords: random unit-size mixes, alignments
and corruptions through the device and the host entry points, every unit compared with the CPU
checker.  SECONDS (default 120) bounds the run; prints one JSON line and exits 1 on a mismatch."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from kvedge_b200 import b64, workload
from oracle import b64 as ob64

budget = float(os.environ.get("SECONDS_BUDGET", "120"))
rng = np.random.default_rng(int(os.environ.get("SEED", "1")))


def sizes(n):
    kind = int(rng.integers(0, 7))
    if kind == 0:
        return rng.integers(0, 64, n)
    if kind == 1:
        return rng.integers(0, 4000, n)
    if kind == 2:
        return rng.integers(2000, 40000, n)
    if kind == 3:
        return (rng.pareto(1.1, n) * 300).clip(0, 300000).astype(np.int64)
    if kind == 4:
        base = rng.choice([3071, 3072, 3073, 4095, 4096, 4097, 12287, 12288, 12289, 16383, 16384, 16385, 9215, 9216, 9217])
        return np.full(n, base) + rng.integers(-2, 3, n)
    if kind == 5:
        a = rng.integers(0, 3000, n); a[rng.random(n) < 0.2] = 0; a[rng.random(n) < 0.1] = rng.integers(13000, 60000)
        return a
    return np.where(rng.random(n) < 0.5, rng.integers(9000, 14000, n), 2403)


rounds = units_total = 0
t0 = time.time()
while time.time() - t0 < budget:
    n = int(rng.choice([1, 2, 7, 33, 200, 1500, 6000]))
    units = [rng.bytes(int(s)) for s in sizes(n)]
    expect = [ob64.encode(u) for u in units]
    # ---- encode, device entry, random alignment
    out_align = int(rng.choice([1, 4, 16])); in_shift = int(rng.integers(0, 20)); out_shift = int(rng.integers(0, 20)) * (1 if out_align == 1 else out_align)
    blob, in_off = workload.pack(units)
    in_off = in_off + in_shift
    out_off = b64.encode_offsets(in_off, align=out_align) + out_shift
    d_in = torch.zeros(int(in_off[-1]) + 64, dtype=torch.uint8, device="cuda")
    d_in[in_shift:in_shift + len(blob)] = torch.from_numpy(blob).cuda()
    d_out = torch.full((int(out_off[-1]) + 64,), 0x7E, dtype=torch.uint8, device="cuda")
    scratch = torch.empty(b64.scratch_bytes(n, int(in_off[-1])), dtype=torch.uint8, device="cuda")
    b64.b64enc_device(d_in, torch.from_numpy(in_off).cuda(), d_out, torch.from_numpy(out_off).cuda(), scratch, total_in_bytes=int(in_off[-1]))
    raw = d_out.cpu().numpy()
    mask = np.ones(len(raw), dtype=bool)
    for i, e in enumerate(expect):
        a = int(out_off[i])
        if raw[a:a + len(e)].tobytes() != e:
            print(json.dumps({"fail": "encode dev", "round": rounds, "unit": i, "len": len(units[i]), "n": n, "align": out_align})); sys.exit(1)
        mask[a:a + len(e)] = False
    if not (raw[mask] == 0x7E).all():
        print(json.dumps({"fail": "encode dev wrote outside its units", "round": rounds})); sys.exit(1)
    # ---- encode, host entry
    if b64.b64enc_many(units) != expect:
        print(json.dumps({"fail": "encode host", "round": rounds})); sys.exit(1)
    # ---- decode, host entry, with corruption
    texts = []
    for i, e in enumerate(expect):
        t = bytearray(e)
        k = int(rng.integers(0, 12))
        if k == 1 and len(t) > 8: t[int(rng.integers(0, len(t)))] = int(rng.choice([42, 10, 13, 61, 200, 0]))
        elif k == 2 and len(t) > 8: p = int(rng.integers(0, len(t))); t[p:p] = b"\n"
        elif k == 3 and len(t) > 4: del t[-1]
        elif k == 4 and len(t) > 4: t += b"="
        texts.append(bytes(t))
    tb, toff = workload.pack(texts)
    doff = b64.decode_offsets(toff, align=int(rng.choice([1, 4, 16])))
    out, _, out_len, status = b64.b64dec_packed(tb, toff, doff)
    r_out, r_len, r_status = ob64.decode_batch(tb, toff, doff)
    if not (np.array_equal(status, r_status) and np.array_equal(out_len, r_len)):
        bad = int(np.nonzero((status != r_status) | (out_len != r_len))[0][0])
        print(json.dumps({"fail": "decode status/len", "round": rounds, "unit": bad, "got": [int(status[bad]), int(out_len[bad])], "want": [int(r_status[bad]), int(r_len[bad])], "text_len": len(texts[bad])})); sys.exit(1)
    for i in range(n):
        a = int(doff[i])
        if not np.array_equal(out[a:a + int(out_len[i])], r_out[a:a + int(r_len[i])]):
            print(json.dumps({"fail": "decode bytes", "round": rounds, "unit": i})); sys.exit(1)
    rounds += 1; units_total += n
print(json.dumps({"ok": True, "rounds": rounds, "units": units_total, "seconds": round(time.time() - t0, 1)}))
