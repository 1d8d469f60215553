"""Small, alignment-heavy encode/decode run for compute-sanitizer memcheck."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from kvedge_b200 import b64, workload
from oracle import b64 as ob64
rng = np.random.default_rng(0)
ok = True
for in_shift, out_shift, align in [(0, 0, 16), (1, 0, 4), (3, 5, 1), (0, 4, 4)]:
    units = [rng.bytes(n) for n in list(range(0, 30)) + [3071, 12288, 12289, 30000]]
    blob, in_off = workload.pack(units)
    in_off = in_off + in_shift
    out_off = b64.encode_offsets(in_off, align=align) + out_shift
    d_in = torch.zeros(int(in_off[-1]), dtype=torch.uint8, device="cuda")   # exact size: no slack
    d_in[in_shift:] = torch.from_numpy(blob).cuda()
    d_out = torch.zeros(int(out_off[-1]), dtype=torch.uint8, device="cuda")
    scratch = torch.empty(b64.scratch_bytes(len(units), int(in_off[-1])), dtype=torch.uint8, device="cuda")
    b64.b64enc_device(d_in, torch.from_numpy(in_off).cuda(), d_out, torch.from_numpy(out_off).cuda(), scratch,
                      total_in_bytes=int(in_off[-1]))
    torch.cuda.synchronize()
    raw = d_out.cpu().numpy().tobytes()
    for i, u in enumerate(units):
        ok &= raw[int(out_off[i]):int(out_off[i]) + b64.encoded_len(len(u))] == ob64.encode(u)
    # decode the same thing back through the host API
    ok &= b64.b64dec_many([ob64.encode(u) for u in units]) == units
print("memcheck workload ok:", ok)
sys.exit(0 if ok else 1)
