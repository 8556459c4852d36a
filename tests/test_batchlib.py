"""CPU tier: the batch tool itself (tests/batchlib.py) — layout and verification — with the oracle standing in
for the decoder, so that the -m gpu configs and bench.py can trust what verify() says."""
import ctypes

import numpy as np
import torch

import batchlib
import orclib
import workloads


def _fill_with_oracle(b):
    chk = orclib.ref() or orclib.orc()
    res = np.zeros(b.n, dtype=batchlib.RES_DT)
    base_in, base_out = b.t_in.data_ptr(), b.t_out.data_ptr()
    for j in range(b.n):
        comp = ctypes.string_at(int(b.desc["inp"][j]), int(b.desc["in_len"][j]))
        r, out = chk.oneshot(comp, int(b.desc["out_cap"][j]), int(b.desc["wb"][j]))
        res[j] = r.astuple()
        off = int(b.desc["out"][j]) - base_out
        b.t_out[off:off + len(out)] = torch.frombuffer(bytearray(out) or bytearray(1), dtype=torch.uint8)[:len(out)]
    b.t_res[:b.n * 32] = torch.from_numpy(res.view(np.uint8).copy())
    assert base_in


def test_layout_and_verification_of_a_shard():
    w = workloads.c5(distinct=24, tiles=5, lo=200, hi=5000)
    n = 24 * 5
    members = np.array([g for g in range(n) if g % 3 != 1 or g < 30], dtype=np.int64)      # a ragged share
    b = batchlib.DeviceBatch(w, torch.device("cpu"), members=members).to_device()
    assert b.n == len(members) and sorted(b.global_ids.tolist()) == members.tolist()
    # no two members share input or output bytes
    iv = sorted((int(o), int(o + l)) for o, l in zip(b.in_off, b.in_len))
    ov = sorted((int(o), int(o + c)) for o, c in zip(b.out_off, b.out_cap))
    assert all(a[1] <= c[0] for a, c in zip(iv, iv[1:])) and all(a[1] <= c[0] for a, c in zip(ov, ov[1:]))
    _fill_with_oracle(b)
    assert b.verify() == []
    # a flipped output byte and a changed record are both reported
    j = b.n // 2
    off = int(b.desc["out"][j]) - b.t_out.data_ptr() + 3
    b.t_out[off] ^= 1
    assert any(m.startswith("bytes") for m in b.verify())
    b.t_out[off] ^= 1
    b.t_res[j * 32 + 8] ^= 1                         # total_in of member j
    assert any(m.startswith("tin") for m in b.verify())
