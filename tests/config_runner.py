"""TEST TOOL: runs one workload of tests/workloads.py on cuda:<dev> through zsb_inflate_batch_device() and
checks EVERY member: the result record (ret, msg, total_in, total_out, check, data_type) against the oracle's
for its distinct member, and the output bytes [0, total_out) of every tile against the oracle's output.
Returns timing (CUDA events on the launching stream) for scripts/bench_configs.py."""
import ctypes
from concurrent.futures import ThreadPoolExecutor

import numpy as np

import orclib
import zsblib

RES_DT = np.dtype([("ret", "<i4"), ("msg", "<i4"), ("tin", "<u8"), ("tout", "<u8"), ("chk", "<u4"), ("dt", "<i4")])
MEM_DT = np.dtype([("inp", "<u8"), ("in_len", "<u8"), ("out", "<u8"), ("out_cap", "<u8"), ("wb", "<i4"), ("rsv", "<i4")])


def expected(distinct):
    """Oracle results for the distinct members: (records, outputs)."""
    chk = orclib.ref() or orclib.orc()

    def one(m):
        comp, wb, plain = m
        r, out = chk.oneshot(comp, len(plain), wb)
        return r.astuple(), out
    with ThreadPoolExecutor(16) as ex:            # the C library runs without the GIL
        got = list(ex.map(one, distinct))
    rec = np.zeros(len(distinct), dtype=RES_DT)
    for i, (t, _) in enumerate(got):
        rec[i] = t
    return rec, [o for _, o in got]


def run(w, eng, dev=0, steps=2, warmup=1, flags=None):
    import torch
    distinct, tiles = w["distinct"], w["tiles"]
    K = len(distinct)
    n = K * tiles
    rec, outs = expected(distinct)
    lens = np.array([len(m[0]) for m in distinct], dtype=np.int64)
    caps = np.array([len(m[2]) for m in distinct], dtype=np.int64)
    wbs = np.array([m[1] for m in distinct], dtype=np.int32)
    ioff = np.concatenate(([0], np.cumsum((lens + 15) & ~15)))
    ooff = np.concatenate(([0], np.cumsum(caps)))
    in_blob = np.zeros(int(ioff[-1]), dtype=np.uint8)
    exp_blob = np.zeros(int(ooff[-1]), dtype=np.uint8)
    mask = np.zeros(int(ooff[-1]), dtype=np.uint8)
    for i, m in enumerate(distinct):
        in_blob[ioff[i]:ioff[i] + lens[i]] = np.frombuffer(m[0], dtype=np.uint8)
        o = outs[i]
        exp_blob[ooff[i]:ooff[i] + len(o)] = np.frombuffer(o, dtype=np.uint8)
        mask[ooff[i]:ooff[i] + len(o)] = 0xff
    device = torch.device("cuda", dev)
    t_in = torch.from_numpy(in_blob).to(device).repeat(tiles)
    t_exp = torch.from_numpy(exp_blob).to(device)
    t_mask = torch.from_numpy(mask).to(device)
    t_out = torch.zeros(int(ooff[-1]) * tiles, dtype=torch.uint8, device=device)
    desc = np.zeros(n, dtype=MEM_DT)
    idx = np.arange(n)
    d, t = idx % K, idx // K
    desc["inp"] = t_in.data_ptr() + ioff[d] + t * int(ioff[-1])
    desc["in_len"] = lens[d]
    desc["out"] = t_out.data_ptr() + ooff[d] + t * int(ooff[-1])
    desc["out_cap"] = caps[d]
    desc["wb"] = wbs[d]
    t_desc = torch.from_numpy(desc.view(np.uint8).copy()).to(device)
    t_res = torch.zeros(n * 32, dtype=torch.uint8, device=device)
    stream = torch.cuda.current_stream(device)
    if flags is None:
        flags = zsblib.ZSB_RAW_ONLY if (wbs < 0).all() else zsblib.ZSB_HAS_WRAPPED

    def step():
        rc = eng.lib.zsb_inflate_batch_device(eng.ctx, 0, n, t_desc.data_ptr(), desc.ctypes.data, t_res.data_ptr(), None,
                                              flags, ctypes.c_void_p(stream.cuda_stream))
        if rc:
            raise RuntimeError("zsb_inflate_batch_device rc=%d %s" % (rc, eng.error()))

    for _ in range(warmup):
        step()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.set_option("profile", 1)
    e0.record(stream)
    for _ in range(steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize(device)
    ms = e0.elapsed_time(e1) / steps
    pa, pb = ctypes.c_double(0), ctypes.c_double(0)
    pairs = eng.lib.zsb_profile_read(eng.ctx, 0, ctypes.byref(pa), ctypes.byref(pb))
    eng.set_option("profile", 0)
    launches = eng.lib.zsb_last_launches(eng.ctx)

    # every member's record
    res = t_res.cpu().numpy().view(RES_DT).reshape(tiles, K)
    bad = []
    for f in RES_DT.names:
        ne = np.argwhere(res[f] != rec[f][None, :])
        if len(ne):
            tt, dd = ne[0]
            bad.append("%s: tile %d member %d got %r want %r" % (f, tt, dd, res[f][tt, dd], rec[f][dd]))
    # every member's bytes
    B = int(ooff[-1])
    for tt in range(tiles):
        diff = (t_out[tt * B:(tt + 1) * B] ^ t_exp) & t_mask
        if bool(diff.any()):
            pos = int(torch.nonzero(diff)[0])
            dd = int(np.searchsorted(ooff, pos, side="right") - 1)
            bad.append("bytes: tile %d member %d offset %d" % (tt, dd, pos - int(ooff[dd])))
            break
    sum_in, sum_out = int(lens.sum()) * tiles, int(rec["tout"].sum()) * tiles
    return {"workload": w["desc"], "members": n, "distinct": K, "ms_per_step": ms, "in_bytes": sum_in, "out_bytes": sum_out,
            "decompressed_gbs": sum_out / ms / 1e6, "in_plus_out_gbs": (sum_in + sum_out) / ms / 1e6,
            "gpu_launches_per_step": int(launches), "decode_kernels_ms_per_step": pa.value / steps,
            "k_lz_ms_per_step": pb.value / steps, "decode_lz_pairs_per_step": pairs / steps, "non_ok_members": int((rec["ret"] != 1).sum()) * tiles,
            "mismatches": bad}
