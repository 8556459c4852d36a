#!/usr/bin/env python
"""Diagnostic: pace of ONE large member through k_big (option split_big) against a lane of k_lane."""
import ctypes, os, sys, time, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import gen, workloads, zsblib, random

def run(comp, plain, wb, split, count=1):
    eng = zsblib.Zsb(devices=[0])
    eng.set_option("split_big", split)
    n = count
    t_in = torch.frombuffer(bytearray(comp), dtype=torch.uint8).cuda()
    t_out = torch.zeros(len(plain) * n, dtype=torch.uint8, device="cuda")
    desc = np.zeros(n, dtype=np.dtype([("inp", "<u8"), ("in_len", "<u8"), ("out", "<u8"), ("out_cap", "<u8"), ("wb", "<i4"), ("fl", "<i4")]))
    desc["inp"] = t_in.data_ptr(); desc["in_len"] = len(comp)
    desc["out"] = t_out.data_ptr() + np.arange(n) * len(plain); desc["out_cap"] = len(plain); desc["wb"] = wb
    t_desc = torch.from_numpy(desc.view(np.uint8).copy()).cuda()
    t_res = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream()
    def step():
        rc = eng.lib.zsb_inflate_batch_device(eng.ctx, 0, n, t_desc.data_ptr(), desc.ctypes.data, t_res.data_ptr(), None, 4, ctypes.c_void_p(st.cuda_stream))
        assert rc == 0, eng.error()
    step(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.set_option("profile", 1)
    e0.record(); step(); e1.record(); torch.cuda.synchronize()
    a, b = ctypes.c_double(0), ctypes.c_double(0)
    eng.lib.zsb_profile_read(eng.ctx, 0, ctypes.byref(a), ctypes.byref(b))
    global last_split
    last_split = (a.value, b.value)
    ok = t_out[:len(plain)].cpu().numpy().tobytes() == plain
    eng.close()
    return e0.elapsed_time(e1), ok

pool = workloads.soup_pool(64)
rnd = random.Random(1)
plain = workloads.soup_bytes(int(os.environ.get('BIG_MIB', '8')) << 20, rnd, pool)
CASES = [("soup level 6", gen.deflate(plain, 6, -15), -15), ("soup Z_FIXED", gen.deflate(plain, 6, -15, zlib.Z_FIXED), -15),
         ("soup level 1", gen.deflate(plain, 1, -15), -15)]
COUNTS = [int(x) for x in os.environ.get("BIG_COUNTS", "1,512").split(",")]
SPLITS = [int(x) for x in os.environ.get("BIG_SPLITS", "1,0").split(",")]
for name, comp, wb in CASES[:int(os.environ.get("BIG_CASES", "3"))]:
    for split in SPLITS:
        for count in COUNTS:
            ms, ok = run(comp, plain, wb, split, count)
            print("%-14s %s x%-4d: %8.1f ms  %6.1f MB/s out per member  ok=%s decode %.1f ms, k_lz %.1f ms" % (name, "k_big " if split else "k_lane", count, ms, len(plain) / ms / 1e3, ok, last_split[0], last_split[1]), flush=True)
