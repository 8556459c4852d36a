"""BASELINE.json configs C2, C3, C4, C5 at their full member counts on the GPU: every member's
result record and output bytes are compared with the oracle's for its distinct member (tests/batchlib.py).
The default engine path is used: lane path, members handed out largest first, members far larger than the
rest of their batch decoded by the large-member path, members of 128 KiB compressed or more in segments; C4 also
runs without the split and without the segments."""
import os

import pytest

import batchlib
import workloads
import zsblib

pytestmark = pytest.mark.gpu
SCALE = os.environ.get("ZSB_CONFIG_SCALE", "full")      # "small" shrinks the batches 16x for quick runs


def _run(w, split=True, options=()):
    import torch
    eng = zsblib.Zsb(devices=[0])
    eng.set_option("split_big", 1 if split else 0)
    for k, v in options:
        eng.set_option(k, v)
    try:
        dev = torch.device("cuda", 0)
        b = batchlib.DeviceBatch(w, dev).to_device()
        st = torch.cuda.current_stream(dev)
        b.step(eng, st)
        torch.cuda.synchronize(dev)
        b.clear_outputs()
        b.step(eng, st)
        torch.cuda.synchronize(dev)
        launches = eng.lib.zsb_last_launches(eng.ctx)
        bad = b.verify()
        non_ok = int((b.rec["ret"][b.d] != 1).sum())
    finally:
        eng.close()
    assert not bad, bad
    assert launches >= 2
    return {"gpu_launches_per_step": int(launches), "non_ok_members": non_ok, "members": b.n}


def test_c2_64kib_raw_deflate_members():
    """The headline configuration: 65,536 x 64 KiB level-6 raw-deflate members, 1,024 distinct."""
    w = workloads.c2(distinct=1024, tiles=64 if SCALE == "full" else 4)
    r = _run(w)
    assert r["non_ok_members"] == 0 and r["members"] == 1024 * w["tiles"]


def test_c3_deflate64_4mib_members():
    w = workloads.c3(distinct=64 if SCALE == "full" else 8, tiles=64 if SCALE == "full" else 4)
    r = _run(w)
    assert r["non_ok_members"] == 0


def test_c4_wrapped_mixed_sizes_with_checksums():
    w = workloads.c4(distinct=512, tiles=32 if SCALE == "full" else 2)
    r = _run(w)
    assert r["non_ok_members"] > 0          # the damaged ~1 % report the oracle's error codes


def test_c4_large_members_get_their_own_path():
    w = workloads.c4(distinct=512, tiles=2)
    r = _run(w, split=True)
    assert r["gpu_launches_per_step"] >= 12  # the segment kernels (zsb_seg.cuh) and k_big beside k_lane + k_lz, then k_check_parts + k_check
    r = _run(w, split=True, options=(("segment", 0),))
    assert r["gpu_launches_per_step"] == 5   # k_big beside k_lane + k_lz, then k_check_parts + k_check
    r = _run(w, split=False, options=(("segment", 0),))
    assert r["gpu_launches_per_step"] == 4
    r = _run(w, split=False)
    assert r["gpu_launches_per_step"] >= 10


def test_c5_zip_like_mix():
    w = workloads.c5(distinct=8192 if SCALE == "full" else 512, tiles=128 if SCALE == "full" else 64)
    r = _run(w)
    assert r["non_ok_members"] == 0
