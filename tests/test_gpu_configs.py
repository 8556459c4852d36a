"""BASELINE.json configs C3, C4, C5 at their full member counts on the GPU: every member's
result record and output bytes are compared with the oracle's for its distinct member (tests/config_runner.py).
The default engine path is used: lane path, members handed out largest first, members far larger than the
rest of their batch decoded by k_big (a warp each); C4 also runs without that split."""
import os

import pytest

import config_runner
import workloads
import zsblib

pytestmark = pytest.mark.gpu
SCALE = os.environ.get("ZSB_CONFIG_SCALE", "full")      # "small" shrinks the batches 16x for quick runs


def _run(w, split=True):
    eng = zsblib.Zsb(devices=[0])
    eng.set_option("split_big", 1 if split else 0)
    try:
        r = config_runner.run(w, eng, steps=1, warmup=1)
    finally:
        eng.close()
    assert not r["mismatches"], r["mismatches"]
    assert r["gpu_launches_per_step"] >= 2
    return r


def test_c3_deflate64_4mib_members():
    w = workloads.c3(distinct=8, tiles=512 if SCALE == "full" else 32)
    r = _run(w)
    assert r["non_ok_members"] == 0


def test_c4_wrapped_mixed_sizes_with_checksums():
    w = workloads.c4(distinct=512, tiles=32 if SCALE == "full" else 2)
    r = _run(w)
    assert r["non_ok_members"] > 0          # the damaged ~1 % report the oracle's error codes


def test_c4_large_members_get_a_warp_each():
    w = workloads.c4(distinct=512, tiles=2)
    r = _run(w, split=True)
    assert r["gpu_launches_per_step"] >= 4   # k_big beside k_lane + k_lz, then k_check
    r = _run(w, split=False)
    assert r["gpu_launches_per_step"] == 3


def test_c5_zip_like_mix():
    w = workloads.c5(distinct=2048, tiles=512 if SCALE == "full" else 32)
    r = _run(w)
    assert r["non_ok_members"] == 0
