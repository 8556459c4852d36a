"""CPU tier: the C-ABI library loads, exports every symbol include/zsb.h declares, its pure-host entry
points work, and the GPU entry points fail loudly (never fall back) where no CUDA device exists."""
import ctypes
import os
import re

import pytest

import zsblib


def test_library_exports_declared_symbols():
    lib = zsblib.load_product()
    hdr = open(os.path.join(zsblib.ROOT, "include", "zsb.h")).read()
    declared = re.findall(r"ZSB_API[^;]*?\b(zsb_\w+)\s*\(", hdr)
    assert sorted(set(declared)) == sorted(zsblib.API_SYMBOLS)
    for name in declared + zsblib.ZLIB_SYMBOLS:
        assert hasattr(lib, name), name


def test_messages_match_reference_strings():
    lib = zsblib.load_product()
    import orclib
    o = orclib.orc().lib
    arr = (ctypes.c_char_p * 19).in_dll(o, "orc_messages")
    assert lib.zsb_message(0) is None
    for i in range(1, 19):
        assert lib.zsb_message(i) == arr[i]


def test_plan_shards_is_balanced_and_deterministic():
    lib = zsblib.load_product()
    import random
    rnd = random.Random(3)
    costs = [int(2 ** rnd.uniform(10, 24)) for _ in range(5000)]
    for ns in (1, 2, 4, 8):
        so, load = zsblib.plan_shards(lib, costs, ns)
        so2, _ = zsblib.plan_shards(lib, costs, ns)
        assert so == so2 and set(so) <= set(range(ns))
        assert sum(load) == sum(costs)
        assert max(load) - min(load) <= max(costs)
        assert max(load) <= sum(costs) / ns * 1.02 + max(costs)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(RuntimeError, match="no CUDA device|zsb_create failed"):
        zsblib.Zsb()


def test_plan_big_picks_only_outliers():
    """The length-aware split inside a GPU (zsb_plan_big): uniform batches keep every member on the lane path,
    outliers get a warp each, and never more members than two rounds of warps."""
    import random
    lib = zsblib.load_product()

    def plan(sizes):
        sizes = sorted(sizes, reverse=True)
        n = len(sizes)
        a = (ctypes.c_uint64 * n)(*[s // 2 for s in sizes])
        b = (ctypes.c_uint64 * n)(*sizes)
        return lib.zsb_plan_big(n, a, b, 148)
    assert plan([65536] * 65536) == 0                          # C2
    assert plan([65536] * 4096) == 0
    assert plan([16 << 20] + [65536] * 65535) == 1             # one outlier
    assert plan([16 << 20] * 5 + [65536] * 20000) == 5
    rnd = random.Random(4)
    c4 = [int(1024 * (16384 ** rnd.random())) for _ in range(16384)]       # 1 KiB .. 16 MiB log-uniform
    k = plan(c4)
    assert 100 < k <= 2 * 16 * 148
    c5 = [int(1024 * (256 ** rnd.random())) for _ in range(1 << 20)]       # 1 KiB .. 256 KiB
    assert plan(c5) == 0
    assert plan([]) == 0
