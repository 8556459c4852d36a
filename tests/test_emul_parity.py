"""CPU tier: the decoders' DEVICE source (zlib-streams_b200/csrc/zsb_device.cuh = group path,
zsb_lane.cuh + zsb_lz.cuh = lane path), compiled for the host by tests/emul with a warp of one lane,
against the oracle: every observable field and every output byte.  The same cases run through the
real kernels on the GPU in test_gpu_parity.py."""
import hashlib

import pytest

import cases
import orclib
import zsblib

CHK = orclib.ref() or orclib.orc()
EM = zsblib.emul()


PATHS = pytest.mark.parametrize("lane", [True, False, "big"], ids=["lane", "group", "big"])


def same(name, data, wb, cap, lane):
    a, ao = CHK.oneshot(data, cap, wb)
    b, bo = EM.oneshot(data, cap, wb, lane)
    assert a.astuple() == b.astuple(), (name, wb, cap, a, b)
    assert ao == bo, (name, "output differs")


@PATHS
def test_golden_fixtures(lane):
    for name, data, m in cases.golden_fixtures():
        r, out = EM.oneshot(data, m["out_len"] + 64, -16, lane)
        assert (r.ret, r.msg_id, r.total_in, r.total_out) == (1, 0, m["in_len"], m["out_len"]), (name, r)
        assert hashlib.sha256(out).hexdigest() == m["out_sha256"], name
        same(name, data, -16, m["out_len"], lane)
        same(name, data, -15, m["out_len"], lane)      # Deflate64 data read as deflate: error parity


@PATHS
def test_probe_vectors(lane):
    for name, data, wb, cap, want in cases.probe_vectors():
        r, out = EM.oneshot(data, cap, wb, lane)
        assert (r.ret, r.msg_id, r.total_in, r.total_out) == want, (name, r)
        same(name, data, wb, cap, lane)


@PATHS
def test_valid_streams(lane):
    for name, data, wb, plain in cases.valid_streams():
        for cap in (len(plain) + 16, len(plain), max(len(plain) - 1, 0)):
            same(name, data, wb, cap, lane)


@PATHS
def test_damaged_streams(lane):
    for name, data, wb, cap in cases.damaged_streams(1500):
        same(name, data, wb, cap, lane)
