"""GPU tier, collected last: zsb_deflate_batch_static() — the entry point that allocates nothing (no counterpart in
the reference).  Its scratch is one file-scope device array, so the
test checks (1) that the compressed bytes and result records are IDENTICAL to zsb_deflate_batch()'s on the same
device-resident members — on two calls in a row, with different members, since the array keeps the first call's
contents — (2) that the streams decode to their plaintexts on the oracle, (3) that the device's free memory does
not change across the calls, and (4) that batches beyond the stated limits are refused with a message."""
import ctypes

import pytest

import gen
import orclib
import zsblib

pytestmark = pytest.mark.gpu
CHK = orclib.ref() or orclib.orc()


def device_members(torch, lib, plains, wbits, levels):
    dev = torch.device("cuda", 0)
    n = len(plains)
    caps = [lib.zsb_deflate_bound(len(p), w) for p, w in zip(plains, wbits)]
    t_in = [torch.frombuffer(bytearray(p or b"\0"), dtype=torch.uint8).to(dev) for p in plains]
    t_out = [torch.zeros(c, dtype=torch.uint8, device=dev) for c in caps]
    arr = (zsblib.DeflateMember * n)()
    for i in range(n):
        arr[i].inp, arr[i].in_len = t_in[i].data_ptr(), len(plains[i])
        arr[i].out, arr[i].out_cap = t_out[i].data_ptr(), caps[i]
        arr[i].window_bits, arr[i].level = wbits[i], levels[i]
    return arr, t_in, t_out


def record(r):
    return (r.ret, r.blocks, r.total_in, r.total_out, r.check)


def test_static_scratch_matches_the_allocating_entry_point():
    import torch
    eng = zsblib.Zsb(devices=[0])
    lib = eng.lib
    mm, mb = ctypes.c_size_t(), ctypes.c_size_t()
    nbytes = lib.zsb_deflate_static_limits(ctypes.byref(mm), ctypes.byref(mb))
    assert mm.value == 64 and mb.value == 64 and 0 < nbytes <= 16 << 20
    batches = [
        [gen.word_soup(65536, 11), gen.word_soup(200000, 12), b"", b"abc", bytes(70000), gen.mixed_plain(150000, 13),
         gen.word_soup(65537, 14), gen.word_soup(1 << 20, 15)],
        [gen.word_soup(3000, 21), gen.mixed_plain(300000, 22), gen.word_soup(65535, 23), bytes(range(256)) * 700, b"x"],
    ]
    stream = torch.cuda.Stream(torch.device("cuda", 0))
    for k, plains in enumerate(batches):                      # two calls in a row: the second meets the first one's leftovers
        n = len(plains)
        wbits = [(-15, 15, 31)[(i + k) % 3] for i in range(n)]
        levels = [(6, 1, 9, 0)[(i + k) % 4] for i in range(n)]
        arr_a, in_a, out_a = device_members(torch, lib, plains, wbits, levels)
        arr_b, in_b, out_b = device_members(torch, lib, plains, wbits, levels)
        res_a, res_b = (zsblib.DeflateResult * n)(), (zsblib.DeflateResult * n)()
        assert lib.zsb_deflate_batch(eng.ctx, n, arr_a, res_a) == 0, eng.error()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(0)[0]
        err = ctypes.create_string_buffer(256)
        rc = lib.zsb_deflate_batch_static(0, n, arr_b, res_b, ctypes.c_void_p(stream.cuda_stream), err, 256)
        assert rc == 0, err.value
        assert torch.cuda.mem_get_info(0)[0] == free0, "the call changed the device's free memory"
        for i in range(n):
            assert record(res_a[i]) == record(res_b[i]), (k, i, record(res_a[i]), record(res_b[i]))
            assert res_b[i].ret == 1
            a = out_a[i][:res_a[i].total_out].cpu().numpy().tobytes()
            b = out_b[i][:res_b[i].total_out].cpu().numpy().tobytes()
            assert a == b, (k, i)                             # byte-identical (integer work: no tolerance)
            rr, plain = CHK.oneshot(b, len(plains[i]) + 8, wbits[i])
            assert rr.ret == 1 and plain == plains[i], (k, i, rr.ret)
    # limits: 65 members; one member of 65 blocks; bad parameters answer per member like the other entry point
    n = 65
    arr, t_in, t_out = device_members(torch, lib, [b"abc"] * n, [-15] * n, [6] * n)
    res = (zsblib.DeflateResult * n)()
    err = ctypes.create_string_buffer(256)
    assert lib.zsb_deflate_batch_static(0, n, arr, res, None, err, 256) == -104 and b"at most 64" in err.value
    arr, t_in, t_out = device_members(torch, lib, [bytes(64 * 65536 + 1)], [15], [6])
    assert lib.zsb_deflate_batch_static(0, 1, arr, res, None, err, 256) == -104 and b"blocks" in err.value
    arr, t_in, t_out = device_members(torch, lib, [bytes(64 * 65536), b"abc"], [15, -15], [6, 6])
    assert lib.zsb_deflate_batch_static(0, 1, arr, res, None, err, 256) == 0, err.value       # exactly 64 blocks fit
    assert res[0].ret == 1 and res[0].blocks == 64
    assert lib.zsb_deflate_batch_static(0, 2, arr, res, None, err, 256) == -104                # 64 + 1 do not
    arr, t_in, t_out = device_members(torch, lib, [b"abc", b"abcd"], [14, -15], [6, 6])
    assert lib.zsb_deflate_batch_static(0, 2, arr, res, None, err, 256) == 0, err.value
    assert res[0].ret == -2 and res[1].ret == 1
    host = ctypes.create_string_buffer(b"abc", 3)
    arr[1].inp = ctypes.addressof(host)
    assert lib.zsb_deflate_batch_static(0, 2, arr, res, None, err, 256) == -100 and b"device pointers" in err.value
    eng.close()
