"""TEST / BENCH TOOL: one rank's share of a synthetic batch, resident in HBM, and its verification.

A workload (tests/workloads.py) is K distinct seeded members and a member count n = tiles * K; global member
g is distinct member g % K (tile g // K).  `DeviceBatch` takes the list of global members ONE rank decodes
(all of them at N = 1, its zsb_plan_shards() share otherwise), gives every member its own compressed bytes in
HBM (the k-th member of a rank that is distinct member d reads copy k of the distinct blob, so no two members
share input cache lines) and its own output range, laid out copy by copy ("tile" by tile, the order of the descriptors too, so that neighbouring
members have adjacent buffers as in a real archive); the outputs of one distinct member are then a strided column
block of the (copies x tile bytes) output matrix.  verify() then checks EVERY member: the whole result record
against the oracle's record for its distinct member, and every output byte [0, total_out) against the
oracle's bytes (one broadcast comparison per distinct member on the GPU).  Nothing here is used by the product.
"""
import ctypes
import hashlib
import os
import pickle
from concurrent.futures import ThreadPoolExecutor

import numpy as np

import orclib
import zsblib

RES_DT = np.dtype([("ret", "<i4"), ("msg", "<i4"), ("tin", "<u8"), ("tout", "<u8"), ("chk", "<u4"), ("dt", "<i4")])
MEM_DT = np.dtype([("inp", "<u8"), ("in_len", "<u8"), ("out", "<u8"), ("out_cap", "<u8"), ("wb", "<i4"), ("flags", "<i4")])


def expected(distinct, threads=None):
    """Oracle records and outputs of the distinct members (the reference build where present, else the port)."""
    chk = orclib.ref() or orclib.orc()

    def one(m):
        comp, wb, plain = m
        r, out = chk.oneshot(comp, len(plain), wb)
        return r.astuple(), out
    with ThreadPoolExecutor(threads or min(32, os.cpu_count() or 1)) as ex:    # the C library runs without the GIL
        got = list(ex.map(one, distinct))
    rec = np.zeros(len(distinct), dtype=RES_DT)
    for i, (t, _) in enumerate(got):
        rec[i] = t
    return rec, [o for _, o in got]


def shared_workload(name, build, rank, world, barrier):
    """Build a workload once per box: rank 0 builds (all host cores) and leaves a pickle in /dev/shm (or /tmp),
    the other ranks load it after the barrier."""
    if world == 1:
        return build()
    d = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
    path = os.path.join(d, "zsb_workload_%s_%d.pkl" % (hashlib.sha1(name.encode()).hexdigest()[:12], os.getppid()))
    if rank == 0:
        w = build()
        with open(path + ".tmp", "wb") as f:
            pickle.dump(w, f, protocol=4)
        os.replace(path + ".tmp", path)
    barrier()
    if rank != 0:
        with open(path, "rb") as f:
            w = pickle.load(f)
    barrier()
    if rank == 0:
        try:
            os.remove(path)
        except OSError:
            pass
    return w


class DeviceBatch:
    def __init__(self, w, dev, members=None, exp=None):
        """w: workload; dev: torch device; members: global member indices of this rank (None = all);
        exp: expected(w["distinct"]) if the caller has it already."""
        import torch
        self.torch = torch
        self.dev = dev
        distinct, tiles = w["distinct"], w["tiles"]
        K = len(distinct)
        self.K = K
        self.desc_text = w["desc"]
        if members is None:
            members = np.arange(K * tiles, dtype=np.int64)
        members = np.asarray(members, dtype=np.int64)
        self.rec, self.outs = exp if exp is not None else expected(distinct)
        lens = np.array([len(m[0]) for m in distinct], dtype=np.int64)
        caps = np.array([len(m[2]) for m in distinct], dtype=np.int64)
        wbs = np.array([m[1] for m in distinct], dtype=np.int32)
        # copy index = occurrence number of the distinct member on this rank; members are ordered copy-major
        # (the order of a tiled batch), so that consecutive members have adjacent input and output buffers
        d0 = members % K
        o0 = np.argsort(d0, kind="stable")
        count = np.bincount(d0, minlength=K).astype(np.int64)
        first = np.concatenate(([0], np.cumsum(count)))[:-1]
        copy0 = np.empty(len(members), dtype=np.int64)
        copy0[o0] = np.arange(len(members), dtype=np.int64) - first[d0[o0]]
        order = np.lexsort((d0, copy0))
        self.global_ids = members[order]
        d, copy = d0[order], copy0[order]
        n = len(d)
        self.n = n
        self.count, self.d = count, d
        ncopies = int(count.max()) if n else 0
        ioff = np.concatenate(([0], np.cumsum((lens + 15) & ~15)))
        blob = np.zeros(int(ioff[-1]), dtype=np.uint8)
        for i, m in enumerate(distinct):
            blob[ioff[i]:ioff[i] + lens[i]] = np.frombuffer(m[0], dtype=np.uint8)
        self.blob, self.blob_len, self.ncopies, self.ioff = blob, int(ioff[-1]), ncopies, ioff
        ooff1 = np.concatenate(([0], np.cumsum(caps)))                   # output offsets inside one copy ("tile")
        self.ooff1, self.tile_bytes = ooff1, int(ooff1[-1])
        self.caps, self.lens, self.wbs = caps, lens, wbs
        self.in_off = ioff[d] + copy * self.blob_len
        self.out_off = ooff1[d] + copy * self.tile_bytes
        self.in_len = lens[d]
        self.out_cap = caps[d]
        self.wb = wbs[d]
        self.sum_in = int(self.in_len.sum())
        self.sum_out = int(self.rec["tout"][d].sum())
        self.out_bytes_total = ncopies * self.tile_bytes
        self.raw_only = bool((wbs < 0).all())
        self.t_in = self.t_out = self.t_desc = self.t_res = None

    # ---- device-resident form ----
    def to_device(self):
        torch = self.torch
        t_blob = torch.from_numpy(self.blob).to(self.dev)
        self.t_in = t_blob.repeat(max(self.ncopies, 1))
        del t_blob
        self.t_out = torch.zeros(max(self.out_bytes_total, 1), dtype=torch.uint8, device=self.dev)
        self.desc = self.descriptors(self.t_in.data_ptr(), self.t_out.data_ptr())
        self.t_desc = torch.from_numpy(self.desc.view(np.uint8).copy()).to(self.dev)
        self.t_res = torch.zeros(max(self.n, 1) * 32, dtype=torch.uint8, device=self.dev)
        return self

    def descriptors(self, in_base, out_base):
        a = np.zeros(self.n, dtype=MEM_DT)
        a["inp"] = in_base + self.in_off
        a["in_len"] = self.in_len
        a["out"] = out_base + self.out_off
        a["out_cap"] = self.out_cap
        a["wb"] = self.wb
        return a

    def step(self, eng, stream, dev_index=0):
        flags = zsblib.ZSB_RAW_ONLY if self.raw_only else zsblib.ZSB_HAS_WRAPPED
        rc = eng.lib.zsb_inflate_batch_device(eng.ctx, dev_index, self.n, self.t_desc.data_ptr(), self.desc.ctypes.data,
                                              self.t_res.data_ptr(), None, flags, ctypes.c_void_p(stream.cuda_stream))
        if rc:
            raise RuntimeError("zsb_inflate_batch_device rc=%d %s" % (rc, eng.error()))

    def clear_outputs(self):
        self.t_out.zero_()
        self.t_res.zero_()

    def verify_records(self, res):
        """res: structured array of this rank's n result records (in this batch's member order)."""
        bad = []
        for f in RES_DT.names:
            ne = np.nonzero(res[f] != self.rec[f][self.d])[0]
            if len(ne):
                j = int(ne[0])
                bad.append("%s: member %d (distinct %d) got %r want %r" % (f, int(self.global_ids[j]), int(self.d[j]),
                                                                            res[f][j], self.rec[f][self.d[j]]))
        return bad

    def verify(self):
        """Every record and every output byte of this rank's members; returns a list of mismatch descriptions."""
        torch = self.torch
        res = self.t_res.cpu().numpy()[:self.n * 32].view(RES_DT)
        bad = self.verify_records(res)
        exp_cat = np.frombuffer(b"".join(self.outs), dtype=np.uint8)
        t_exp = torch.from_numpy(exp_cat.copy()).to(self.dev) if len(exp_cat) else None
        eoff = np.concatenate(([0], np.cumsum([len(o) for o in self.outs])))
        tiles = self.t_out[:self.ncopies * self.tile_bytes].view(max(self.ncopies, 1), -1) if self.n else None
        for di in range(self.K):
            c, cap, e = int(self.count[di]), int(self.caps[di]), len(self.outs[di])
            if c == 0 or e == 0:
                continue
            rows = tiles[:c, int(self.ooff1[di]):int(self.ooff1[di]) + e]
            want = t_exp[int(eoff[di]):int(eoff[di]) + e]
            if not bool((rows == want.unsqueeze(0)).all()):
                neq = (rows != want.unsqueeze(0)).nonzero()[0]
                bad.append("bytes: distinct %d copy %d offset %d" % (di, int(neq[0]), int(neq[1])))
                break
        return bad

    # ---- host form (the e2e leg): pinned buffers with the same layout ----
    def to_host(self, lib):
        in_bytes = max(self.ncopies, 1) * self.blob_len
        self.h_in = lib.zsb_host_alloc(max(in_bytes, 1))
        self.h_out = lib.zsb_host_alloc(max(self.out_bytes_total, 1))
        if not self.h_in or not self.h_out:
            raise MemoryError("pinned host allocation failed")
        ctypes.memset(self.h_out, 0, max(self.out_bytes_total, 1))
        src = self.blob.ctypes.data
        for c in range(max(self.ncopies, 1)):
            ctypes.memmove(self.h_in + c * self.blob_len, src, self.blob_len)
        self.hdesc = self.descriptors(self.h_in, self.h_out)
        self.hres = np.zeros(max(self.n, 1) * 32, dtype=np.uint8)
        return self

    def host_step(self, eng):
        rc = eng.lib.zsb_inflate_batch(eng.ctx, self.n, self.hdesc.ctypes.data, self.hres.ctypes.data)
        if rc:
            raise RuntimeError("zsb_inflate_batch rc=%d %s" % (rc, eng.error()))

    def verify_host(self):
        res = self.hres[:self.n * 32].view(RES_DT)
        bad = self.verify_records(res)
        htiles = np.ctypeslib.as_array((ctypes.c_uint8 * max(self.out_bytes_total, 1)).from_address(self.h_out))
        htiles = htiles[:self.ncopies * self.tile_bytes].reshape(max(self.ncopies, 1), -1)
        for di in range(self.K):
            c, cap, e = int(self.count[di]), int(self.caps[di]), len(self.outs[di])
            if c == 0 or e == 0:
                continue
            rows = htiles[:c, int(self.ooff1[di]):int(self.ooff1[di]) + e]
            want = np.frombuffer(self.outs[di], dtype=np.uint8)
            if not (rows == want[None, :]).all():
                bad.append("host bytes: distinct %d" % di)
                break
        return bad

    def free_host(self, lib):
        lib.zsb_host_free(self.h_in)
        lib.zsb_host_free(self.h_out)
        self.h_in = self.h_out = None
