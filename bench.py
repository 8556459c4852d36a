#!/usr/bin/env python
"""bench.py — batched inflate throughput on B200 (BASELINE.json metric, config C2).

Workload (N=1): 65,536 x 64 KiB synthetic raw-deflate members (zlib level 6, word-soup plaintext,
SURVEY.md §8d), built from DISTINCT seeded members tiled to distinct HBM addresses.
  value     = decompressed GB/s, members resident in HBM, CUDA-event timed (weak scaling: every rank
              decodes its own 65,536 members; value is the sum over ranks / max time).
  e2e       = same metric through zsb_inflate_batch() with HOST (pinned) buffers: host->device copy of
              the compressed members and device->host copy of outputs + results inside the timed region.
  roofline  = (compressed in + decompressed out bytes per launch) / kernel time vs measured HBM peak.
  cpu_baseline / --impl reference = the reference's own C inflate (oracle/_ref, built from
              /root/reference/src/{inflate,inffast,inftrees}.c) one stream per core on the host cores.
"""
import argparse
import ctypes
import json
import os

# more hardware work queues than the default 8, so that every chunk of a pipelined host batch (the e2e leg)
# decodes on a stream of its own; must be in the environment before CUDA starts
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# stdout carries the ONE JSON line: NCCL's own log (the "NCCL version" line of NCCL_DEBUG=VERSION included) goes to stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

MEMBER_BYTES = 65536
# dram__bytes_read.sum + dram__bytes_write.sum per launch on this workload, from the committed ncu --set full
# capture (profiles/r01_ncu_summary.txt); updated whenever the kernel changes
NCU_TRAFFIC = {"k_lane": 8.83e9}


def _make_member(seed):
    import gen
    p = gen.word_soup(MEMBER_BYTES, seed)
    return gen.deflate(p, 6, -15)


def make_distinct(k, rank):
    """k distinct compressed members (seeds 1000+i, offset per rank so ranks hold different data)."""
    from concurrent.futures import ProcessPoolExecutor
    seeds = [1000 + rank * 100000 + i for i in range(k)]
    workers = max(1, min(os.cpu_count() or 1, 32))
    if workers == 1:
        return [_make_member(s) for s in seeds]
    with ProcessPoolExecutor(workers) as ex:
        return list(ex.map(_make_member, seeds, chunksize=16))


def peaks():
    try:
        j = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU while the timed region runs: NVML in-process (a sample every
    few ms, so even a 0.3 s region is covered on a busy 8-GPU box), nvidia-smi as the fallback."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu = gpu
        self.stop = False
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self.how = "nvml"
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[gpu]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else gpu
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None
            self.how = "nvidia-smi"

    def sample_nvml(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
            else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for nm, bit in (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40),
                        ("sw_power_cap", 0x4)):
            if r & bit:
                self.reasons.add(nm)

    def sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        o = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q,
                            "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        f = [x.strip() for x in o.strip().split(",")]
        self.sm.append(float(f[0]))
        self.max_mhz = float(f[1])
        for nm, v in zip(names, f[2:]):
            if v.lower().startswith("active"):
                self.reasons.add(nm)

    def run(self):
        while not self.stop:
            try:
                if self.h is not None:
                    self.sample_nvml()
                else:
                    self.sample_smi()
            except Exception:
                pass
            time.sleep(0.005 if self.h is not None else 0.05)

    def summary(self):
        s = sorted(self.sm)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "source": self.how}


def cpu_reference_run(members, steps, warmup, threads):
    """Times the reference's C inflate (oracle/_ref; else the oracle port) on host cores."""
    import orclib
    chk = orclib.ref()
    kind = "reference"
    if chk is None:
        chk, kind = orclib.orc(), "port"
    caps = [MEMBER_BYTES] * len(members)
    wb = [-15] * len(members)
    out_bytes = None
    for _ in range(warmup):
        chk.bench(members, wb, caps, threads, 1)
    best = []
    for _ in range(steps):
        secs, so, bad = chk.bench(members, wb, caps, threads, 1)
        assert bad == 0
        out_bytes = so
        best.append(secs)
    return kind, out_bytes, best


def cpu_sample(distinct, threads):
    """A bounded sample of the workload: about 2 s of CPU work per pass at ~0.12 GB/s/core."""
    per_pass = 2.0 * threads * 0.12e9
    reps = max(1, int(per_pass / (len(distinct) * MEMBER_BYTES) + 0.5))
    return distinct * reps, "%d members = %d distinct x %d (64 KiB raw-deflate level 6), one stream per thread" % (
        len(distinct) * reps, len(distinct), reps)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--members", type=int, default=65536)
    ap.add_argument("--distinct", type=int, default=1024)
    ap.add_argument("--lanes", type=int, default=0, help="0 = lane path (k_lane + k_lz); else group path k_inflate<G>")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--small-ctas", action="store_true", help="diagnostic: device-resident leg with k_lane in small CTAs")
    ap.add_argument("--chunk-sweep", default="", help="comma list of pipeline depths to time the e2e leg with (diagnostic)")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    metric = "batched inflate decompressed GB/s"
    config = {"workload": "C2: %d x 64 KiB synthetic raw-deflate (zlib level 6, word-soup) members per GPU" % args.members,
              "distinct_members": args.distinct, "window_bits": -15,
              "l2": "inputs+outputs per step (~6.4 GB) exceed the 126 MB L2; no explicit flush"}

    if args.impl == "reference":
        if rank != 0:
            return
        distinct = make_distinct(min(args.distinct, 256), 0)
        sample, desc = cpu_sample(distinct, threads)
        kind, out_bytes, secs = cpu_reference_run(sample, args.steps, args.warmup, threads)
        t = sum(secs) / len(secs)
        v = out_bytes / t / 1e9
        print(json.dumps({"impl": "reference", "metric": metric, "value": v, "unit": "GB/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                          "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": "GB/s", "cores": threads, "kind": kind, "sample": desc},
                          "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return

    import numpy as np
    import torch
    import zsblib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libzsb has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    distinct = make_distinct(args.distinct, rank)
    n = args.members
    k = len(distinct)
    lens = np.array([len(d) for d in distinct], dtype=np.int64)
    pad = (-lens) % 16
    offs = np.concatenate(([0], np.cumsum(lens + pad)))[:-1]
    blob = bytearray(int((lens + pad).sum()))
    for d, o in zip(distinct, offs):
        blob[o:o + len(d)] = d
    blob_len = len(blob)
    tiles = (n + k - 1) // k
    in_len = lens[np.arange(n) % k]
    in_off = offs[np.arange(n) % k] + (np.arange(n) // k) * blob_len
    out_off = np.arange(n, dtype=np.int64) * MEMBER_BYTES
    sum_in, sum_out = int(in_len.sum()), n * MEMBER_BYTES

    eng = zsblib.Zsb(devices=[local_rank])
    if args.lanes:
        eng.set_option("path", 2)
        eng.set_option("lanes_per_member", args.lanes)
    else:
        eng.set_option("path", 0)           # the default: lane path (k_lane + k_lz)
        if args.small_ctas:
            eng.set_option("small_ctas", 1)
    lib = eng.lib

    def descriptors(in_base, out_base):
        a = np.zeros(n, dtype=np.dtype([("inp", "<u8"), ("in_len", "<u8"), ("out", "<u8"), ("out_cap", "<u8"),
                                        ("wb", "<i4"), ("rsv", "<i4")]))
        a["inp"] = in_base + in_off
        a["in_len"] = in_len
        a["out"] = out_base + out_off
        a["out_cap"] = MEMBER_BYTES
        a["wb"] = -15
        return a

    # ---------------- device-resident leg ----------------
    t_blob = torch.frombuffer(blob, dtype=torch.uint8).to(dev)
    t_in = t_blob.repeat(tiles)
    t_out = torch.empty(n * MEMBER_BYTES, dtype=torch.uint8, device=dev)
    desc = descriptors(t_in.data_ptr(), t_out.data_ptr())
    t_desc = torch.from_numpy(desc.view(np.uint8).copy()).to(dev)
    t_res = torch.zeros(n * 32, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    def step():
        rc = lib.zsb_inflate_batch_device(eng.ctx, 0, n, t_desc.data_ptr(), desc.ctypes.data, t_res.data_ptr(), None,
                                          zsblib.ZSB_RAW_ONLY, ctypes.c_void_p(stream.cuda_stream))
        if rc:
            raise RuntimeError("zsb_inflate_batch_device rc=%d %s" % (rc, eng.error()))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    launches_per_step = lib.zsb_last_launches(eng.ctx)
    if not args.lanes:
        eng.set_option("profile", 1)        # CUDA events around every k_lane / k_lz launch, on `stream`
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    sampler.stop = True
    ms = ev0.elapsed_time(ev1)
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    lane_ms = lz_ms = None
    if not args.lanes:
        a, b = ctypes.c_double(0), ctypes.c_double(0)
        pairs = lib.zsb_profile_read(eng.ctx, 0, ctypes.byref(a), ctypes.byref(b))
        eng.set_option("profile", 0)
        if pairs > 0:
            lane_ms, lz_ms = a.value / pairs, b.value / pairs

    # results: every member must have ended cleanly with 64 KiB out; spot-check bytes with Python's zlib
    res = t_res.cpu().numpy().view(np.dtype([("ret", "<i4"), ("msg", "<i4"), ("tin", "<u8"), ("tout", "<u8"),
                                             ("chk", "<u4"), ("dt", "<i4")]))
    ok = bool((res["ret"] == 1).all() and (res["tout"] == MEMBER_BYTES).all() and (res["tin"] == in_len).all())
    import zlib
    for i in (0, 1, n // 2, n - 1):
        got = t_out[i * MEMBER_BYTES:(i + 1) * MEMBER_BYTES].cpu().numpy().tobytes()
        ok = ok and got == zlib.decompress(distinct[i % k], -15)
    if not ok:
        raise SystemExit("bench.py: GPU results differ from expectation (ret/total_out/bytes)")

    kernel_name = ("k_inflate<%d>" % eng.get_option("lanes_per_member")) if args.lanes else "k_lane"
    value = world * sum_out / (ms_per_step * 1e-3) / 1e9
    peak, peak_src = peaks()
    # the dominant kernel's own launch duration (events on the launching stream); the group path is one kernel
    kernel_ms = lane_ms if lane_ms else ms_per_step
    achieved = (sum_in + sum_out) / (kernel_ms * 1e-3) / 1e9
    out = {"metric": metric, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": NCU_TRAFFIC.get(kernel_name), "peak_source": peak_src, "kernel": kernel_name,
                        "kernel_ms": kernel_ms, "kernel_share_of_step": kernel_ms / ms_per_step,
                        "other_kernels_ms": {"k_lz": lz_ms} if lz_ms else {},
                        "step_achieved": (sum_in + sum_out) / (ms_per_step * 1e-3) / 1e9,
                        "algorithmic_bytes_per_launch": sum_in + sum_out},
           "clocks": sampler.summary(), "gpu_launches": int(launches_per_step * args.steps),
           "path": "group" if args.lanes else "lane", "lanes_per_member": int(args.lanes)}
    del t_in, t_out, t_blob
    torch.cuda.empty_cache()

    # ---------------- end-to-end leg: host buffers through the public batch call ----------------
    if not args.no_e2e:
        in_bytes = tiles * blob_len
        # pinned buffers on the GPU's own NUMA node: pages are placed where the allocating thread runs
        saved_aff, numa = None, "n/a"
        try:
            props = torch.cuda.get_device_properties(local_rank)
            bdf = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
            with open("/sys/bus/pci/devices/%s/local_cpulist" % bdf) as f:
                cpus = set()
                for part in f.read().strip().split(","):
                    a, _, b = part.partition("-")
                    cpus.update(range(int(a), int(b or a) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                saved_aff = os.sched_getaffinity(0)
                os.sched_setaffinity(0, cpus)
                numa = "gpu-local cpus %d" % len(cpus)
        except Exception as e:          # no sysfs / no such attribute: allocate wherever we run
            numa = "unpinned (%s)" % type(e).__name__
        h_in = lib.zsb_host_alloc(in_bytes)
        h_out = lib.zsb_host_alloc(n * MEMBER_BYTES)
        if h_out:
            ctypes.memset(h_out, 0, n * MEMBER_BYTES)
        if saved_aff is not None:
            os.sched_setaffinity(0, saved_aff)
        if not h_in or not h_out:
            raise SystemExit("bench.py: pinned host allocation failed")
        src = (ctypes.c_char * blob_len).from_buffer(blob)
        for t_ in range(tiles):
            ctypes.memmove(h_in + t_ * blob_len, src, blob_len)
        hdesc = descriptors(h_in, h_out)
        hres = np.zeros(n * 32, dtype=np.uint8)
        e2e_steps = max(2, min(args.steps, 3))

        def e2e_step():
            rc = lib.zsb_inflate_batch(eng.ctx, n, hdesc.ctypes.data, hres.ctypes.data)
            if rc:
                raise RuntimeError("zsb_inflate_batch rc=%d %s" % (rc, eng.error()))

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        t1 = time.perf_counter()
        es = (t1 - t0) / e2e_steps
        if dist is not None:
            t = torch.tensor([es], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            es = float(t.item())
        if args.chunk_sweep:
            sweep = {}
            for kk in [int(x) for x in args.chunk_sweep.split(",")]:
                eng.set_option("chunks", kk)
                e2e_step()
                t0 = time.perf_counter()
                for _ in range(2):
                    e2e_step()
                sweep[kk] = (time.perf_counter() - t0) / 2 * 1e3
            eng.set_option("chunks", 0)
            out["e2e_chunk_sweep_ms"] = sweep
        r2 = hres.view(res.dtype)
        got = ctypes.string_at(h_out + (n - 1) * MEMBER_BYTES, MEMBER_BYTES)
        if not ((r2["ret"] == 1).all() and (r2["tout"] == MEMBER_BYTES).all() and
                got == zlib.decompress(distinct[(n - 1) % k], -15)):
            raise SystemExit("bench.py: e2e results differ from expectation")
        out["e2e"] = {"value": world * sum_out / es / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(sum_in + n * 40),
                      "d2h_bytes_per_step": int(sum_out + n * 32), "ms_per_step": es * 1e3, "steps": e2e_steps,
                      "host_buffers": "pinned (zsb_host_alloc), " + numa,
                      "timer": "host clock around the synchronous zsb_inflate_batch() call, max over ranks",
                      "gpu_launches": int(lib.zsb_last_launches(eng.ctx))}
        lib.zsb_host_free(h_in)
        lib.zsb_host_free(h_out)

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    if rank == 0 and world == 1 and not args.no_cpu:
        sample, sdesc = cpu_sample(distinct[:256], threads)
        kind, out_bytes, secs = cpu_reference_run(sample, 3, 1, threads)
        out["cpu_baseline"] = {"value": out_bytes / min(secs) / 1e9, "unit": "GB/s", "cores": threads, "kind": kind,
                               "sample": sdesc}
    if rank == 0:
        print(json.dumps(out))
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
