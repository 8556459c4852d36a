#!/usr/bin/env python
"""bench.py — batched inflate throughput on B200 (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W] [--config C2|C3|C4|C5] [--impl reference]

Default workload (the configuration the metric is quoted on): C2 = 65,536 x 64 KiB synthetic raw-deflate
members (zlib level 6, word-soup plaintext) PER GPU, built from distinct seeded members placed
at distinct HBM addresses.  --config C3 / C4 / C5 runs BASELINE.json's other configurations the same way.
  value     = decompressed GB/s, members resident in HBM, CUDA-event timed on the launching stream (weak
              scaling: every rank decodes its own batch; value = sum over ranks / max time).
  e2e       = same metric through zsb_inflate_batch() with HOST (pinned) buffers: host->device copy of the
              compressed members and device->host copy of outputs + results inside the timed region.
  roofline  = (compressed in + decompressed out bytes of a step) / the STEP's device time (all decode kernels:
              k_lane + k_lz [+ k_big, k_check]) against the measured HBM peak; traffic = the summed DRAM bytes
              of those kernels from the committed ncu capture (profiles/r02_traffic.json).
  strong    = (C2 runs, unless --no-strong) ONE fixed C5 batch of 1,048,576 members split over the N ranks by
              the product's scheduler zsb_plan_shards(), device-resident, timed barrier to barrier; every
              rank's members are verified record by record and byte by byte.
  cpu_baseline / --impl reference = the reference's own C inflate (oracle/_ref, built from
              /root/reference/src/{inflate,inffast,inftrees}.c) one stream per core on the host cores.
Every member decoded in the timed legs is verified against the oracle (records + bytes) after the timing.
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

METRIC = "batched inflate decompressed GB/s"


def peaks():
    try:
        j = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(config):
    """Summed dram__bytes_read + dram__bytes_write of the step's decode kernels, from the committed ncu
    --set full capture of this command (profiles/r02_traffic.json, written by scripts/ncu_traffic.py)."""
    try:
        j = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        return j.get(config)
    except Exception:
        return None


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


# ---------------------------------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------------------------------
def build_workload(config, args, rank):
    import workloads
    if config == "C2":
        tiles = max(1, args.members // args.distinct)
        return workloads.c2(distinct=args.distinct, tiles=tiles, seed_base=1000 + rank * 100000)
    if config == "C3":
        return workloads.c3(distinct=args.distinct_c3, tiles=max(1, 4096 // args.distinct_c3))
    if config == "C4":
        return workloads.c4(distinct=512, tiles=32)
    return workloads.c5(distinct=args.distinct_c5, tiles=max(1, (1 << 20) // args.distinct_c5))


def config_desc(config, w, per):
    return {"workload": "%s: %s%s" % (config, w["desc"], per), "distinct_members": len(w["distinct"]),
            "l2": "inputs + outputs of a step (GBs) exceed the 126 MB L2; no explicit flush",
            "verified": "every member of every timed leg: result record and output bytes against the oracle"}


def cpu_reference_run(sample, steps, warmup, threads):
    """Times the reference's C inflate (oracle/_ref; else the oracle port) on host cores.
    sample: list of (compressed, wbits, out_cap)."""
    import orclib
    chk = orclib.ref()
    kind = "reference"
    if chk is None:
        chk, kind = orclib.orc(), "port"
    members = [m[0] for m in sample]
    wb = [m[1] for m in sample]
    caps = [m[2] for m in sample]
    out_bytes = None
    for _ in range(warmup):
        chk.bench(members, wb, caps, threads, 1)
    secs = []
    for _ in range(steps):
        s, so, _bad = chk.bench(members, wb, caps, threads, 1)
        out_bytes = so
        secs.append(s)
    return kind, out_bytes, secs


def cpu_sample(w, threads, seconds=2.0):
    """A bounded sample of the workload's distinct members: about `seconds` of CPU work per pass at
    ~0.12 GB/s/core, whole distinct members in order (repeated when the distinct set is smaller)."""
    budget = seconds * threads * 0.12e9
    distinct = w["distinct"]
    out, total, i = [], 0, 0
    while total < budget and i < 4 * len(distinct) + 64:
        comp, wb, plain = distinct[i % len(distinct)]
        out.append((comp, wb, len(plain)))
        total += len(plain)
        i += 1
    return out, "%d members (%.2f GB out) drawn in order from the %d distinct members of %s, one stream per thread" % (
        len(out), total / 1e9, len(distinct), w["name"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=["C2", "C3", "C4", "C5"])
    ap.add_argument("--members", type=int, default=65536)
    ap.add_argument("--distinct", type=int, default=1024)
    ap.add_argument("--distinct-c3", type=int, default=64)
    ap.add_argument("--distinct-c5", type=int, default=4096)
    ap.add_argument("--lanes", type=int, default=0, help="0 = lane path (k_lane + k_lz); else group path k_inflate<G>")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--option", action="append", default=[], help="name=value for zsb_set_option (diagnostic)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    config = args.config
    if config != "C2":
        args.steps = min(args.steps, 5)          # seconds per step: keep the default run within minutes

    if args.impl == "reference":
        if rank != 0:
            return
        if config == "C2":
            args.distinct = min(args.distinct, 256)
        elif config == "C3":
            args.distinct_c3 = min(args.distinct_c3, 16)
        elif config == "C5":
            args.distinct_c5 = min(args.distinct_c5, 2048)
        w = build_workload(config, args, 0)
        sample, desc = cpu_sample(w, threads)
        kind, out_bytes, secs = cpu_reference_run(sample, args.steps, args.warmup, threads)
        t = sum(secs) / len(secs)
        v = out_bytes / t / 1e9
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "GB/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                          "data": "synthetic", "config": config_desc(config, w, " per GPU (bounded sample on the host cores)"),
                          "cpu_baseline": {"value": v, "unit": "GB/s", "cores": threads, "kind": kind, "sample": desc},
                          "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return

    import numpy as np
    import torch
    import batchlib
    import zsblib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libzsb has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    eng = zsblib.Zsb(devices=[local_rank])
    if args.lanes:
        eng.set_option("path", 2)
        eng.set_option("lanes_per_member", args.lanes)
    for kv in args.option:
        k, _, v = kv.partition("=")
        eng.set_option(k, int(v))
    lib = eng.lib
    stream = torch.cuda.current_stream()

    # ---------------- device-resident leg: every rank its own batch (weak scaling) ----------------
    if config == "C2":
        w = build_workload(config, args, rank)                 # per-rank seeds: the ranks hold different data
    else:
        w = batchlib.shared_workload(config, lambda: build_workload(config, args, 0), rank, world, barrier)
    exp = batchlib.expected(w["distinct"])
    batch = batchlib.DeviceBatch(w, dev, exp=exp).to_device()
    n = batch.n
    for _ in range(args.warmup):
        batch.step(eng, stream)
    barrier()
    launches_per_step = lib.zsb_last_launches(eng.ctx)
    batch.clear_outputs()
    if not args.lanes:
        eng.set_option("profile", 1)        # CUDA events around every k_lane / k_lz launch, on `stream`
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        batch.step(eng, stream)
    ev1.record()
    barrier()
    sampler.stop = True
    ms_per_step = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    lane_ms = lz_ms = check_ms = seg_ms = None
    if not args.lanes:
        a, b = ctypes.c_double(0), ctypes.c_double(0)
        pairs = lib.zsb_profile_read(eng.ctx, 0, ctypes.byref(a), ctypes.byref(b))
        eng.set_option("profile", 0)
        if pairs > 0:
            lane_ms, lz_ms = a.value / args.steps, b.value / args.steps
        c = ctypes.c_double(0)
        if lib.zsb_profile_read_check(eng.ctx, 0, ctypes.byref(c)) > 0:
            check_ms = c.value / args.steps
        c = ctypes.c_double(0)
        if lib.zsb_profile_read_seg(eng.ctx, 0, ctypes.byref(c)) > 0:
            seg_ms = c.value / args.steps
    bad = batch.verify()
    if bad:
        raise SystemExit("bench.py: GPU results differ from the oracle: %s" % bad[:3])
    sum_in, sum_out = batch.sum_in, batch.sum_out
    total_out = sum_over_ranks(float(sum_out))
    value = total_out / (ms_per_step * 1e-3) / 1e9
    peak, peak_src = peaks()
    # the roofline is the whole decode step's: k_lane writes no output byte, k_lz (and k_big / k_check) finish the job
    kernels = ("k_inflate<%d>" % eng.get_option("lanes_per_member")) if args.lanes else \
        ("k_lane+k_lz" if batch.raw_only and config in ("C2", "C5") else
         "k_seg_*+k_lane_seg+k_lane+k_lz+k_big" + ("" if batch.raw_only else "+k_check"))
    achieved = (sum_in + sum_out) / (ms_per_step * 1e-3) / 1e9
    out = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config_desc(config, w, " per GPU"),
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": ncu_traffic(config), "peak_source": peak_src, "kernel": kernels,
                        "kernel_ms": ms_per_step,
                        "kernel_ms_split": {"k_lane(+k_big, concurrent streams)": lane_ms, "k_lz": lz_ms, "k_check": check_ms,
                                            "segmented members (k_seg_* + k_lane_seg, own stream beside k_lane)": seg_ms},
                        "k_check_gbs": (sum_out / (check_ms * 1e-3) / 1e9) if check_ms else None,
                        "algorithmic_bytes_per_launch": sum_in + sum_out},
           "clocks": sampler.summary(), "gpu_launches": int(launches_per_step * args.steps),
           "path": "group" if args.lanes else "lane", "lanes_per_member": int(args.lanes),
           "members_verified": int(n)}
    h2d, d2h = int(sum_in + n * 40), int(batch.out_bytes_total + n * 32)
    del batch
    torch.cuda.empty_cache()

    # ---------------- end-to-end leg: host buffers through the public batch call ----------------
    if not args.no_e2e:
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
        # a host batch of at most ~8 GB of output: the whole of C2, a tile-wise cut of the larger configurations
        hw = dict(w)
        per_tile = sum(len(m[2]) for m in w["distinct"])
        hw["tiles"] = max(1, min(w["tiles"], int(8e9 // max(per_tile, 1))))
        hb = batchlib.DeviceBatch(hw, dev, exp=exp).to_host(lib)
        if saved_aff is not None:
            os.sched_setaffinity(0, saved_aff)
        e2e_steps = max(2, min(args.steps, 3))
        hb.host_step(eng)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            hb.host_step(eng)
        t1 = time.perf_counter()
        es = max_over_ranks((t1 - t0) / e2e_steps)
        badh = hb.verify_host()
        if badh:
            raise SystemExit("bench.py: e2e results differ from the oracle: %s" % badh[:3])
        e2e_out = sum_over_ranks(float(hb.sum_out))
        out["e2e"] = {"value": e2e_out / es / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(hb.sum_in + hb.n * 40),
                      "d2h_bytes_per_step": int(hb.out_bytes_total + hb.n * 32), "ms_per_step": es * 1e3, "steps": e2e_steps,
                      "members": int(hb.n), "host_buffers": "pinned (zsb_host_alloc), " + numa,
                      "timer": "host clock around the synchronous zsb_inflate_batch() call, max over ranks",
                      "gpu_launches": int(lib.zsb_last_launches(eng.ctx))}
        hb.free_host(lib)
        del hb

    # ---------------- strong scaling through the product's scheduler: one C5 batch over N ranks ----------------
    if config == "C2" and not args.no_strong and not args.lanes:
        import workloads
        sw = batchlib.shared_workload("C5strong", lambda: workloads.c5(distinct=args.distinct_c5,
                                                                       tiles=max(1, (1 << 20) // args.distinct_c5)),
                                      rank, world, barrier)
        sexp = batchlib.expected(sw["distinct"])
        K, nn = len(sw["distinct"]), len(sw["distinct"]) * sw["tiles"]
        g = np.arange(nn, dtype=np.int64)
        cost = np.array([len(m[0]) + len(m[2]) for m in sw["distinct"]], dtype=np.uint64)[g % K]
        t_plan = time.perf_counter()
        shard_of = np.zeros(nn, dtype=np.int32)
        load = np.zeros(world, dtype=np.uint64)
        rc = lib.zsb_plan_shards(nn, cost.ctypes.data, world, shard_of.ctypes.data, load.ctypes.data)
        if rc:
            raise SystemExit("zsb_plan_shards rc=%d" % rc)
        plan_ms = (time.perf_counter() - t_plan) * 1e3
        mine = g[shard_of == rank]
        sb = batchlib.DeviceBatch(sw, dev, members=mine, exp=sexp).to_device()
        for _ in range(2):
            sb.step(eng, stream)
        barrier()
        sb.clear_outputs()
        ssteps = 3
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(ssteps):
            sb.step(eng, stream)
        e1.record()
        barrier()
        my_ms = e0.elapsed_time(e1) / ssteps
        sms = max_over_ranks(my_ms)
        sbad = sb.verify()
        if sbad:
            raise SystemExit("bench.py: strong-scaling results differ from the oracle: %s" % sbad[:3])
        rank_ms = [0.0] * world
        if dist is not None:
            t = torch.zeros(world, device=dev, dtype=torch.float64)
            t[rank] = my_ms
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            rank_ms = [float(x) for x in t.tolist()]
        else:
            rank_ms = [my_ms]
        tot_out = sum_over_ranks(float(sb.sum_out))
        tot_in = sum_over_ranks(float(sb.sum_in))
        out["strong"] = {"workload": "C5: " + sw["desc"] + ", ONE batch split over the ranks by zsb_plan_shards()",
                         "members": int(nn), "distinct_members": K, "ms": sms, "value": tot_out / (sms * 1e-3) / 1e9,
                         "unit": "GB/s", "in_plus_out_gbs": (tot_in + tot_out) / (sms * 1e-3) / 1e9,
                         "frac_of_hbm_peak_per_gpu": (tot_in + tot_out) / (sms * 1e-3) / 1e9 / world / peak,
                         "rank_load_bytes": [int(x) for x in load.tolist()], "rank_ms": rank_ms,
                         "plan_ms_host": plan_ms, "scaling": "strong", "members_verified": int(nn),
                         "note": "efficiency = t_1 / (N * t_N) across the driver's N = 1, 2, 4, 8 runs of this line"}
        del sb
        torch.cuda.empty_cache()
        # ---- one context over all GPUs of the box (rank 0; the other ranks wait): ONE zsb_inflate_batch() call with
        # host buffers, members sharded over the devices by the library itself, results back in the caller's order
        if world > 1:
            barrier()
            if rank == 0:
                allctx = zsblib.Zsb(devices=list(range(world)))
                sub = np.arange(min(nn, 65536), dtype=np.int64)      # whole tiles: neighbouring members have adjacent buffers
                ob = batchlib.DeviceBatch(sw, dev, members=sub, exp=sexp).to_host(lib)
                ob.host_step(allctx)
                t0 = time.perf_counter()
                ob.host_step(allctx)
                oms = (time.perf_counter() - t0) * 1e3
                obad = ob.verify_host()
                if obad:
                    raise SystemExit("bench.py: one-context results differ from the oracle: %s" % obad[:3])
                out["one_context"] = {"devices": int(lib.zsb_num_devices(allctx.ctx)), "members": int(ob.n), "ms": oms,
                                      "value": ob.sum_out / (oms * 1e-3) / 1e9, "unit": "GB/s",
                                      "what": "zsb_create(all GPUs) + one zsb_inflate_batch() over host buffers while the "
                                              "other ranks idle; every member verified",
                                      "members_verified": int(ob.n)}
                ob.free_host(lib)
                allctx.close()
            barrier()

    # ---------------- CPU baseline (rank 0, N=1 only) ----------------
    if rank == 0 and world == 1 and not args.no_cpu:
        sample, sdesc = cpu_sample(w, threads)
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
