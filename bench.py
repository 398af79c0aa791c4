#!/usr/bin/env python
"""bench.py -- JPEG decode throughput of the B200-native path (BASELINE.json metric).

A "step" = one pass of the hot path over one batch: `--images` 1920x1080 baseline 4:2:0 q90 JPEGs
with restart interval 16 (BASELINE.json configs[1]) per GPU.  Independent image batches are sharded
across ranks with no collective (weak scaling: per-GPU work fixed).

  value   decoded megapixels/s, whole job, bitstreams + descriptors already resident in HBM
          (entropy decode + fused dequant/IDCT/upsample/colour; CUDA events on the library's stream)
  e2e     same metric through stbi_jpeg_decode_batch with pinned HOST buffers: marker parse,
          H2D, kernels and D2H of every pixel inside the timed region
  roofline      the fused IDCT+colour kernel: algorithmic bytes / its event-timed duration
  cpu_baseline  the reference's own jpeg.c (oracle/_ref, SSE2 build) on the host cores, bounded sample

`--impl reference` times the reference CPU decode instead (all host threads, bounded sample).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "jpeg_decode_megapixels_per_s"
UNIT = "MP/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--images", type=int, default=1024, help="images per GPU per step")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--distinct", type=int, default=64, help="distinct seeded streams, cycled to fill the batch")
    ap.add_argument("--restart", type=int, default=16)
    ap.add_argument("--subsampling", default="420")
    ap.add_argument("--progressive", type=int, default=0)
    ap.add_argument("--gray", type=int, default=0)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--e2e-images", type=int, default=0, help="images per e2e step (0 = same as --images)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


def make_streams(args, rank):
    from tools import testenc as te

    kw = dict(quality=90, subsampling=args.subsampling, restart_interval=args.restart, progressive=args.progressive)
    if args.gray:
        kw["colorspace"] = te.CS_GRAY
    d = max(1, min(args.distinct, args.images))
    return te.encode_many(d, args.width, args.height, kind=-1, base_seed=1000 + 7919 * rank, **kw)


def workload_name(args):
    fmt = "gray" if args.gray else args.subsampling
    return (f"{args.images}x {args.width}x{args.height} {'progressive' if args.progressive else 'baseline'} "
            f"{fmt} q90 restart{args.restart} per GPU")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks line (B200_PROFILING.md) sampled during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.proc = None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_run(streams, nthreads, target_s, req_comp=3):
    """Time the reference's jpeg.c on the host cores: returns (MP/s, kind, sample description)."""
    from oracle import oracle_lib as ol

    kind = "reference"
    try:
        o = ol.load("reference_sse2")
    except Exception:
        try:
            o = ol.load("reference")
        except Exception:
            o = ol.load("port")
            kind = "port"
    sample = streams[: max(nthreads, min(len(streams), 2 * nthreads))]
    if len(sample) < nthreads:
        sample = (sample * (nthreads // len(sample) + 1))[:nthreads]
    secs, px, fails = o.time_decode(sample, req_comp, 1, nthreads)
    reps = max(1, int(target_s / max(secs, 1e-3)))
    if reps > 1:
        secs, px, fails = o.time_decode(sample, req_comp, reps, nthreads)
    else:
        reps = 1
    mps = px / 1e6 / secs
    desc = f"{len(sample)} streams x {reps} reps, {nthreads} threads, {secs:.1f}s, {o.kind} build" + (
        f", {fails} failures" if fails else "")
    return mps, kind, desc


def bind_to_gpu_numa(local_rank):
    """Pin this rank (and the library threads it will start) to the cores of the NUMA node its GPU hangs off, so
    that the pinned staging buffers (first touch) and the parser threads sit next to the PCIe root port the
    copies go through.  Returns a description for the JSON line, or None where the topology is not exposed."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return {"gpu_bus": bus, "node": node, "bound": False}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return {"gpu_bus": bus, "node": node, "bound": False}
        os.sched_setaffinity(0, cpus)
        return {"gpu_bus": bus, "node": node, "bound": True, "cpus": len(cpus)}
    except Exception as e:  # no sysfs / nvml: stay unbound
        return {"bound": False, "why": str(e)[:80]}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ncores = os.cpu_count() or 1

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        streams = make_streams(args, 0)
        vals = []
        desc = kind = None
        for i in range(args.warmup + args.steps):
            mps, kind, desc = cpu_reference_run(streams, ncores, target_s=4.0 if i >= args.warmup else 1.0)
            if i >= args.warmup:
                vals.append(mps)
        v = float(np.mean(vals)) if vals else 0.0
        px_step = args.images * args.width * args.height / 1e6
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * px_step / v if v else None,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32",
                "data": "synthetic", "config": {"workload": workload_name(args), "bounded_sample": desc},
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": ncores, "kind": kind, "sample": desc},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return

    # ------------------------------------------------------------------ our arm
    import torch
    import torch.distributed as dist

    import image_codecs_b200 as ic
    from image_codecs_b200.api import PinnedBuffer

    if ic.cuda_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- this path has no CPU fallback")
    numa = bind_to_gpu_numa(local_rank) if world > 1 and not os.environ.get("BENCH_NO_NUMA") else None
    if numa is not None:
        print(f"[bench] rank {rank}: numa {numa}", file=sys.stderr)
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # stdout carries the one JSON line only
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    distinct = make_streams(args, rank)
    streams = [distinct[i % len(distinct)] for i in range(args.images)]
    px_rank = args.images * args.width * args.height

    dec = ic.BatchDecoder(device=local_rank, profile=True)
    nok = dec.prepare(streams, 3)
    if nok != args.images:
        raise SystemExit(f"bench.py: {args.images - nok} streams failed to parse")

    for _ in range(max(args.warmup, 3)):
        dec.run()
        dec.sync()
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    barrier()
    t0 = time.perf_counter()
    dev_ms = ent_ms = idct_ms = 0.0
    launches = 0
    for _ in range(args.steps):
        dec.run()
        dec.sync()
        tm = dec.timings()
        dev_ms += tm["total_ms"]
        ent_ms += tm["entropy_ms"]
        idct_ms += tm["idct_ms"]
        launches += tm["launches"]
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop()
    tm = dec.timings()
    for i in range(0, args.images, max(1, args.images // 8)):
        if not dec.status(i):
            raise SystemExit(f"bench.py: image {i} failed on the device: {dec.reason(i)}")

    dev_ms_max = max_over_ranks(dev_ms)
    total_px = sum_over_ranks(float(px_rank))
    ms_per_step = dev_ms_max / args.steps
    value = total_px / 1e6 / (ms_per_step / 1e3)

    # roofline of the fused kernel (rank-local, it does not shard differently)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s"
    algo_bytes = tm["coef_bytes"] + tm["out_bytes"]
    traffic = None
    try:  # dram__bytes_read+write of one ncu --set full capture (profiles/), scaled per image to this launch
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        fmt = "gray" if args.gray else args.subsampling
        if fmt in tj and not args.progressive and (args.width, args.height) == (1920, 1080):
            traffic = tj[fmt]["dram_bytes_per_image"] * args.images
    except Exception:
        pass
    idct_step_ms = idct_ms / args.steps
    achieved = algo_bytes / 1e9 / (idct_step_ms / 1e3) if idct_step_ms > 0 else 0.0
    ent_step_ms = ent_ms / args.steps
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    bits_per_s_sm = tm["entropy_bits"] / (ent_step_ms / 1e3) / sms if ent_step_ms > 0 else 0.0

    # ---------------- e2e: host buffers in, host buffers out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        n_e = args.e2e_images or args.images
        e_streams = streams[:n_e]
        in_bytes = sum(len(s) for s in e_streams)
        pin_in = PinnedBuffer(in_bytes + 64)
        bufs = (C.c_void_p * n_e)()
        lens = (C.c_int * n_e)()
        off = 0
        for i, s in enumerate(e_streams):
            pin_in.array[off:off + len(s)] = np.frombuffer(s, dtype=np.uint8)
            bufs[i] = pin_in.ptr + off
            lens[i] = len(s)
            off += len(s)
        per = args.width * args.height * 3
        pin_out = PinnedBuffer(per * n_e)
        outs = (C.c_void_p * n_e)()
        caps = (C.c_size_t * n_e)()
        for i in range(n_e):
            outs[i] = pin_out.ptr + i * per
            caps[i] = per
        st = (C.c_int * n_e)()
        # chunk size of the library's host pipeline (a public tuning parameter): about 3 GB of pixels per chunk
        e2e_chunk = int(min(1024, max(64, 2 ** round(np.log2(3.2e9 / (args.width * args.height * 3))))))
        e_dec = ic.BatchDecoder(device=local_rank, profile=False, chunk_images=int(os.environ.get("BENCH_CHUNK", "0")) or e2e_chunk,
                                host_threads=max(2, min(ncores // world, len(os.sched_getaffinity(0)))))  # ranks share the host cores
        e_dec.decode_into(bufs, lens, n_e, 3, outs, caps, None, None, None, st)  # warm-up (allocations)
        barrier()
        t0 = time.perf_counter()
        for _ in range(max(1, args.e2e_steps)):
            rc = e_dec.decode_into(bufs, lens, n_e, 3, outs, caps, None, None, None, st)
            if rc != n_e:
                raise SystemExit(f"bench.py: e2e decoded {rc}/{n_e}")
        barrier()
        e_ms = max_over_ranks(1e3 * (time.perf_counter() - t0)) / max(1, args.e2e_steps)
        e_px = sum_over_ranks(float(n_e * args.width * args.height))
        e2e = {"value": e_px / 1e6 / (e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": in_bytes,
               "d2h_bytes_per_step": per * n_e, "ms_per_step": e_ms, "images_per_step": n_e}
        e_dec.close()
        pin_in.free()
        pin_out.free()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            mps, kind, desc = cpu_reference_run(distinct, ncores, target_s=10.0)
            cpu = {"value": mps, "unit": UNIT, "cores": ncores, "kind": kind, "sample": desc}
        except Exception as e:  # the checker is optional for the bench line
            cpu = {"value": None, "unit": UNIT, "cores": ncores, "kind": "unavailable", "sample": str(e)}

    if rank == 0:
        deq = "true" if args.progressive else "false"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int16 coefficients, int32 arithmetic", "data": "synthetic",
            "config": {"workload": workload_name(args), "distinct_streams": len(distinct),
                       "cache": "inputs larger than L2 (coefficient slab %.1f GB, pixels %.1f GB per GPU)" % (
                           tm["coef_bytes"] / 1e9, tm["out_bytes"] / 1e9),
                       "req_comp": 3, "parallelism": f"image-sharded x{world}, no collective",
                       "numa_binding": numa},
            "clocks": clocks,
            "e2e": e2e,
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": ("k_fused1x1_strip<3,1,%s>" % deq) if args.gray
                         else ("k_fused420_strip<3,%s>" % deq) if args.subsampling == "420"
                         else ("k_fused1x1_strip<3,3,%s>" % deq) if args.subsampling == "444" else "k_generic_idct + k_generic_color", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                         "traffic_source": "ncu dram__bytes_read+write per image (profiles/traffic.json) x images",
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": algo_bytes,
                         "ms_per_launch": idct_step_ms},
            "cpu_baseline": cpu,
            "stages": {"entropy_ms": ent_step_ms, "idct_color_ms": idct_step_ms,
                       "entropy_bits_per_s_per_sm": bits_per_s_sm, "entropy_gbit_per_s": bits_per_s_sm * sms / 1e9,
                       "selfsync_scans": tm.get("selfsync_scans"), "selfsync_rounds": tm.get("selfsync_rounds"),
                       "wall_ms_per_step": wall_ms / args.steps},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
