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

`--config` picks the BASELINE.json configuration (default c2, the one the metric is quoted on):
  c1  one 512x512 4:2:0 q90 stream written by the jpeg_write.c-identical GPU writer, through the per-image
      stbi_load_from_memory (p50 / p99 latency, reference beside it on 1 host thread)
  c2  1024 x 1920x1080 baseline 4:2:0 q90, restart interval 16, per GPU
  c3  1080p 4:4:4 colour + 8-bit grayscale WITHOUT restart markers (self-synchronising entropy path)
  c4  1080p progressive 4:2:0 without restart markers
  c5  8192 x 3840x2160 4:2:0 q90 restart 16 in total, sharded over the ranks, each shard processed in
      device-resident chunks (a step = the whole shard)
After the timed region every run fetches images back and compares them with the oracle byte for byte; a
mismatch aborts the line.
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
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c4", "c5"])
    ap.add_argument("--total-images", type=int, default=8192, help="c5: images of the whole job (all ranks)")
    ap.add_argument("--chunk", type=int, default=0, help="device-resident images per chunk (0 = all of --images)")
    ap.add_argument("--mix-gray", type=int, default=0, help="every second image of the batch is a grayscale file")
    ap.add_argument("--check-images", type=int, default=4, help="images compared with the oracle after the timed region")
    ap.add_argument("--no-pageable", action="store_true")
    ap.add_argument("--e2e-images", type=int, default=0, help="images per e2e step (0 = same as --images)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    given = {a.split("=")[0] for a in sys.argv[1:] if a.startswith("--")}
    world = int(os.environ.get("WORLD_SIZE", "1"))

    def preset(**kw):
        for k, v in kw.items():
            if "--" + k.replace("_", "-") not in given:
                setattr(args, k, v)

    if args.config == "c1":
        preset(images=1, width=512, height=512, distinct=1, restart=0)
    elif args.config == "c3":
        preset(images=256, subsampling="444", restart=0, mix_gray=1, steps=10)
    elif args.config == "c4":
        preset(images=256, progressive=1, restart=0, steps=5, e2e_steps=2)
    elif args.config == "c5":
        preset(images=max(1, args.total_images // world), width=3840, height=2160, distinct=64, chunk=256, steps=3,
               e2e_steps=1)
    return args


def make_streams(args, rank):
    from tools import testenc as te

    kw = dict(quality=90, subsampling=args.subsampling, restart_interval=args.restart, progressive=args.progressive)
    if args.gray:
        kw["colorspace"] = te.CS_GRAY
    d = max(1, min(args.distinct, args.images))
    if args.mix_gray and not args.gray:
        # c3: 4:4:4 colour and 8-bit grayscale files alternate in one batch
        col = te.encode_many((d + 1) // 2, args.width, args.height, kind=-1, base_seed=1000 + 7919 * rank, **kw)
        kw["colorspace"] = te.CS_GRAY
        kw.pop("subsampling")
        gry = te.encode_many(d // 2, args.width, args.height, kind=-1, base_seed=1500 + 7919 * rank, **kw)
        out = []
        for i in range(d):
            out.append(col[i // 2] if i % 2 == 0 else gry[i // 2])
        return out
    return te.encode_many(d, args.width, args.height, kind=-1, base_seed=1000 + 7919 * rank, **kw)


def workload_name(args):
    if args.config == "c1":
        return "single 512x512 baseline 4:2:0 q90 JPEG (jpeg_write.c byte stream) via stbi_load_from_memory"
    fmt = "gray" if args.gray else (args.subsampling + "+gray" if args.mix_gray else args.subsampling)
    name = (f"{args.images}x {args.width}x{args.height} {'progressive' if args.progressive else 'baseline'} "
            f"{fmt} q90 restart{args.restart} per GPU")
    if args.config == "c5":
        name += f" ({args.total_images} images over all ranks, device-resident chunks of {args.chunk or args.images})"
    return name


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


_ZIGZAG = [0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35,
           42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63]


def huffman_symbols_per_bit(streams, max_streams=8):
    """SURVEY 8(d) also asks for symbols/s of the entropy stage.  For baseline files the number of Huffman symbols
    follows from the coefficients (stbi__jpeg_decode_block, jpeg.c:308-370): per block one DC symbol, one per non-zero
    AC coefficient, one ZRL per 16 zeros inside a run, and an EOB unless position 63 is non-zero.  Counted with the
    oracle's coefficient capture (checker use, outside every timed region) on a sample of the distinct streams;
    returns (symbols per entropy-coded bit, description)."""
    from oracle import oracle_lib as ol
    import image_codecs_b200 as ic

    o = ol.load()
    zz = np.array(_ZIGZAG)
    k = np.arange(64)
    syms = bits = 0
    n = 0
    for st in streams[:max_streams]:
        cap = o.decode_capture(st, 3)
        info = ic.inspect(st, 3)
        if cap is None or not info or info.get("progressive"):
            return None, "baseline files only"
        for coef in cap[1]:
            a = coef.reshape(-1, 64)[:, zz] != 0
            a[:, 0] = True  # the DC position anchors the first run
            last = np.maximum.accumulate(np.where(a, k, 0), axis=1)  # index of the last non-zero at or before k
            run = k[1:] - last[:, :-1] - 1  # zeros in front of position k (k = 1..63)
            ac = a[:, 1:]
            syms += a.shape[0] + int(ac.sum()) + int((run // 16)[ac].sum()) + int((last[:, 63] < 63).sum())
        bits += 8 * info["entropy_bytes"]
        n += 1
    return syms / bits, f"{n} of the distinct streams, symbols counted from the oracle's coefficients (MCU-padded blocks)"


def compare_with_oracle(stream, got, tag):
    """Checker use of the oracle (never timed here): abort the bench line when a decoded image differs."""
    from oracle import oracle_lib as ol

    want, *_ = ol.load().load_from_memory(stream, 3)
    if want is None or got is None or got.shape != want.shape or not np.array_equal(got, want):
        raise SystemExit(f"bench.py: PIXEL MISMATCH against the oracle on {tag} -- no bench line")


def check_pixels(args, dec, streams, rank):
    n = len(streams)
    if args.check_images < 0:  # kernel timing experiments with deliberately broken variants (tools/bench_variants.sh)
        return "SKIPPED"
    picks = sorted({(n * k) // max(1, args.check_images) for k in range(max(1, args.check_images))} | {n - 1})
    for i in picks:
        compare_with_oracle(streams[i], dec.fetch(i), f"image {i} of rank {rank}")
    return f"{len(picks)} images per rank bit-exact vs the oracle after the timed region"


def d2h_ceiling(torch, local_rank, nbytes, barrier, max_over_ranks):
    """GB/s this rank reaches copying nbytes device -> pinned host in 64 MB pieces while all ranks do the same."""
    dev = torch.device("cuda", local_rank)
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    piece = 64 << 20

    def once():
        for off in range(0, nbytes, piece):
            h[off:off + piece].copy_(d[off:off + piece], non_blocking=True)
        torch.cuda.synchronize()

    once()
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        once()
    barrier()
    ms = max_over_ranks(1e3 * (time.perf_counter() - t0)) / 3
    del d, h
    return nbytes / 1e6 / ms


def run_c1(args, ic, distinct, ncores, local_rank):
    """BASELINE.json configs[0]: ONE 512x512 4:2:0 q90 stream as jpeg_write.c writes it (here: the byte-identical GPU
    writer of this library), loaded through the per-image stbi_load_from_memory; latency percentiles, with the
    reference's own CPU decode of the same stream on one host thread beside it."""
    from tools import testenc as te

    img = te.synth_rgb(args.width, args.height, 0, 12345)
    stream = ic.write_jpg_to_bytes(img, 90)
    if not stream:
        raise SystemExit("bench.py: the GPU writer failed")
    for _ in range(20):
        got, *_ = ic.stbi_load_from_memory(stream, 3)
    compare_with_oracle(stream, got, "c1 image")
    lat = []
    n = max(200, args.steps * 50)
    for _ in range(n):
        t0 = time.perf_counter()
        got, *_ = ic.stbi_load_from_memory(stream, 3)
        lat.append(1e3 * (time.perf_counter() - t0))
    lat = np.sort(np.array(lat))
    p50, p99 = float(lat[len(lat) // 2]), float(lat[int(len(lat) * 0.99)])
    mp = args.width * args.height / 1e6
    cpu = None
    if not args.no_cpu:
        from oracle import oracle_lib as ol

        try:
            o = ol.load("reference_sse2")
            kind = "reference"
        except Exception:
            o = ol.load()
            kind = "reference" if o.kind.startswith("reference") else "port"
        secs, px, fails = o.time_decode([stream], 3, 2000, 1)
        cpu = {"value": px / 1e6 / secs, "unit": UNIT, "cores": 1, "kind": kind,
               "sample": f"1 stream x 2000 reps, 1 thread, {secs:.1f}s, {o.kind} build", "ms_per_image": 1e3 * secs / 2000}
    d = ic.BatchDecoder(device=local_rank, profile=True)  # same stream through a batch object: kernels per load
    d.prepare([stream], 3)
    d.run()
    d.sync()
    per_load = d.timings()["launches"]
    d.close()
    line = {"metric": METRIC, "value": mp / (p50 / 1e3), "unit": UNIT, "n_gpus": 1, "steps": n, "warmup": 20,
            "ms_per_step": p50, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/int16 coefficients, int32 arithmetic", "data": "synthetic",
            "config": {"workload": workload_name(args), "name": "c1", "stream_bytes": len(stream), "pixel_check": "bit-exact vs the oracle"},
            "latency_ms": {"p50": p50, "p99": p99, "min": float(lat[0]), "mean": float(lat.mean())},
            "e2e": {"value": mp / (float(lat.mean()) / 1e3), "unit": UNIT, "h2d_bytes_per_step": len(stream),
                    "d2h_bytes_per_step": args.width * args.height * 3, "ms_per_step": float(lat.mean()),
                    "buffers": "pageable (malloc'ed result, as the reference returns)"},
            "gpu_launches": per_load * n, "roofline": None, "cpu_baseline": cpu}
    emit(line)


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


_REAL_STDOUT = None


def emit(line: dict) -> None:
    """The ONE JSON line goes to the real stdout; everything else this process (NCCL, torch, the libraries) prints on
    fd 1 was pointed at stderr at start-up."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)  # e.g. "NCCL version ..." is written to fd 1 by the library itself
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ncores = os.cpu_count() or 1

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        streams = make_streams(args, 0)
        nthreads = ncores
        if args.config == "c1":  # configs[0]: one jpeg_write.c stream, one host thread (the reference is single-threaded)
            from oracle import oracle_lib as ol
            from tools import testenc as te

            streams = [ol.load().write_jpg(te.synth_rgb(args.width, args.height, 0, 12345), 90)]
            nthreads = 1
        vals = []
        desc = kind = None
        for i in range(args.warmup + args.steps):
            mps, kind, desc = cpu_reference_run(streams, nthreads, target_s=4.0 if i >= args.warmup else 1.0)
            if i >= args.warmup:
                vals.append(mps)
        v = float(np.mean(vals)) if vals else 0.0
        px_step = args.images * args.width * args.height / 1e6
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * px_step / v if v else None,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32",
                "data": "synthetic", "config": {"workload": workload_name(args), "name": args.config, "bounded_sample": desc},
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": nthreads, "kind": kind, "sample": desc},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        emit(line)
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
    if args.config == "c1":
        run_c1(args, ic, distinct, ncores, local_rank)
        return
    # device-resident chunk: the shard of this rank is processed `chunk` images at a time (c5: 8192 x 2160p needs
    # 199 GB of coefficients + 204 GB of pixels in total); the streams cycle through the distinct ones, so every
    # chunk holds the same files and ONE prepared chunk stands for all of them
    chunk = min(args.images, args.chunk) if args.chunk > 0 else args.images
    reps = (args.images + chunk - 1) // chunk
    n_img = reps * chunk  # images per rank per step (rounded up to whole chunks)
    streams = [distinct[i % len(distinct)] for i in range(chunk)]
    px_rank = n_img * args.width * args.height

    dec = ic.BatchDecoder(device=local_rank, profile=True)
    nok = dec.prepare(streams, 3)
    if nok != chunk:
        raise SystemExit(f"bench.py: {chunk - nok} streams failed to parse")

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
    per_launch_ms = []
    for _ in range(args.steps * reps):
        dec.run()
        dec.sync()
        tm = dec.timings()
        per_launch_ms.append(tm["total_ms"])
        dev_ms += tm["total_ms"]
        ent_ms += tm["entropy_ms"]
        idct_ms += tm["idct_ms"]
        launches += tm["launches"]
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop()
    tm = dec.timings()
    for i in range(chunk):
        if not dec.status(i):
            raise SystemExit(f"bench.py: image {i} failed on the device: {dec.reason(i)}")
    # ---- pixel check: what the timed runs left in HBM must be the reference's bytes (checker use of the oracle)
    checked = check_pixels(args, dec, streams, rank)

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
        if fmt in tj and not args.progressive and not args.mix_gray and (args.width, args.height) == (1920, 1080):
            traffic = tj[fmt]["dram_bytes_per_image"] * chunk
    except Exception:
        pass
    idct_launch_ms = idct_ms / (args.steps * reps)
    achieved = algo_bytes / 1e9 / (idct_launch_ms / 1e3) if idct_launch_ms > 0 else 0.0
    ent_step_ms = ent_ms / args.steps
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    bits_per_s_sm = tm["entropy_bits"] * reps / (ent_step_ms / 1e3) / sms if ent_step_ms > 0 else 0.0
    sym_per_bit, sym_desc = (None, "skipped (--no-cpu)")
    if rank == 0 and not args.no_cpu and not args.progressive:
        try:
            sym_per_bit, sym_desc = huffman_symbols_per_bit(distinct)
        except Exception as e:  # the symbol rate is an extra, never a reason to lose the line
            sym_desc = f"unavailable: {e}"
    dec.close()

    # ---------------- e2e: host buffers in, host buffers out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        # one call of the public batched entry point per sub-batch; c5: the shard goes through in sub-batches of
        # `chunk` images whose pinned output buffers are reused (host RAM bounds them, not the library)
        n_call = args.e2e_images or chunk
        calls = max(1, (args.images + n_call - 1) // n_call) if not args.e2e_images else 1
        e_streams = [distinct[i % len(distinct)] for i in range(n_call)]
        in_bytes = sum(len(s) for s in e_streams)
        per = args.width * args.height * 3
        e2e_chunk = int(min(1024, max(64, 2 ** round(np.log2(3.2e9 / per)))))
        host_threads = max(2, min(ncores // world, len(os.sched_getaffinity(0))))  # ranks share the host cores

        h2d_ms = [None]

        def run_e2e(pinned, steps):
            if pinned:
                buf_in, buf_out = PinnedBuffer(in_bytes + 64), PinnedBuffer(per * n_call)
                a_in, a_out, p_in, p_out = buf_in.array, buf_out.array, buf_in.ptr, buf_out.ptr
            else:  # what a plain stbi-style caller passes: malloc'ed (pageable) memory on both sides
                a_in, a_out = np.empty(in_bytes + 64, dtype=np.uint8), np.empty(per * n_call, dtype=np.uint8)
                a_out[::4096] = 0  # touch the pages outside the timed region
                p_in, p_out = a_in.ctypes.data, a_out.ctypes.data
            bufs, lens = (C.c_void_p * n_call)(), (C.c_int * n_call)()
            outs, caps = (C.c_void_p * n_call)(), (C.c_size_t * n_call)()
            off = 0
            for i, s_ in enumerate(e_streams):
                a_in[off:off + len(s_)] = np.frombuffer(s_, dtype=np.uint8)
                bufs[i], lens[i] = p_in + off, len(s_)
                outs[i], caps[i] = p_out + i * per, per
                off += len(s_)
            st = (C.c_int * n_call)()
            e_dec = ic.BatchDecoder(device=local_rank, profile=False, chunk_images=int(os.environ.get("BENCH_CHUNK", "0")) or e2e_chunk,
                                    host_threads=host_threads)
            e_dec.decode_into(bufs, lens, n_call, 3, outs, caps, None, None, None, st)  # warm-up (allocations)
            barrier()
            t0 = time.perf_counter()
            for _ in range(max(1, steps) * calls):
                rc = e_dec.decode_into(bufs, lens, n_call, 3, outs, caps, None, None, None, st)
                if rc != n_call:
                    raise SystemExit(f"bench.py: e2e decoded {rc}/{n_call}")
            barrier()
            ms = max_over_ranks(1e3 * (time.perf_counter() - t0)) / max(1, steps)
            # the bytes that arrived in the caller's buffers are checked too
            got = np.array(a_out[:per]).reshape(args.height, args.width, 3)
            compare_with_oracle(e_streams[0], got, "e2e image 0 (%s buffers)" % ("pinned" if pinned else "pageable"))
            e_dec.close()
            if pinned and n_call * per * 2.1 < 60e9:
                # SURVEY 8(d) protocol point (b): parse + H2D + kernels, pixels stay in HBM (the staged form of the same API)
                s_dec = ic.BatchDecoder(device=local_rank, profile=False, host_threads=host_threads)
                xs, ys, cs = ((C.c_int * n_call)() for _ in range(3))
                L = ic.api.lib()

                def staged():
                    if L.stbi_jpeg_batch_prepare(s_dec._h, bufs, lens, n_call, 3, xs, ys, cs, st) != n_call:
                        raise SystemExit("bench.py: staged decode failed")
                    s_dec.count = n_call
                    s_dec.run()
                    s_dec.sync()

                staged()
                barrier()
                t0 = time.perf_counter()
                for _ in range(max(1, steps) * calls):
                    staged()
                barrier()
                h2d_ms[0] = max_over_ranks(1e3 * (time.perf_counter() - t0)) / max(1, steps)
                s_dec.close()
            if pinned:
                buf_in.free()
                buf_out.free()
            return ms

        e_ms = run_e2e(True, args.e2e_steps)
        e_px = sum_over_ranks(float(n_call * calls * args.width * args.height))
        e2e = {"value": e_px / 1e6 / (e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": in_bytes * calls,
               "d2h_bytes_per_step": per * n_call * calls, "ms_per_step": e_ms, "images_per_step": n_call * calls,
               "steps": max(1, args.e2e_steps), "calls_per_step": calls, "buffers": "pinned host memory in and out"}
        # ceiling of that path on this box: every rank copies the same number of result bytes device -> pinned host
        # at the same time (64 MB pieces on one stream), nothing else running
        ceil_gbs = d2h_ceiling(torch, local_rank, min(per * n_call, 1 << 31), barrier, max_over_ranks) * world
        e2e["d2h_ceiling_gbs_all_ranks"] = ceil_gbs
        e2e["d2h_achieved_gbs_all_ranks"] = e_px * 3 / 1e9 / (e_ms / 1e3)
        e2e["frac_of_d2h_ceiling"] = e2e["d2h_achieved_gbs_all_ranks"] / ceil_gbs if ceil_gbs else None
        if h2d_ms[0]:
            e2e["h2d_decode"] = {"value": e_px / 1e6 / (h2d_ms[0] / 1e3), "unit": UNIT, "ms_per_step": h2d_ms[0],
                                 "what": "marker parse + H2D + kernels through the staged calls, pixels left in HBM"}
        if not args.no_pageable:
            p_ms = run_e2e(False, max(1, min(2, args.e2e_steps)))
            e2e["pageable"] = {"value": e_px / 1e6 / (p_ms / 1e3), "unit": UNIT, "ms_per_step": p_ms,
                               "buffers": "malloc'ed (pageable) host memory in and out"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            mps, kind, desc = cpu_reference_run(distinct, ncores, target_s=10.0)
            cpu = {"value": mps, "unit": UNIT, "cores": ncores, "kind": kind, "sample": desc}
        except Exception as e:  # the checker is optional for the bench line
            cpu = {"value": None, "unit": UNIT, "cores": ncores, "kind": "unavailable", "sample": str(e)}

    if rank == 0:
        deq = "true" if args.progressive else "false"
        k444, kgray = "k_fused1x1_strip<3,3,%s>" % deq, "k_fused1x1_strip<3,1,%s>" % deq
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8/int16 coefficients, int32 arithmetic", "data": "synthetic",
            "config": {"workload": workload_name(args), "name": args.config, "distinct_streams": len(distinct),
                       "cache": "inputs larger than L2 (coefficient slab %.1f GB, pixels %.1f GB per launch)" % (
                           tm["coef_bytes"] / 1e9, tm["out_bytes"] / 1e9),
                       "req_comp": 3, "parallelism": f"image-sharded x{world}, no collective",
                       "chunks_per_step": reps, "pixel_check": checked, "numa_binding": numa},
            "clocks": clocks,
            "e2e": e2e,
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": kgray if args.gray else (k444 + " + " + kgray) if args.mix_gray
                         else ("k_fused420_strip<3,%s>" % deq) if args.subsampling == "420"
                         else k444 if args.subsampling == "444" else ("k_fused422_strip<3,%s>" % deq) if args.subsampling == "422"
                         else "k_generic_idct + k_generic_color", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                         "traffic_source": "ncu dram__bytes_read+write per image (profiles/traffic.json) x images",
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": algo_bytes,
                         "ms_per_launch": idct_launch_ms},
            "cpu_baseline": cpu,
            "stages": {"entropy_ms": ent_step_ms, "idct_color_ms": idct_ms / args.steps,
                       "entropy_bits_per_s_per_sm": bits_per_s_sm, "entropy_gbit_per_s": bits_per_s_sm * sms / 1e9,
                       "entropy_gsymbols_per_s": (bits_per_s_sm * sms * sym_per_bit / 1e9) if sym_per_bit else None,
                       "entropy_symbols_sample": sym_desc,
                       "selfsync_scans": tm.get("selfsync_scans"), "selfsync_rounds": tm.get("selfsync_rounds"),
                       "wall_ms_per_step": wall_ms / args.steps,
                       # best and median of the timed launches (device events around one whole batch / chunk)
                       "device_ms_per_launch_best": min(per_launch_ms), "device_ms_per_launch_median": float(np.median(per_launch_ms)),
                       "timed_launches": len(per_launch_ms)},
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
