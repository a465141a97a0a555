#!/usr/bin/env python
"""bench.py -- headline benchmark of the JPG target-size hot path on B200.

Metric (BASELINE.json): megapixels/s and images/s at a 1 MB target on 1/2/4/8 B200,
next to the reference's host-CPU path.  Workload = BASELINE.json configs[0]'s image
set (synthetic 12 MP 4000x3000 JPG pixels, -t 1048576 -q 0.25), the configuration the
metric's "1 MB target" names; it is sharded by image over the N GPUs (weak scaling:
fixed images per GPU per step, no collective on the data path).

A "step" = one pass of compressJpgWithTargetSize (core/ImageCompressionJpg.java:77-122)
over a batch of images per GPU.  `value` times it with the decoded pixels already
resident in HBM; `e2e` times the same call through the C ABI with pinned HOST pixel
buffers (H2D of the pixels and D2H of the JFIF bytes inside the timed region).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (w, h, target, init_q, description)
    "cfg1": (4000, 3000, 1048576, 0.25, "BASELINE configs[0]: synthetic 12MP (4000x3000) JPG pixels, -t 1048576 -q 0.25"),
    "cfg2": (6000, 4000, 524288, 0.25, "BASELINE configs[1]: synthetic 24MP (6000x4000) JPG pixels, 512KB target, -q 0.25"),
}
POOL = 8            # distinct images per GPU: 8 x 36 MB = 288 MB > 126 MB L2


def make_pool(w, h, n, seed0):
    """n distinct natural-statistics images (half generated, half mirrored copies)."""
    from image_compression_b200 import synth
    base = [synth.natural_image(seed0 + i, w, h) for i in range((n + 1) // 2)]
    pool = []
    for im in base:
        pool.append(im)
        if len(pool) < n:
            pool.append(np.ascontiguousarray(im[:, ::-1]))
    return pool[:n]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        torch.cuda.set_device(local)
        dist_mod.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    return rank, local, world, dist


def barrier(dist):
    if dist is not None:
        dist.barrier()


def max_over_ranks(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(dist, local, x: float) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ---------------------------------------------------------------- CPU arm (oracle)
def cpu_fit_batch(pool, n_images, threads, init_q, target):
    """The reference's work pattern on the host cores: per image a float32 bisection with a
    FULL re-encode per step plus the final save encode (ImageCompressionJpg.java:158-200,
    :255-260), one image per pool thread (CompressionBatch.java:64-88).  Returns seconds."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as O            # allowed here: this IS the timed CPU baseline
    O.lib()

    def one(i):
        rc, data, bq, sc, _, ne = O.jpg_fit(pool[i % len(pool)], init_q, target)
        return rc, len(data), ne

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        res = list(ex.map(one, range(n_images)))
    dt = time.perf_counter() - t0
    return dt, res


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w, h, target, q, desc = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    threads = cores
    n_images = max(threads, 8)
    pool = make_pool(w, h, 4, seed0=0)
    for _ in range(args.warmup):
        cpu_fit_batch(pool, min(n_images, threads), threads, q, target)
    t = 0.0
    for _ in range(args.steps):
        dt, _ = cpu_fit_batch(pool, n_images, threads, q, target)
        t += dt
    imgs = n_images * args.steps
    ips = imgs / t
    mp = ips * w * h / 1e6
    sample = f"{n_images} images/step x {args.steps} steps, {threads} threads, oracle port of IJG-6b/JDK encode, 5+1 encodes per image"
    line = {
        "impl": "reference", "metric": "megapixels_per_s", "value": mp, "unit": "MP/s", "images_per_s": ips,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": desc, "images_per_step": n_images, "note":
                   "reference jar unavailable on this box (no JVM/Gradle); CPU port of the same algorithm"},
        "cpu_baseline": {"value": mp, "unit": "MP/s", "images_per_s": ips, "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": mp, "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------- GPU arm
def run_step(eng, submit, n_images, slots, outs, cursor):
    """Pipelined batch: keep `slots` fits in flight, FIFO wait.  Returns list of results."""
    inflight, results = [], []
    for i in range(n_images):
        if len(inflight) == slots:
            results.append(eng.wait(inflight.pop(0)))
        inflight.append(submit((cursor + i), outs[i % slots]))
    while inflight:
        results.append(eng.wait(inflight.pop(0)))
    return results


def timed_region(eng, dist, local, steps, step_fn):
    barrier(dist)
    eng.mark(0, 0)
    t0 = time.perf_counter()
    for s in range(steps):
        step_fn(s)
    eng.mark(0, 1)
    ms = eng.elapsed_ms(0)                       # CUDA events over every slot stream
    wall = (time.perf_counter() - t0) * 1e3
    barrier(dist)
    return max_over_ranks(dist, local, ms), wall


def run_gpu(args):
    rank, local, world, dist = dist_setup(args.gpus)
    import image_compression_b200 as pkg
    w, h, target, q, desc = WORKLOADS[args.workload]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s"

    slots = args.slots
    eng = pkg.Engine(devices=[local], slots_per_gpu=slots, max_w=w, max_h=h)
    pool = make_pool(w, h, POOL, seed0=1000 * rank)
    resident = [eng.upload_image(p, 0) for p in pool]
    pinned = []
    for p in pool:
        a = eng.alloc_pinned((h, w, 3))
        a[...] = p
        pinned.append(a)
    outs = [eng.alloc_pinned((target + 16,)) for _ in range(slots)]
    batch = args.batch

    def submit_res(i, out):
        return eng.submit_jpg_fit_image(resident[i % POOL], q, target, out)

    def submit_host(i, out):
        return eng.submit_jpg_fit(pinned[i % POOL], q, target, out)

    # correctness of what is being timed: every image must fit, one is checked against the oracle
    res = run_step(eng, submit_res, POOL, slots, outs, 0)
    assert all(r.status == 0 and 0 < r.out_len <= target for r in res), [(r.status, r.out_len) for r in res]
    parity = "skipped"
    if rank == 0 and not args.no_check:
        from oracle import oracle as O          # checker only, outside every timed region
        r = eng.jpg_fit(pool[0], q, target)
        rc, ref, rq, rs, _, _ = O.jpg_fit(pool[0], q, target)
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and r.data == ref, "GPU result differs from oracle"
        parity = "bit-exact vs oracle on pool[0] (%d bytes, q=%.7f)" % (len(ref), rq)

    for s in range(args.warmup):
        run_step(eng, submit_res, batch, slots, outs, s * batch)
    k0 = eng.stats()["kernels_launched"]
    sampler = ClockSampler(local)
    sampler.start()
    ms, wall = timed_region(eng, dist, local, args.steps,
                            lambda s: run_step(eng, submit_res, batch, slots, outs, s * batch))
    launches = eng.stats()["kernels_launched"] - k0

    for s in range(max(1, args.warmup // 2)):
        run_step(eng, submit_host, batch, slots, outs, s * batch)
    st0 = eng.stats()
    ms_e2e, wall_e2e = timed_region(eng, dist, local, args.steps,
                                    lambda s: run_step(eng, submit_host, batch, slots, outs, s * batch))
    clocks = sampler.stop()
    st1 = eng.stats()
    h2d_step = (st1["h2d_bytes"] - st0["h2d_bytes"]) / args.steps
    d2h_step = (st1["d2h_bytes"] - st0["d2h_bytes"]) / args.steps

    # per-kernel device times, live, CUDA events on the launching stream; pool > L2 so K1 reads HBM
    prof = [eng.profile_jpg_fit_image(resident[i % POOL], q, target, 1) for i in range(2 * POOL)][POOL:]
    xf_ms = statistics.mean(p["transform_ms"] for p in prof)
    ent_ms = statistics.mean(p["entropy_kernel_ms"] for p in prof)
    ent_n = statistics.mean(p["entropy_kernel_launches"] for p in prof)
    search_ms = statistics.mean(p["search_ms"] for p in prof)
    mean_out = statistics.mean(r.out_len for r in res)
    nblocks = 6 * ((w + 15) // 16) * ((h + 15) // 16)
    xf_bytes = 3 * w * h + 128 * nblocks
    ent_bytes = 128 * nblocks + mean_out
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass

    def roof(name, nbytes, ms_launch, share):
        a = nbytes / (ms_launch * 1e-3) / 1e9
        return {"kernel": name, "bound": "hbm", "achieved": a, "peak": peak_gbs, "unit": "GB/s",
                "frac": a / peak_gbs, "traffic": traffic.get(name), "algorithmic_bytes_per_launch": nbytes,
                "ms_per_launch": ms_launch, "share_of_image_device_time": share, "peak_source": peak_src}

    tot = xf_ms + search_ms
    r_xf = roof("k_transform", xf_bytes, xf_ms, xf_ms / tot)
    r_ent = roof("k_entropy", ent_bytes, ent_ms / max(ent_n, 1), ent_ms / tot)
    r_ent["note"] = "integer-pipe bound (quantise + Huffman bit work on L2-resident coefficients), not HBM bound"
    dominant = r_ent if ent_ms > xf_ms else r_xf

    n_img = batch * args.steps
    ips_local = n_img / (ms * 1e-3)
    ips = ips_local * world                       # every rank runs the same per-GPU batch (weak scaling)
    ips_e2e = n_img * world / (ms_e2e * 1e-3)
    mpx = w * h / 1e6
    launches_all = int(sum_over_ranks(dist, local, float(launches)))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        n_cpu = max(cores, 8)
        dt, _ = cpu_fit_batch(pool, n_cpu, cores, q, target)
        cpu = {"value": n_cpu / dt * mpx, "unit": "MP/s", "images_per_s": n_cpu / dt, "cores": cores, "kind": "port",
               "sample": f"{n_cpu} of the same 12MP images, one per thread, {cores} threads, {dt:.1f}s wall; "
                         "oracle port (scalar IJG-6b restatement), 5 search encodes + 1 save encode per image"}

    if rank == 0:
        line = {
            "metric": "megapixels_per_s", "value": ips * mpx, "unit": "MP/s", "images_per_s": ips,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": {"workload": desc, "images_per_gpu_per_step": batch, "slots_per_gpu": slots,
                       "l2": f"pool of {POOL} distinct {w}x{h} images per GPU ({POOL * w * h * 3 >> 20} MiB) > 126 MB L2",
                       "parallelism": f"images sharded over {world} GPU(s), no collective", "parity_check": parity},
            "clocks": clocks,
            "e2e": {"value": ips_e2e * mpx, "unit": "MP/s", "images_per_s": ips_e2e,
                    "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches_all,
            "roofline": dominant, "roofline_transform": r_xf, "roofline_entropy": r_ent,
            "device_ms_per_image": {"transform": xf_ms, "search_total": search_ms, "entropy_kernels": ent_ms,
                                    "entropy_passes": ent_n},
            "host_wall_ms_per_step": {"resident": wall / args.steps, "e2e": wall_e2e / args.steps},
            "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    for im in resident:
        eng.free_image(im)
    eng.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg1", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--slots", type=int, default=6, help="stream work-queue slots per GPU")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
