#!/usr/bin/env python
"""bench.py -- headline benchmark of the JPG target-size hot path on B200.

Metric (BASELINE.json): megapixels/s and images/s at a 1 MB target on 1/2/4/8 B200,
next to the reference's host-CPU path.  Workload = BASELINE.json configs[0]'s image
set (synthetic 12 MP 4000x3000 JPG pixels, -t 1048576 -q 0.25), the configuration the
metric's "1 MB target" names; it is sharded by image over the N GPUs (weak scaling:
fixed images per GPU per step, no collective on the data path).

A "step" = one pass of compressJpgWithTargetSize (core/ImageCompressionJpg.java:77-122)
over a batch of images per GPU, driven by the library's own batch executor
(icb_run_list = CompressionBatch.java:64-108 in native code: worker threads, one cursor,
several jobs in flight per thread).  `value` times it with the decoded pixels already
resident in HBM; `e2e` times the same call with pinned HOST pixel buffers (H2D of the
pixels and D2H of the JFIF bytes inside the timed region).

The same invocation then runs short legs for the other BASELINE configs (24 MP / 512 KB,
4K and 8K PNG pixels, the warm-cache mixed list, the 100 MP latency case), each checked
against the oracle in the run, and reports them under `other_configs`; with N > 1 it
also measures ONE process driving all N GPUs (the Java host's shape) under `one_process`.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
from __future__ import annotations

import argparse
import json
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # before anything starts CUDA (icb_init asks for the same; torch may come first)
import statistics
import subprocess
import sys
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
    cache = os.environ.get("ICB_POOL_CACHE")            # A/B sessions: synthesise once per box, not once per run
    path = cache and os.path.join(cache, f"pool_{w}x{h}_{n}_{seed0}.npy")
    if path and os.path.exists(path):
        return list(np.load(path))
    base = [synth.natural_image(seed0 + i, w, h) for i in range((n + 1) // 2)]
    pool = []
    for im in base:
        pool.append(im)
        if len(pool) < n:
            pool.append(np.ascontiguousarray(im[:, ::-1]))
    if path:
        os.makedirs(cache, exist_ok=True)
        np.save(path, np.stack(pool[:n]))
    return pool[:n]


def compose(pool, w, h, channels=3, k=0):
    """A w x h image of the same statistics, tiled from the 12 MP pool (FFT synthesis of a 33 MP image takes
    ten seconds; the legs only need pixels of the right statistics and size).  Alpha = another pool image's
    green channel (smooth, full range)."""
    ph, pw = pool[0].shape[:2]
    out = np.empty((h, w, channels), np.uint8)
    for ty in range(0, h, ph):
        for tx in range(0, w, pw):
            src = pool[(k + ty // ph * 3 + tx // pw) % len(pool)]
            hh, ww = min(ph, h - ty), min(pw, w - tx)
            out[ty:ty + hh, tx:tx + ww, :3] = src[:hh, :ww]
            if channels == 4:
                out[ty:ty + hh, tx:tx + ww, 3] = pool[(k + 1 + ty // ph + tx // pw) % len(pool)][:hh, :ww, 1]
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md); one sampler
    (rank 0), 500 ms period, so that the sampling itself does not load the host."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, enabled: bool = True):
        self.gpu = gpu_index
        self.enabled = enabled
        self.proc = None

    def start(self):
        if not self.enabled:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "500"],
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


def dist_setup():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        import datetime
        torch.cuda.set_device(local)
        dist_mod.init_process_group(backend="nccl", device_id=torch.device("cuda", local),
                                    timeout=datetime.timedelta(seconds=300))
        dist = dist_mod
        # host-side barrier for the stretch in which rank 0 alone drives EVERY GPU: an NCCL barrier would park a
        # spinning kernel of the waiting ranks on the very GPUs rank 0 is using (measured r02: the leg never ended)
        global _HOST_GROUP
        _HOST_GROUP = dist_mod.new_group(backend="gloo", timeout=datetime.timedelta(seconds=300))
    return rank, local, world, dist


_HOST_GROUP = None


def host_barrier(dist):
    if dist is not None:
        dist.barrier(group=_HOST_GROUP)


def barrier(dist):
    if dist is not None:
        dist.barrier()


def _reduce(dist, local, x: float, op: str) -> float:
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
    return float(t.item())


def gather_ranks(dist, local, world, x: float):
    if dist is None:
        return [x]
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=torch.device("cuda", local))
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def peaks():
    p = {}
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    return (float(p.get("hbm_gbs", 6650.0)),
            "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in p else "fallback 6.65 TB/s")


def traffic_table():
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        return {}


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


def oracle_build_note():
    try:
        from oracle import oracle as O
        return O.build_flags()
    except Exception:
        return "?"


def fast_cpu_lines(img, target, q):
    """SURVEY 8d sanity lines, one thread each: libjpeg-turbo (Pillow) single 4:2:0 encode of the same
    pixels -- how fast a SIMD CPU encoder is, so the GPU/CPU ratio can be read against the scalar port --
    and the host decode of a q=92 JPEG of the image ("input staging", which stays on the host)."""
    try:
        import io
        from PIL import Image
        im = Image.fromarray(img)
        buf = io.BytesIO()
        t0 = time.perf_counter()
        n = 3
        for _ in range(n):
            buf.seek(0); buf.truncate()
            im.save(buf, format="JPEG", quality=25, subsampling=2, optimize=False)
        enc = (time.perf_counter() - t0) / n
        src = io.BytesIO()
        im.save(src, format="JPEG", quality=92, subsampling=2)
        t0 = time.perf_counter()
        for _ in range(n):
            src.seek(0)
            Image.open(src).convert("RGB").load()
        dec = (time.perf_counter() - t0) / n
        mp = img.shape[0] * img.shape[1] / 1e6
        return {"libjpeg_turbo_single_encode": {"value": mp / enc, "unit": "MP/s", "threads": 1, "ms": enc * 1e3,
                                                "note": "Pillow/libjpeg-turbo (SIMD) ONE q=25 4:2:0 encode; the reference "
                                                        "does 6 such encodes per image with the JDK's scalar libjpeg 6b"},
                "host_decode_input_staging": {"value": mp / dec, "unit": "MP/s", "threads": 1, "ms": dec * 1e3,
                                              "file_bytes": src.getbuffer().nbytes,
                                              "note": "Pillow decode of a q=92 4:2:0 JPEG of the same image; timed "
                                                      "separately, decode stays on the host (north_star)"}}
    except Exception as e:                                  # Pillow is optional
        return {"unavailable": repr(e)}


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
    sample = (f"{n_images} images/step x {args.steps} steps, {threads} threads, oracle port of IJG-6b/JDK encode "
              f"({oracle_build_note()}), 5+1 encodes per image")
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


# ---------------------------------------------------------------- GPU arm: timing helpers
def timed(eng, devs, fn, dist=None, local=0):
    """fn() between CUDA-event marks over every slot stream of the listed device indices; -> (device ms =
    max over devices and ranks, host wall ms)."""
    barrier(dist)
    for d in devs:
        eng.mark(d, 0)
    t0 = time.perf_counter()
    fn()
    for d in devs:
        eng.mark(d, 1)
    ms = max(eng.elapsed_ms(d) for d in devs)
    wall = (time.perf_counter() - t0) * 1e3
    barrier(dist)
    return _reduce(dist, local, ms, "MAX"), wall


def steps_for(eng, devs, fn, seconds=1.0, min_steps=3):
    """Run fn() once to size the leg, then as many times as fit `seconds`; -> (ms per step, steps)."""
    ms1, _ = timed(eng, devs, fn)
    n = max(min_steps, min(200, int(seconds * 1e3 / max(ms1, 1e-3))))
    ms, _ = timed(eng, devs, lambda: [fn() for _ in range(n)])
    return ms / n, n


def fit_items(eng, images, n, q, target, outs, cached=None, host=False):
    """n JPG fit items cycling over `images` (resident handles, or pinned host arrays with host=True)."""
    lst = []
    for i in range(n):
        im = images[i % len(images)]
        lst.append(dict(kind="jpg_fit" if host else "jpg_fit_image", image=im, init_q=q, target=target,
                        cached=cached[i % len(images)] if cached else None, out=outs[i]))
    return eng.make_items(lst)


def pinned_copy(eng, arr):
    a = eng.alloc_pinned(arr.shape)
    a[...] = arr
    return a


# ---------------------------------------------------------------- legs: the other BASELINE configs
def leg_jpg(eng, pool, name, threads, inflight, peak_gbs):
    """cfg2: 24 MP pixels, 512 KB target -- same call as the headline, other size (needs the 0.85 scale loop
    or the pruned lists depending on the image)."""
    from oracle import oracle as O
    w, h, target, q, desc = WORKLOADS[name]
    imgs = [compose(pool, w, h, 3, k) for k in range(3)]             # 3 x 72 MB > L2
    res = [eng.upload_image(im, 0) for im in imgs]
    pins = [pinned_copy(eng, im) for im in imgs]
    n, n_e2e = 24, 12
    outs = [eng.alloc_pinned((target + 16,)) for _ in range(n)]
    r = eng.jpg_fit(imgs[0], q, target)
    rc, ref, rq, rs, _, _ = O.jpg_fit(imgs[0], q, target)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs) and r.data == ref, name + ": differs from the oracle"
    items = fit_items(eng, res, n, q, target, outs)
    eng.run_list(items, threads, inflight)
    ms, steps = steps_for(eng, [0], lambda: eng.run_list(items, threads, inflight))
    assert all(x.status == 0 for x in items.results())
    st0 = eng.stats()
    items_h = fit_items(eng, pins, n_e2e, q, target, outs, host=True)
    ms_e, steps_e = steps_for(eng, [0], lambda: eng.run_list(items_h, threads, inflight))
    st1 = eng.stats()
    mpx = w * h / 1e6
    out = {"workload": desc, "value": n / ms * 1e3 * mpx, "unit": "MP/s", "images_per_s": n / ms * 1e3,
           "ms_per_step": ms, "steps": steps, "images_per_step": n,
           "e2e": {"value": n_e2e / ms_e * 1e3 * mpx, "unit": "MP/s", "images_per_s": n_e2e / ms_e * 1e3,
                   "h2d_bytes_per_step": (st1["h2d_bytes"] - st0["h2d_bytes"]) / (steps_e + 1),
                   "d2h_bytes_per_step": (st1["d2h_bytes"] - st0["d2h_bytes"]) / (steps_e + 1)},
           "parity_check": "bit-exact vs oracle on image 0 (%d bytes, q=%.7f, scale=%.4f)" % (len(ref or b""), rq, rs)}
    try:        # the same leg from JPEG files (decode on the GPU): what "100,000 synthetic 24MP JPGs" start from
        import io
        from PIL import Image
        files = []
        for im in imgs:
            bio = io.BytesIO()
            Image.fromarray(im).save(bio, "JPEG", quality=92, subsampling=2)
            files.append(np.frombuffer(bio.getvalue(), np.uint8))
        fpins = []
        for f in files:
            a = eng.alloc_pinned((f.size,))
            a[:] = f
            fpins.append(a)
        items_f = eng.make_items([dict(kind="jpg_file_fit", file=fpins[i % len(fpins)], init_q=q, target=target, out=outs[i])
                                  for i in range(n_e2e)])
        eng.run_list(items_f, threads, inflight)
        rf = items_f.results()[0]
        px = O.jpeg_decode(files[0].tobytes())
        rc2, ref2, rq2, rs2, _, _ = O.jpg_fit(px, q, target)
        assert (rf.status, rf.best_q, rf.scale) == (rc2, rq2, rs2) and bytes(outs[0][: rf.out_len]) == ref2, name + " from files: differs from the oracle"
        sf0 = eng.stats()
        ms_f, steps_f = steps_for(eng, [0], lambda: eng.run_list(items_f, threads, inflight))
        sf1 = eng.stats()
        out["e2e_file"] = {"value": n_e2e / ms_f * 1e3 * mpx, "unit": "MP/s", "images_per_s": n_e2e / ms_f * 1e3,
                           "h2d_bytes_per_step": (sf1["h2d_bytes"] - sf0["h2d_bytes"]) / (steps_f + 1),
                           "file_bytes_mean": float(np.mean([f.size for f in files])),
                           "parity_check": "bit-exact vs oracle decode + fit of file 0 (q=%.7f, scale=%.4f)" % (rq2, rs2)}
        for a in fpins:
            eng.free_pinned(a)
    except Exception as e:
        out["e2e_file"] = {"error": repr(e)}
    xf_ms, _ = eng.profile_transform_series(res, 3, list_q=q)
    nblocks = 6 * ((w + 15) // 16) * ((h + 15) // 16)
    xb = 3 * w * h + 128 * nblocks
    out["roofline"] = {"kernel": "k_transform", "bound": "hbm", "achieved": xb / xf_ms / 1e6, "peak": peak_gbs,
                       "unit": "GB/s", "frac": xb / xf_ms / 1e6 / peak_gbs, "traffic": None,
                       "algorithmic_bytes_per_launch": xb, "ms_per_launch": xf_ms}
    for x in res:
        eng.free_image(x)
    for a in pins + outs:
        eng.free_pinned(a)
    return out


def leg_png(eng, pool, w, h, threads, inflight, peak_gbs, traffic):
    """cfg3: 4K / 8K RGBA8 PNG pixels into the 1920x1920 box (core/ImageCompressionPng.java:49-66 ->
    tools/ImageTools.java:7-26), then the PNG file itself (:70) on the GPU.  MP/s counts SOURCE pixels."""
    import image_compression_b200 as pkg
    from oracle import oracle as O
    dw, dh, _ = pkg.png_target_dims(w, h, 1920, 1920)
    npool = 2 if w > 4000 else 4                                     # 2 x 133 MB / 4 x 33 MB > L2
    imgs = [compose(pool, w, h, 4, k) for k in range(npool)]
    res = [eng.upload_image(im, 0) for im in imgs]
    pins = [pinned_copy(eng, im) for im in imgs]
    got = eng.resize_to(imgs[0], dw, dh)
    assert np.array_equal(got, O.resize(imgs[0], dw, dh, fmt=O.FMT_RGBA8)), "resize differs from the oracle"
    rs_ms = statistics.median(eng.profile_resize_series(x, dw, dh, 20) for x in res)
    png_ms = eng.profile_resize_series(res[0], dw, dh, 6, with_png=True)
    n = 12
    douts = [eng.alloc_pinned((dw * dh * 4,)) for _ in range(n)]
    items = eng.make_items([dict(kind="png_fit", image=pins[i % npool], min_w=1920, min_h=1920, out=douts[i])
                            for i in range(n)])
    st0 = eng.stats()
    ms_e, steps_e = steps_for(eng, [0], lambda: eng.run_list(items, threads, inflight))
    st1 = eng.stats()
    assert all(x.status == 0 for x in items.results())
    fouts = [eng.alloc_pinned((dw * dh * 5,)) for _ in range(n)]
    items_f = eng.make_items([dict(kind="png_compress", image=pins[i % npool], min_w=1920, min_h=1920, out=fouts[i])
                              for i in range(n)])
    ms_f, _ = steps_for(eng, [0], lambda: eng.run_list(items_f, threads, inflight), seconds=0.6)
    rf = items_f.results()
    assert all(x.status == 0 for x in rf)
    import zlib
    png = bytes(fouts[0][: rf[0].out_len])
    idat, pos = [], 8
    while pos < len(png):
        ln = int.from_bytes(png[pos:pos + 4], "big")
        if png[pos + 4:pos + 8] == b"IDAT":
            idat.append(png[pos + 8:pos + 8 + ln])
        pos += 12 + ln
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), np.uint8).reshape(dh, 1 + dw * 4)
    assert raw.shape[0] == dh                                        # inflates, Adler-32 verified by zlib
    # sector-granular source bytes the 2x2 taps touch (scale .5: every pixel; scale .25: half of every 32 B sector row pair)
    touched = (w * h * 4 if w < 4000 else w * h * 2) + dw * dh * 4
    mpx = w * h / 1e6
    out = {"workload": f"BASELINE configs[2]: {w}x{h} RGBA8 PNG pixels -> {dw}x{dh} (1920x1920 box)",
           "value": mpx / rs_ms * 1e3, "unit": "MP/s", "images_per_s": 1e3 / rs_ms,
           "note": "value = resize kernel, source and result in HBM (back-to-back launches, CUDA events)",
           "e2e": {"value": n / ms_e * 1e3 * mpx, "unit": "MP/s", "images_per_s": n / ms_e * 1e3,
                   "h2d_bytes_per_step": (st1["h2d_bytes"] - st0["h2d_bytes"]) / (steps_e + 1),
                   "d2h_bytes_per_step": (st1["d2h_bytes"] - st0["d2h_bytes"]) / (steps_e + 1),
                   "note": "pinned host pixels in, resized pixels out (icb_submit_png_fit)"},
           "png_file": {"resize_plus_encode_device_ms": png_ms, "e2e_images_per_s": n / ms_f * 1e3,
                        "file_bytes": int(rf[0].out_len), "raw_bytes": dw * dh * 4,
                        "note": "section 8 f3: the PNG file made on the GPU (icb_submit_png_compress); it inflates to the "
                                "resized pixels' filtered rows"},
           "roofline": {"kernel": "k_resize", "bound": "hbm", "achieved": touched / rs_ms / 1e6, "peak": peak_gbs,
                        "unit": "GB/s", "frac": touched / rs_ms / 1e6 / peak_gbs,
                        "traffic": traffic.get("k_resize_8k" if w > 4000 else "k_resize_4k"),
                        "algorithmic_bytes_per_launch": touched, "ms_per_launch": rs_ms},
           "parity_check": "resized pixels bit-exact vs oracle on image 0"}
    for x in res:
        eng.free_image(x)
    for a in pins + douts + fouts:
        eng.free_pinned(a)
    return out


def leg_mixed(eng, pool, threads, inflight):
    """cfg4: a mixed JPG/PNG list with a WARM learned cache.  70 % JPG pixels (12 MP and 24 MP, half each,
    1 MB target) and 30 % PNG pixels (4K and 8K RGBA8 into the 1920x1920 box), interleaved.  The H2 cache stays
    a host-side map; it is warmed by one search per distinct key, so every timed JPG takes the cache-hit path of
    core/ImageCompressionJpg.java:82-89, 216-238 (ONE encode at the learned quality).  e2e only."""
    from image_compression_b200 import LearnedParams
    from oracle import oracle as O
    import image_compression_b200 as pkg
    target, q = 1048576, 0.25
    kinds = []
    for (w, h) in [(4000, 3000), (6000, 4000)]:
        img = pool[0] if w == 4000 else compose(pool, w, h, 3, 1)
        r = eng.jpg_fit(img, q, target)                              # the warm-up search that fills the cache
        assert r.status == 0
        lp = LearnedParams(r.best_q, r.scale)
        rc, ref, rq, rs, uc, _ = O.jpg_fit(img, q, target, cached=(lp.quality, lp.scale))
        hit = eng.jpg_fit(img, q, target, cached=lp)
        assert (hit.status, hit.used_cache, hit.data) == (rc, bool(uc), ref) and uc, "cached fit differs from the oracle"
        kinds.append(("jpg_fit", pinned_copy(eng, img), lp))
    for (w, h) in [(3840, 2160), (7680, 4320)]:
        kinds.append(("png_fit", pinned_copy(eng, compose(pool, w, h, 4, 2)), None))
    order = [0, 1, 2, 0, 1, 3, 0, 1, 0, 2]                           # 7 JPG : 3 PNG per ten lines
    n = 20
    lst, bufs = [], []
    for i in range(n):
        kind, pin, lp = kinds[order[i % len(order)]]
        if kind == "jpg_fit":
            o = eng.alloc_pinned((target + 16,))
            lst.append(dict(kind=kind, image=pin, init_q=q, target=target, cached=lp, out=o))
        else:
            o = eng.alloc_pinned((1920 * 1080 * 4,))
            lst.append(dict(kind=kind, image=pin, min_w=1920, min_h=1920, out=o))
        bufs.append(o)
    items = eng.make_items(lst)
    st0 = eng.stats()
    ms, steps = steps_for(eng, [0], lambda: eng.run_list(items, threads, inflight), seconds=1.5)
    st1 = eng.stats()
    rr = items.results()
    assert all(x.status == 0 for x in rr)
    assert all(x.used_cache for x, d in zip(rr, lst) if d["kind"] == "jpg_fit"), "a cached fit fell back to the search"
    mp_step = sum(d["image"].shape[0] * d["image"].shape[1] for d in lst) / 1e6
    out = {"workload": "BASELINE configs[3]: mixed list, 70% JPG (12/24 MP, cache hit: one encode at the learned quality)"
                       " + 30% PNG (4K/8K RGBA8 -> 1920x1080), warm host-side cache",
           "value": mp_step / ms * 1e3, "unit": "MP/s", "images_per_s": n / ms * 1e3, "ms_per_step": ms, "steps": steps,
           "e2e": {"value": mp_step / ms * 1e3, "unit": "MP/s", "images_per_s": n / ms * 1e3,
                   "h2d_bytes_per_step": (st1["h2d_bytes"] - st0["h2d_bytes"]) / (steps + 1),
                   "d2h_bytes_per_step": (st1["d2h_bytes"] - st0["d2h_bytes"]) / (steps + 1)},
           "note": "e2e only: value == e2e (pinned host pixels in, JFIF bytes / resized pixels out)",
           "roofline": None,
           "parity_check": "both cached JPG kinds bit-exact vs oracle (used_cache, bytes); every timed JPG used the cache"}
    try:        # the same list with the JPG lines as FILES (decode on the GPU); the PNG lines stay decoded pixels
        import io
        from PIL import Image
        fkinds = []
        for (w, h) in [(4000, 3000), (6000, 4000)]:
            img = pool[0] if w == 4000 else compose(pool, w, h, 3, 1)
            bio = io.BytesIO()
            Image.fromarray(img).save(bio, "JPEG", quality=92, subsampling=2)
            fbytes = np.frombuffer(bio.getvalue(), np.uint8)
            r = eng.jpg_file_fit(fbytes, q, target)                      # warm-up search on the decoded pixels
            assert r.status == 0
            lp = LearnedParams(r.best_q, r.scale)
            px = O.jpeg_decode(fbytes.tobytes())
            rc, ref, rq, rs, uc, _ = O.jpg_fit(px, q, target, cached=(lp.quality, lp.scale))
            hit = eng.jpg_file_fit(fbytes, q, target, cached=lp)
            assert (hit.status, hit.used_cache, hit.data) == (rc, bool(uc), ref) and uc, "cached file fit differs from the oracle"
            fp = eng.alloc_pinned((fbytes.size,))
            fp[:] = fbytes
            fkinds.append((fp, lp))
        lst_f, bytes_in = [], 0
        for i, d in enumerate(lst):
            if d["kind"] == "jpg_fit":
                fp, lp = fkinds[order[i % len(order)]]
                lst_f.append(dict(kind="jpg_file_fit", file=fp, init_q=q, target=target, cached=lp, out=d["out"]))
            else:
                lst_f.append(d)
        items_f = eng.make_items(lst_f)
        eng.run_list(items_f, threads, inflight)
        sf0 = eng.stats()
        ms_f, steps_f = steps_for(eng, [0], lambda: eng.run_list(items_f, threads, inflight), seconds=1.5)
        sf1 = eng.stats()
        rf = items_f.results()
        assert all(x.status == 0 for x in rf)
        assert all(x.used_cache for x, d in zip(rf, lst_f) if d["kind"] == "jpg_file_fit"), "a cached file fit fell back to the search"
        out["e2e_file"] = {"value": mp_step / ms_f * 1e3, "unit": "MP/s", "images_per_s": n / ms_f * 1e3,
                           "h2d_bytes_per_step": (sf1["h2d_bytes"] - sf0["h2d_bytes"]) / (steps_f + 1),
                           "parity_check": "both cached JPG kinds from files bit-exact vs oracle decode + cached fit; every timed JPG used the cache",
                           "note": "JPG lines as JPEG files decoded on the GPU, PNG lines as decoded RGBA pixels"}
        for fp, _ in fkinds:
            eng.free_pinned(fp)
    except Exception as e:
        out["e2e_file"] = {"error": repr(e)}
    for k in kinds:
        eng.free_pinned(k[1])
    for o in bufs:
        eng.free_pinned(o)
    return out


def leg_latency(eng, pool, reps=5):
    """cfg5: ONE 100 MP source (decoded 5774x4330 after the s=2 subsampling of core/ImageCompression.java:140-153),
    256 KB target, -q 1.0 = the full 7-level sweep and the scale loop.  Latency, lower is better."""
    from oracle import oracle as O
    w, h, target, q = 5774, 4330, 262144, 1.0
    img = compose(pool, w, h, 3, 3)
    res = eng.upload_image(img, 0)
    pin = pinned_copy(eng, img)
    out = eng.alloc_pinned((target + 16,))
    r = None
    for _ in range(2):
        r = eng.wait(eng.submit_jpg_fit_image(res, q, target, out))
    ms, _ = timed(eng, [0], lambda: [eng.wait(eng.submit_jpg_fit_image(res, q, target, out)) for _ in range(reps)])
    st0 = eng.stats()
    ms_e, _ = timed(eng, [0], lambda: [eng.wait(eng.submit_jpg_fit(pin, q, target, out)) for _ in range(reps)])
    st1 = eng.stats()
    t0 = time.perf_counter()
    rc, ref, rq, rs, _, ne = O.jpg_fit(img, q, target)
    dt = time.perf_counter() - t0
    assert (r.status, r.best_q, r.scale, r.out_len) == (rc, rq, rs, len(ref)) and bytes(out[: r.out_len]) == ref, \
        "cfg5 differs from the oracle"
    o = {"workload": "BASELINE configs[4]: one 100MP source decoded 5774x4330, -t 262144 -q 1.0",
         "metric": "latency_ms_per_image", "value": ms / reps, "unit": "ms", "higher_is_better": False,
         "result": {"best_q": r.best_q, "scale": r.scale, "bytes": r.out_len, "entropy_passes": r.n_passes},
         "e2e": {"value": ms_e / reps, "unit": "ms", "h2d_bytes_per_step": (st1["h2d_bytes"] - st0["h2d_bytes"]) / reps,
                 "d2h_bytes_per_step": (st1["d2h_bytes"] - st0["d2h_bytes"]) / reps},
         "roofline": None,
         "cpu_baseline": {"value": dt * 1e3, "unit": "ms", "cores": 1, "kind": "port",
                          "sample": f"the same image once, 1 thread, {ne} full encodes + resizes (oracle port)"},
         "parity_check": "bit-exact vs oracle (%d bytes, q=%.7f, scale=%.4f)" % (len(ref), rq, rs)}
    eng.free_image(res)
    eng.free_pinned(pin)
    eng.free_pinned(out)
    return o


def leg_files(eng, pool, w, h, q, target, batch, args, dist, local, rank, local_dev=0):
    """The headline workload with JPEG FILES as input (what the reference's batch really starts from: `1,000 synthetic
    12MP JPGs`): Pillow / libjpeg-turbo writes each pool image as a q=92 4:2:0 baseline file; the file bytes sit in
    page-locked host memory, go up, are decoded on the GPU (icb_submit_jpg_file_fit) and fitted there."""
    import io
    import image_compression_b200 as pkg
    from PIL import Image
    files = []
    for im in pool:
        bio = io.BytesIO()
        Image.fromarray(im).save(bio, "JPEG", quality=92, subsampling=2)
        files.append(np.frombuffer(bio.getvalue(), np.uint8))
    fslots = max(args.slots, 24)                     # the decode kernels are latency bound: more jobs in flight
    T = max(args.threads, 3)                         # (r02 probes: 3 workers x 8 jobs is steadier than 2 x 12)
    inflight = max(1, fslots // T)
    engf = pkg.Engine(devices=[local_dev], slots_per_gpu=fslots, max_w=w, max_h=h)
    try:
        pins = []
        for f in files:
            a = engf.alloc_pinned((f.size,))
            a[:] = f
            pins.append(a)
        outs = [engf.alloc_pinned((target + 16,)) for _ in range(batch)]
        items = engf.make_items([dict(kind="jpg_file_fit", file=pins[i % len(pins)], init_q=q, target=target, out=outs[i])
                                 for i in range(batch)])
        engf.run_list(items, T, inflight)
        res = items.results()
        assert all(r.status == 0 and 0 < r.out_len <= target for r in res), [(r.status, r.out_len) for r in res][:8]
        parity = "skipped"
        if rank == 0 and not args.no_check:
            from oracle import oracle as O       # checker only, outside the timed region
            px = O.jpeg_decode(files[0].tobytes())
            rc, ref, rq, rs, _, _ = O.jpg_fit(px, q, target)
            assert (res[0].status, res[0].best_q, res[0].scale) == (rc, rq, rs) and bytes(outs[0][: res[0].out_len]) == ref, \
                "file path differs from the oracle"
            parity = ("bit-exact vs oracle decode + fit of file 0 (%d file bytes -> %d bytes, q=%.7f); decoded pixels equal "
                      "libjpeg-turbo's: %s" % (files[0].size, len(ref), rq,
                                               bool(np.array_equal(px, np.asarray(Image.open(io.BytesIO(files[0].tobytes())).convert("RGB"))))))
        for _ in range(max(1, args.warmup // 2)):
            engf.run_list(items, T, inflight)
        s0 = engf.stats()
        ms, wall = timed(engf, [0], lambda: [engf.run_list(items, T, inflight) for _ in range(args.steps)], dist, local)
        s1 = engf.stats()
        n = batch * args.steps
        return {"ms": ms, "wall_ms_per_step": wall / args.steps, "images": n,
                "h2d_bytes_per_step": (s1["h2d_bytes"] - s0["h2d_bytes"]) / args.steps,
                "d2h_bytes_per_step": (s1["d2h_bytes"] - s0["d2h_bytes"]) / args.steps,
                "file_bytes_mean": float(np.mean([f.size for f in files])), "slots_per_gpu": fslots,
                "gpu_launches": s1["kernels_launched"] - s0["kernels_launched"], "parity_check": parity}
    finally:
        engf.close()


# ---------------------------------------------------------------- GPU arm: the headline
def run_gpu(args):
    rank, local, world, dist = dist_setup()
    import image_compression_b200 as pkg
    w, h, target, q, desc = WORKLOADS[args.workload]
    peak_gbs, peak_src = peaks()
    traffic = traffic_table()
    slots, batch, T = args.slots, args.batch, args.threads
    inflight = max(1, slots // T)
    eng = pkg.Engine(devices=[local], slots_per_gpu=slots, max_w=w, max_h=h)
    pool = make_pool(w, h, POOL, seed0=1000 * rank)
    resident = [eng.upload_image(p, 0) for p in pool]
    pinned = [pinned_copy(eng, p) for p in pool]
    outs = [eng.alloc_pinned((target + 16,)) for _ in range(batch)]
    items_res = fit_items(eng, resident, batch, q, target, outs)
    items_host = fit_items(eng, pinned, batch, q, target, outs, host=True)

    # correctness of what is being timed: every image must fit, one is checked against the oracle
    eng.run_list(items_res, T, inflight)
    res = items_res.results()
    assert all(r.status == 0 and 0 < r.out_len <= target for r in res), [(r.status, r.out_len) for r in res]
    parity = "skipped"
    if rank == 0 and not args.no_check:
        from oracle import oracle as O          # checker only, outside every timed region
        rc, ref, rq, rs, _, _ = O.jpg_fit(pool[0], q, target)
        assert (res[0].status, res[0].best_q, res[0].scale) == (rc, rq, rs) and bytes(outs[0][: res[0].out_len]) == ref, \
            "GPU result differs from oracle"
        parity = "bit-exact vs oracle on pool[0] (%d bytes, q=%.7f), taken from the batch call that is timed" % (len(ref), rq)

    for s in range(args.warmup):
        eng.run_list(items_res, T, inflight)
    sampler = ClockSampler(local, enabled=rank == 0)
    sampler.start()
    st0 = eng.stats()
    ms, wall = timed(eng, [0], lambda: [eng.run_list(items_res, T, inflight) for _ in range(args.steps)], dist, local)
    st1 = eng.stats()
    launches = st1["kernels_launched"] - st0["kernels_launched"]

    for s in range(max(1, args.warmup // 2)):
        eng.run_list(items_host, T, inflight)
    se0 = eng.stats()
    ms_e2e, wall_e2e = timed(eng, [0], lambda: [eng.run_list(items_host, T, inflight) for _ in range(args.steps)],
                             dist, local)
    se1 = eng.stats()
    h2d_step = (se1["h2d_bytes"] - se0["h2d_bytes"]) / args.steps
    d2h_step = (se1["d2h_bytes"] - se0["d2h_bytes"]) / args.steps
    # ---- end to end from the FILES: the JPEG bytes go up, the GPU decodes them
    # (core/ImageCompression.java:119-155) and fits the pixels it decoded; nothing but the file crosses PCIe
    e2e_file = None
    try:
        e2e_file = leg_files(eng, pool, w, h, q, target, batch, args, dist, local, rank, local_dev=local)
    except Exception as e:                                           # never lose the headline over this leg
        e2e_file = {"error": repr(e)}
    clocks = sampler.stop()
    ms_ranks = gather_ranks(dist, local, world, eng.elapsed_ms(0))

    # ---- per-kernel device times.  (1) alone: live CUDA events on the launching stream, pool > L2;
    prof = [eng.profile_jpg_fit_image(resident[i % POOL], q, target, 1) for i in range(2 * POOL)][POOL:]
    xf_ms, _ = eng.profile_transform_series(resident[:POOL], 4, list_q=q)
    xf_nolist_ms, _ = eng.profile_transform_series(resident[:POOL], 4, list_q=0.0)
    ent_ms = statistics.mean(p["entropy_kernel_ms"] for p in prof)
    ent_n = statistics.mean(p["entropy_kernel_launches"] for p in prof)
    search_ms = statistics.mean(p["search_ms"] for p in prof)
    # (2) in the step: the kernel clocks (GPU global timer, first CTA start -> last CTA end) of every job of
    # the timed region, with the other slots' kernels sharing the GPU
    kns = [b - a for a, b in zip(st0["kernel_clock_ns"], st1["kernel_clock_ns"])]
    kn = [b - a for a, b in zip(st0["kernel_clock_launches"], st1["kernel_clock_launches"])]
    in_step = [kns[i] / kn[i] * 1e-6 if kn[i] else None for i in range(3)]       # ms per launch
    busy_ms = sum(kns) * 1e-6                                                     # summed kernel spans of the region
    nblocks = 6 * ((w + 15) // 16) * ((h + 15) // 16)
    xf_bytes = 3 * w * h + 128 * nblocks
    ent_bytes = 128 * nblocks                      # one coefficient read per visited node

    def roof(name, nbytes, ms_alone, k):
        a = nbytes / (ms_alone * 1e-3) / 1e9
        r = {"kernel": name, "bound": "hbm", "achieved": a, "peak": peak_gbs, "unit": "GB/s", "frac": a / peak_gbs,
             "traffic": traffic.get(name), "algorithmic_bytes_per_launch": nbytes, "ms_per_launch": ms_alone,
             "ms_per_launch_in_step": in_step[k], "launches_in_timed_region": kn[k], "peak_source": peak_src}
        if in_step[k] and busy_ms > 0:
            # device time the step really spends per launch of this kernel: the step's device time split by the
            # kernels' shares of the summed spans (several jobs overlap, so spans add up to more than the step)
            eff = ms * (kns[k] * 1e-6 / busy_ms) / kn[k]
            r["ms_per_launch_effective"] = eff
            r["frac_effective"] = nbytes / (eff * 1e-3) / 1e9 / peak_gbs
        return r

    r_xf = roof("k_transform", xf_bytes, xf_ms, 0)
    r_xf["timing"] = ("alone = the pool's transforms back to back on one stream between two CUDA events, WITH the "
                      "survivor lists it writes inside a fit (%.4f ms without them)" % xf_nolist_ms)
    r_ent = roof("k_entropy_pass", ent_bytes, ent_ms / max(ent_n, 1), 1)
    r_ent["note"] = ("integer-issue bound on L2-resident survivor lists, not HBM bound: the roofline fraction against "
                     "HBM is low by construction; see profiles/r02 for smsp__issue_active")
    dominant = r_ent if kns[1] >= kns[0] else r_xf
    concurrency = busy_ms / ms if ms else None

    n_img = batch * args.steps
    ips = n_img / (ms * 1e-3) * world             # every rank runs the same per-GPU batch (weak scaling)
    ips_e2e = n_img * world / (ms_e2e * 1e-3)
    mpx = w * h / 1e6
    e2e_pixels = {"value": ips_e2e * mpx, "unit": "MP/s", "images_per_s": ips_e2e, "h2d_bytes_per_step": h2d_step,
                  "d2h_bytes_per_step": d2h_step, "ms_per_step": ms_e2e / args.steps,
                  "input": "decoded RGB pixels in page-locked host memory (icb_submit_jpg_fit): 36 MB per image over PCIe"}
    e2e = e2e_pixels
    if e2e_file and "ms" in e2e_file:
        ips_f = e2e_file["images"] * world / (e2e_file["ms"] * 1e-3)
        e2e = {"value": ips_f * mpx, "unit": "MP/s", "images_per_s": ips_f,
               "h2d_bytes_per_step": e2e_file["h2d_bytes_per_step"], "d2h_bytes_per_step": e2e_file["d2h_bytes_per_step"],
               "ms_per_step": e2e_file["ms"] / args.steps,
               "input": "JPEG files (Pillow q=92 4:2:0, mean %.2f MB) in page-locked host memory, decoded on the GPU "
                        "(icb_submit_jpg_file_fit): the timed region holds MORE than the pixel path -- the decode the "
                        "reference does on the host and the CPU arm does not time" % (e2e_file["file_bytes_mean"] / 1e6),
               "slots_per_gpu": e2e_file["slots_per_gpu"], "gpu_launches": e2e_file["gpu_launches"],
               "parity_check": e2e_file["parity_check"]}
    launches_all = int(_reduce(dist, local, float(launches), "SUM"))

    cpu = fast = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        n_cpu = max(cores, 8)
        dt, _ = cpu_fit_batch(pool, n_cpu, cores, q, target)
        cpu = {"value": n_cpu / dt * mpx, "unit": "MP/s", "images_per_s": n_cpu / dt, "cores": cores, "kind": "port",
               "sample": f"{n_cpu} of the same 12MP images, one per thread, {cores} threads, {dt:.1f}s wall; oracle port "
                         f"(scalar IJG-6b restatement, {oracle_build_note()}), 5 search encodes + 1 save encode per image"}
        fast = fast_cpu_lines(pool[0], target, q)

    # ---- the other BASELINE configs, rank 0, each checked against the oracle in the run
    for im in resident:
        eng.free_image(im)
    for a in pinned + outs:
        eng.free_pinned(a)
    other = {}
    if rank == 0 and not args.no_legs and args.workload == "cfg1":
        for name, fn in [("cfg2", lambda: leg_jpg(eng, pool, "cfg2", T, inflight, peak_gbs)),
                         ("png4k", lambda: leg_png(eng, pool, 3840, 2160, T, inflight, peak_gbs, traffic)),
                         ("png8k", lambda: leg_png(eng, pool, 7680, 4320, T, inflight, peak_gbs, traffic)),
                         ("cfg4", lambda: leg_mixed(eng, pool, T, inflight)),
                         ("cfg5", lambda: leg_latency(eng, pool))]:
            try:
                t0 = time.perf_counter()
                other[name] = fn()
                other[name]["leg_wall_s"] = round(time.perf_counter() - t0, 1)
            except Exception as e:                                   # a leg must never lose the headline
                other[name] = {"error": repr(e)}
    eng.close()
    host_barrier(dist)                  # (the other ranks wait on the host while rank 0 runs its legs)

    # ---- one process, all GPUs (the Java host's shape, CompressionBatch.java:64-90)
    one = None
    if world > 1 and not args.no_legs:
        if rank == 0:
            try:
                one = one_process_all_gpus(world, pool, w, h, q, target, slots, batch)
            except Exception as e:
                one = {"error": repr(e)}
        host_barrier(dist)

    if rank == 0:
        line = {
            "metric": "megapixels_per_s", "value": ips * mpx, "unit": "MP/s", "images_per_s": ips,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic",
            "config": {"workload": desc, "images_per_gpu_per_step": batch, "slots_per_gpu": slots,
                       "executor_threads": T, "jobs_in_flight_per_thread": inflight,
                       "l2": f"pool of {POOL} distinct {w}x{h} images per GPU ({POOL * w * h * 3 >> 20} MiB) > 126 MB L2",
                       "parallelism": f"images sharded over {world} GPU(s), no collective", "parity_check": parity},
            "clocks": clocks,
            "e2e": e2e, "e2e_pixels": e2e_pixels,
            "e2e_file_error": e2e_file.get("error") if e2e_file else None,
            "gpu_launches": launches_all,
            "roofline": dominant, "roofline_transform": r_xf, "roofline_entropy": r_ent,
            "device_ms_per_image": {"transform": xf_ms, "search_total": search_ms, "entropy_kernels": ent_ms,
                                    "entropy_passes": ent_n, "emit_in_step": in_step[2]},
            "in_step": {"kernel_span_ms_summed": busy_ms, "device_ms": ms, "concurrency": concurrency,
                        "note": "kernel spans (first CTA start to last CTA end, GPU global timer) of every job of the timed "
                                "region; spans of concurrent jobs overlap, so their sum / device time = average number of "
                                "kernels in flight"},
            "per_rank_device_ms": ms_ranks,
            "host_wall_ms_per_step": {"resident": wall / args.steps, "e2e": wall_e2e / args.steps},
            "cpu_baseline": cpu, "fast_cpu": fast,
            "other_configs": other, "one_process": one,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def one_process_all_gpus(world, pool, w, h, q, target, slots, batch):
    """Rank 0 alone opens every GPU of the box from ONE process and drives them through the native executor:
    the shape of the Java host (one JVM, CompressionBatch.java:64-90).  Same per-GPU batch as the torchrun figure."""
    import image_compression_b200 as pkg
    eng = pkg.Engine(devices=list(range(world)), slots_per_gpu=slots, max_w=w, max_h=h)
    try:
        n = batch * world
        resident = [eng.upload_image(pool[i % len(pool)], d) for d in range(world) for i in range(4)]
        pins = [pinned_copy(eng, p) for p in pool[:4]]
        outs = [eng.alloc_pinned((target + 16,)) for _ in range(n)]
        devs = list(range(world))
        threads = min(2 * world, 16)
        infl = max(1, slots * world // threads)
        # resident: item i runs on the GPU that holds its image -> interleave the GPUs along the list
        imgs = [resident[(i % world) * 4 + (i // world) % 4] for i in range(n)]
        items = fit_items(eng, imgs, n, q, target, outs)
        eng.run_list(items, threads, infl)
        assert all(r.status == 0 for r in items.results())
        used = sorted({r.device for r in items.results()})
        ms, steps = steps_for(eng, devs, lambda: eng.run_list(items, threads, infl), seconds=1.5)
        items_h = fit_items(eng, pins, n, q, target, outs, host=True)
        eng.run_list(items_h, threads, infl)
        ms_e, steps_e = steps_for(eng, devs, lambda: eng.run_list(items_h, threads, infl), seconds=2.0)
        return {"gpus": world, "devices_used": used, "executor_threads": threads, "jobs_in_flight_per_thread": infl,
                "images_per_s": n / ms * 1e3, "e2e_images_per_s": n / ms_e * 1e3, "images_per_step": n,
                "steps": [steps, steps_e],
                "note": "ONE host process, icb_init over all GPUs, icb_run_list (atomic cursor, worker threads inside the "
                        "library); device time = max over the GPUs' CUDA-event spans"}
    finally:
        eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg1", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=128, help="images per GPU per step")
    ap.add_argument("--slots", type=int, default=8, help="stream work-queue slots per GPU (r02 A/B: 6 -> 12.6k, 8 -> 13.0k, 12 -> 13.1k img/s resident)")
    ap.add_argument("--threads", type=int, default=2, help="executor threads per GPU (each keeps slots/threads jobs in flight)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--no-legs", action="store_true", help="headline only: skip other_configs and one_process")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
