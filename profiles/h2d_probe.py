"""Where is the host-to-device ceiling?  (e2e efficiency 0.41 at 8 GPUs.)

Plain pinned -> device cudaMemcpyAsync of 36 MB buffers (one decoded 12 MP image), nothing of this
library involved, on 1 / 2 / 4 / 8 GPUs of the box:

  python profiles/h2d_probe.py                     one process drives every GPU (one stream per GPU)
  torchrun --nproc-per-node N profiles/h2d_probe.py   one process per GPU (how bench.py --gpus N runs)

Prints one JSON line per configuration: aggregate GB/s (device-timed, max over GPUs), per-GPU GB/s, and
the NUMA node / CPU affinity the process sees.  If this shows the same ceiling as bench.py's e2e figure,
the limit is the platform's (host memory / PCIe root), not the staging code.
"""
import json
import os
import sys
import time

import torch

NBYTES = 36_000_000
REPS = 40
NBUF = 4


def probe_one_process(devs, write_combined=False):
    bufs, dsts, streams, evs = {}, {}, {}, {}
    for d in devs:
        torch.cuda.set_device(d)
        bufs[d] = [torch.empty(NBYTES, dtype=torch.uint8).pin_memory() for _ in range(NBUF)]
        dsts[d] = [torch.empty(NBYTES, dtype=torch.uint8, device=f"cuda:{d}") for _ in range(NBUF)]
        streams[d] = torch.cuda.Stream(device=d)
        evs[d] = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
    for phase in ("warm", "timed"):
        t0 = time.perf_counter()
        for d in devs:
            with torch.cuda.stream(streams[d]):
                if phase == "timed":
                    evs[d][0].record(streams[d])
        for r in range(REPS if phase == "timed" else 4):
            for d in devs:
                with torch.cuda.stream(streams[d]):
                    dsts[d][r % NBUF].copy_(bufs[d][r % NBUF], non_blocking=True)
        for d in devs:
            with torch.cuda.stream(streams[d]):
                if phase == "timed":
                    evs[d][1].record(streams[d])
        for d in devs:
            streams[d].synchronize()
        wall = time.perf_counter() - t0
    per = {d: NBYTES * REPS / (evs[d][0].elapsed_time(evs[d][1]) * 1e-3) / 1e9 for d in devs}
    return {"aggregate_gbs_wall": NBYTES * REPS * len(devs) / wall / 1e9, "per_gpu_gbs": per}


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    try:
        aff = sorted(os.sched_getaffinity(0))
        aff = f"{aff[0]}-{aff[-1]} ({len(aff)} cpus)"
    except Exception:
        aff = "?"
    if world > 1:
        import torch.distributed as dist
        rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
        r = probe_one_process([local])
        t = torch.tensor([r["per_gpu_gbs"][local]], dtype=torch.float64, device=f"cuda:{local}")
        mn = t.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        if rank == 0:
            print(json.dumps({"mode": "process_per_gpu", "gpus": world, "sum_gbs": float(sm.item()),
                              "min_gpu_gbs": float(mn.item()), "cpu_affinity": aff}), flush=True)
        dist.destroy_process_group()
        return
    n = torch.cuda.device_count()
    for k in (1, 2, 4, 8):
        if k > n:
            break
        r = probe_one_process(list(range(k)))
        print(json.dumps({"mode": "one_process", "gpus": k, "sum_gbs": sum(r["per_gpu_gbs"].values()),
                          "min_gpu_gbs": min(r["per_gpu_gbs"].values()), "aggregate_gbs_wall": r["aggregate_gbs_wall"],
                          "cpu_affinity": aff}), flush=True)


if __name__ == "__main__":
    sys.exit(main())
