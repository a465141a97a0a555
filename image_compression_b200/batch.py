"""Host side of the batch: how a file list is split over the GPUs and how the report
counters come back together.

The reference runs one JVM with a fixed thread pool, one task per file-list line
(CompressionBatch.java:64-90), AtomicLong counters (:56-62, :85-87) and a final report
(:110-133).  Here every GPU is one process (rank) that owns an Engine with a few stream
work slots; ranks take interleaved lines of the SAME list (images are independent, so no
collective sits on the data path) and the only cross-rank step is the sum of the
counters for the report -- a host-side all-reduce of seven integers.
"""
from __future__ import annotations

import threading
from collections import deque
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import numpy as np

# core/CompressionResult.java:3-16
COMPRESSED_SUCCESS = "COMPRESSED_SUCCESS"
SKIPPED_CONDITION_NOT_MET = "SKIPPED_CONDITION_NOT_MET"
SKIPPED_NOT_FOUND = "SKIPPED_NOT_FOUND"
FAILED_UNSUPPORTED_FORMAT = "FAILED_UNSUPPORTED_FORMAT"
FAILED_COMPRESSION = "FAILED_COMPRESSION"
FAILED_IO_ERROR = "FAILED_IO_ERROR"
FAILED_OUT_OF_MEMORY = "FAILED_OUT_OF_MEMORY"
FAILED_UNKNOWN = "FAILED_UNKNOWN"
RESULTS = (COMPRESSED_SUCCESS, SKIPPED_CONDITION_NOT_MET, SKIPPED_NOT_FOUND, FAILED_UNSUPPORTED_FORMAT,
           FAILED_COMPRESSION, FAILED_IO_ERROR, FAILED_OUT_OF_MEMORY, FAILED_UNKNOWN)


def status_to_result(status: int) -> str:
    """icb_status -> CompressionResult, the mapping include/imgcomp_b200.h documents
    (what core/ImageCompression.java:81-104 does with the seam's boolean / exception)."""
    if status == 0:
        return COMPRESSED_SUCCESS
    if status == 1:
        return FAILED_COMPRESSION
    if status == -2:
        return FAILED_OUT_OF_MEMORY
    return FAILED_UNKNOWN


def shard_indices(n_items: int, rank: int, world: int) -> range:
    """Lines of the file list that `rank` of `world` processes: interleaved, so that a list
    sorted by size or format still balances (the reference's pool pulls lines in order)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return range(rank, n_items, world)


@dataclass
class BatchCounters:
    """CompressionBatch.java:56-62 -- per-result counts and the two byte totals."""
    counts: Dict[str, int] = field(default_factory=lambda: {r: 0 for r in RESULTS})
    total_original_size: int = 0
    total_compressed_size: int = 0

    def add(self, result: str, original: int, compressed: int) -> None:
        # accounting quirks kept: skipped -> (orig, orig), not found -> (0, 0), failures -> (orig, 0)
        # (core/ImageCompression.java:57, :70-73, :92-103; CompressionBatch.java:85-87)
        self.counts[result] += 1
        self.total_original_size += original
        self.total_compressed_size += compressed

    def as_vector(self) -> List[int]:
        return [self.counts[r] for r in RESULTS] + [self.total_original_size, self.total_compressed_size]

    @classmethod
    def from_vector(cls, v: Sequence[int]) -> "BatchCounters":
        c = cls()
        for r, x in zip(RESULTS, v[: len(RESULTS)]):
            c.counts[r] = int(x)
        c.total_original_size = int(v[len(RESULTS)])
        c.total_compressed_size = int(v[len(RESULTS) + 1])
        return c

    def merged(self, other: "BatchCounters") -> "BatchCounters":
        return BatchCounters.from_vector([a + b for a, b in zip(self.as_vector(), other.as_vector())])

    def saved_percent(self) -> float:
        if self.total_original_size <= 0:
            return 0.0
        return 100.0 * (self.total_original_size - self.total_compressed_size) / self.total_original_size


def reduce_counters(local: BatchCounters, dist=None, device=None) -> BatchCounters:
    """Sum over ranks (the reference's AtomicLongs).  `dist` is torch.distributed (gloo on CPU,
    nccl on GPUs) or None for a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    import torch
    t = torch.tensor(local.as_vector(), dtype=torch.int64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return BatchCounters.from_vector(t.tolist())


@dataclass
class WorkItem:
    """One line of the file list after host-side decode (the reference's DecodedImage +
    original file size; decode itself is input staging and stays on the host)."""
    index: int
    kind: str                 # "jpg" | "png"
    pixels: object            # uint8 ndarray (h, w, 3|4); may be None for a "jpg" line that carries file_bytes
    original_size: int
    file_bytes: object = None  # the JPEG file itself (bytes / uint8 ndarray): SURVEY 8 f2 -- lines whose file the GPU
    #                            reader takes go up as files (core/ImageCompression.java:119-155 on the device) and
    #                            need no host decode; every other line is decoded on the host into `pixels` as before


class _Cursor:
    """The one shared cursor the workers pull lines from (dynamic, granularity one image --
    the reference's pool threads take the next queued task the same way, CompressionBatch.java:72-90)."""

    def __init__(self, n: int):
        self._n, self._i, self._lock = n, 0, threading.Lock()

    def next(self) -> int:
        with self._lock:
            i = self._i
            self._i += 1
        return i if i < self._n else -1


def _run_sync(engine, lines, cursor, params, cache, on_output) -> BatchCounters:
    """One worker over the synchronous seams (anything with the Engine's two seam methods)."""
    c = BatchCounters()
    while True:
        k = cursor.next()
        if k < 0:
            return c
        it = lines[k]
        try:
            if it.kind == "jpg":
                ok = engine.compress_jpg_with_target_size(it.pixels, it.original_size, None, params, cache)
                out = engine.last_fit.data if ok else None
                size = len(out) if ok else 0
            elif it.kind == "png":
                ok, out = engine.compress_png_with_target_size(it.pixels, "-", params)
                size = int(out.size) if ok else 0           # resized pixels; the PNG writer stays with the host
            else:
                c.add(FAILED_UNSUPPORTED_FORMAT, it.original_size, it.original_size)
                continue
            if ok:
                c.add(COMPRESSED_SUCCESS, it.original_size, size)
                if on_output is not None:
                    on_output(it, out)
            else:
                c.add(FAILED_COMPRESSION, it.original_size, 0)
        except MemoryError:
            c.add(FAILED_OUT_OF_MEMORY, it.original_size, 0)
        except Exception:
            c.add(FAILED_UNKNOWN, it.original_size, 0)


def _run_queued(engine, lines, cursor, params, cache, on_output, inflight: int) -> BatchCounters:
    """One worker over the queue form of the seams: up to `inflight` jobs of this thread sit in the GPU
    work queues while it stages the next line (icb_submit_* / icb_wait); cache.get before the submit and
    cache.put after the wait, as core/ImageCompressionJpg.java:79-80, :111 do around the encode."""
    from .api import IcbError, LearnedParams, create_key, png_target_dims
    c = BatchCounters()
    pending = deque()
    target = int(params.target_max_size_bytes)
    # one output buffer per job in flight, page-locked so the stream lands in it without a bounce
    jpg_outs = [engine.alloc_pinned((max(target, 0) + 16,)) for _ in range(inflight)]
    free_outs = list(range(inflight))

    def finish():
        it, kind, ticket, out, extra = pending.popleft()
        try:
            r = engine.wait(ticket)
            if kind == "jpg":
                key, slot = extra
                if r.status == 0:
                    data = out[: r.out_len].tobytes()
                    if not r.used_cache:
                        cache[key] = LearnedParams(r.best_q, r.scale)
                    c.add(COMPRESSED_SUCCESS, it.original_size, len(data))
                    if on_output is not None:
                        on_output(it, data)
                else:
                    c.add(status_to_result(r.status), it.original_size, 0)
            else:
                dw, dh, ch = extra
                px = out[: dw * dh * ch].reshape(dh, dw, ch)
                c.add(COMPRESSED_SUCCESS, it.original_size, int(px.size))
                if on_output is not None:
                    on_output(it, px)
        except IcbError as e:
            c.add(status_to_result(e.code), it.original_size, 0)
        finally:
            if kind == "jpg":
                free_outs.append(extra[1])

    while True:
        k = cursor.next()
        if k < 0:
            break
        it = lines[k]
        if it.kind not in ("jpg", "png"):
            c.add(FAILED_UNSUPPORTED_FORMAT, it.original_size, it.original_size)
            continue
        if len(pending) == inflight:
            finish()
        try:
            if it.kind == "jpg" and it.file_bytes is not None and hasattr(engine, "submit_jpg_file_fit"):
                from .api import jpg_file_info, subsample_factor
                try:
                    fw, fh = jpg_file_info(it.file_bytes)
                except IcbError:
                    fw = 0                                          # another kind of JPEG: the pixel path below
                if fw:
                    sf = subsample_factor(fw, fh)                   # decode rule :140-153 -> the dims the cache key sees
                    key = create_key((fw + sf - 1) // sf, (fh + sf - 1) // sf, it.original_size)
                    slot = free_outs.pop()
                    try:
                        t = engine.submit_jpg_file_fit(it.file_bytes, params.quality, target, jpg_outs[slot], cached=cache.get(key))
                    except Exception:
                        free_outs.append(slot)
                        raise
                    pending.append((it, "jpg", t, jpg_outs[slot], (key, slot)))
                    continue
                if it.pixels is None:
                    c.add(FAILED_UNSUPPORTED_FORMAT, it.original_size, it.original_size)
                    continue
            h, w = it.pixels.shape[:2]
            if it.kind == "jpg":
                key = create_key(w, h, it.original_size)
                slot = free_outs.pop()
                try:
                    t = engine.submit_jpg_fit(it.pixels, params.quality, target, jpg_outs[slot], cached=cache.get(key))
                except Exception:
                    free_outs.append(slot)
                    raise
                pending.append((it, "jpg", t, jpg_outs[slot], (key, slot)))
            else:
                dims = png_target_dims(w, h, params.min_width, params.min_height)
                if dims is None:
                    c.add(FAILED_COMPRESSION, it.original_size, 0)          # the seam returns false (:49-53)
                    continue
                ch = it.pixels.shape[2] if it.pixels.ndim == 3 else 1
                dst = np.empty(dims[0] * dims[1] * ch, np.uint8)
                t, dw, dh = engine.submit_png_fit(it.pixels, params.min_width, params.min_height, dst)
                pending.append((it, "png", t, dst, (dw, dh, ch)))
        except MemoryError:
            c.add(FAILED_OUT_OF_MEMORY, it.original_size, 0)
        except IcbError as e:
            c.add(status_to_result(e.code), it.original_size, 0)
        except Exception:
            c.add(FAILED_UNKNOWN, it.original_size, 0)
    while pending:
        finish()
    for o in jpg_outs:
        engine.free_pinned(o)
    return c


def run_shard(engine, items: Iterable[WorkItem], params, cache, rank: int = 0, world: int = 1,
              on_output: Optional[Callable[[WorkItem, object], None]] = None, threads: int = 1,
              inflight_per_thread: int = 2) -> BatchCounters:
    """Process this rank's share of `items` and count results like CompressionBatch does.

    Ranks (one process per GPU) take interleaved lines of the list; inside a rank `threads` workers pull
    from ONE shared cursor and, when the engine has the queue form of the seams (icb_submit_* / icb_wait),
    each keeps `inflight_per_thread` jobs in the GPU work queues, so uploads, kernels and downloads of
    different lines overlap.  An engine that only has the two synchronous seam methods is driven through
    those.  An Engine opened on several GPUs spreads the jobs over all of them (one process, the Java
    host's shape)."""
    items = list(items)
    lines = [items[i] for i in shard_indices(len(items), rank, world)]
    cursor = _Cursor(len(lines))
    queued = hasattr(engine, "submit_jpg_fit") and hasattr(engine, "submit_png_fit")
    threads = max(1, int(threads))
    results: List[Optional[BatchCounters]] = [None] * threads

    def work(t: int) -> None:
        if queued:
            results[t] = _run_queued(engine, lines, cursor, params, cache, on_output, max(1, int(inflight_per_thread)))
        else:
            results[t] = _run_sync(engine, lines, cursor, params, cache, on_output)

    if threads == 1:
        work(0)
    else:
        ts = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
    total = BatchCounters()
    for r in results:
        if r is not None:
            total = total.merged(r)
    return total
