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

from dataclasses import dataclass, field, asdict
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

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
    pixels: object            # uint8 ndarray (h, w, 3|4)
    original_size: int


def run_shard(engine, items: Iterable[WorkItem], params, cache, rank: int = 0, world: int = 1,
              on_output: Optional[Callable[[WorkItem, object], None]] = None) -> BatchCounters:
    """Process this rank's share of `items` through the seams (image_compression_b200.api.Engine
    or anything with the same two methods) and count results like CompressionBatch does."""
    counters = BatchCounters()
    items = list(items)
    for i in shard_indices(len(items), rank, world):
        it = items[i]
        try:
            if it.kind == "jpg":
                ok = engine.compress_jpg_with_target_size(it.pixels, it.original_size, None, params, cache)
                out = engine.last_fit.data if ok else None
                size = len(out) if ok else 0
            elif it.kind == "png":
                ok, out = engine.compress_png_with_target_size(it.pixels, "-", params)
                size = int(out.size) if ok else 0           # resized pixels; the PNG writer stays with the host
            else:
                counters.add(FAILED_UNSUPPORTED_FORMAT, it.original_size, it.original_size)
                continue
            if ok:
                counters.add(COMPRESSED_SUCCESS, it.original_size, size)
                if on_output is not None:
                    on_output(it, out)
            else:
                counters.add(FAILED_COMPRESSION, it.original_size, 0)
        except MemoryError:
            counters.add(FAILED_OUT_OF_MEMORY, it.original_size, 0)
        except Exception:
            counters.add(FAILED_UNKNOWN, it.original_size, 0)
    return counters
