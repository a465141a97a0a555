"""CPU suite, part 3: the N>1 host logic.  Two gloo ranks take interleaved lines of one file
list (image_compression_b200/batch.py, the stand-in for CompressionBatch.java:64-133), process
them through a stub with the Engine's two seam methods (no device here) and sum the report
counters; the merged report must equal the single-process one."""
import os
import socket
import sys

import numpy as np
import pytest

from image_compression_b200 import batch, CompressionParams

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class StubEngine:
    """Deterministic stand-in with the seams' contract: JPG 'fits' unless the image is
    all-zero, PNG resizes unless it already fits the box (ImageCompressionPng.java:49-53)."""

    class _Fit:
        def __init__(self, data):
            self.data = data

    def compress_jpg_with_target_size(self, pixels, original_size, output_file, params, cache):
        if not pixels.any():
            return False
        n = int(pixels.sum()) % params.target_max_size_bytes + 1
        self.last_fit = self._Fit(b"\0" * n)
        return True

    def compress_png_with_target_size(self, pixels, output_file, params):
        h, w = pixels.shape[:2]
        if w <= params.min_width and h <= params.min_height:
            return False, None
        return True, pixels[::2, ::2]


def make_items(n=23):
    rng = np.random.default_rng(5)
    items = []
    for i in range(n):
        kind = ("jpg", "png", "gif")[i % 3] if i % 7 else "jpg"
        shape = (8 + i % 5, 6 + i % 4, 3)
        px = rng.integers(0, 255, shape, dtype=np.uint8)
        if i % 11 == 0:
            px[...] = 0                                   # a JPG that cannot be made to fit
        items.append(batch.WorkItem(i, kind, px, 1000 + 17 * i))
    return items


PARAMS = CompressionParams(0.25, 1 << 20, 10, 8, 4096)


def test_shard_indices_partition_the_list():
    for n in (0, 1, 7, 23, 100):
        for world in (1, 2, 3, 8):
            seen = sorted(i for r in range(world) for i in batch.shard_indices(n, r, world))
            assert seen == list(range(n))
    with pytest.raises(ValueError):
        batch.shard_indices(10, 2, 2)


def test_counter_accounting_quirks():
    c = batch.BatchCounters()
    c.add(batch.SKIPPED_CONDITION_NOT_MET, 100, 100)      # ImageCompression.java:70-73
    c.add(batch.SKIPPED_NOT_FOUND, 0, 0)                  # :57
    c.add(batch.FAILED_COMPRESSION, 50, 0)                # :92 -- counts as 100 % saved
    assert c.total_original_size == 150 and c.total_compressed_size == 100
    assert abs(c.saved_percent() - 100.0 * 50 / 150) < 1e-12
    assert batch.status_to_result(0) == batch.COMPRESSED_SUCCESS
    assert batch.status_to_result(1) == batch.FAILED_COMPRESSION
    assert batch.status_to_result(-2) == batch.FAILED_OUT_OF_MEMORY
    assert batch.status_to_result(-3) == batch.FAILED_UNKNOWN


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        local = batch.run_shard(StubEngine(), make_items(), PARAMS, {}, rank=rank, world=world)
        merged = batch.reduce_counters(local, dist)
        q.put((rank, local.as_vector(), merged.as_vector()))
    finally:
        dist.destroy_process_group()


def test_two_gloo_ranks_merge_to_the_single_process_report():
    import torch.multiprocessing as mp
    single = batch.run_shard(StubEngine(), make_items(), PARAMS, {})
    assert single.counts[batch.COMPRESSED_SUCCESS] > 0 and single.counts[batch.FAILED_COMPRESSION] > 0
    assert single.counts[batch.FAILED_UNSUPPORTED_FORMAT] > 0
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    locals_ = {r: v for r, v, _ in got}
    assert locals_[0] != locals_[1]                       # each rank really did a different share
    assert [a + b for a, b in zip(locals_[0], locals_[1])] == single.as_vector()
    for _, _, merged in got:
        assert merged == single.as_vector()               # every rank holds the full report
